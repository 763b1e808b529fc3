//! `extern "C"` mirror of `include/dfgpu.h` (the drop-in boundary): same names, same field order, same arity.
//! `tests/test_rust_binding.py` parses this file and the header and fails on any difference, so a field appended to a C
//! struct cannot silently leave the library writing past a Rust struct.
#![allow(non_camel_case_types)]
use arrow::ffi::{FFI_ArrowArray, FFI_ArrowSchema};
use std::os::raw::{c_char, c_int, c_void};

// opaque handles
#[repr(C)] pub struct dfgpu_ctx { _p: [u8; 0] }
#[repr(C)] pub struct dfgpu_batch { _p: [u8; 0] }
#[repr(C)] pub struct dfgpu_agg { _p: [u8; 0] }
#[repr(C)] pub struct dfgpu_join { _p: [u8; 0] }
#[repr(C)] pub struct dfgpu_coalescer { _p: [u8; 0] }

// dfgpu_status
pub const DFGPU_OK: c_int = 0;
pub const DFGPU_INTERNAL: c_int = 1;
pub const DFGPU_PLAN: c_int = 2;
pub const DFGPU_EXECUTION: c_int = 3;
pub const DFGPU_RESOURCES_EXHAUSTED: c_int = 4;
pub const DFGPU_NOT_IMPLEMENTED: c_int = 5;
pub const DFGPU_CUDA: c_int = 6;
pub const DFGPU_NCCL: c_int = 7;
// AggregateMode (physical-plan/src/aggregates/mod.rs:63-100)
pub const DFGPU_AGG_PARTIAL: i32 = 0;
pub const DFGPU_AGG_FINAL: i32 = 1;
pub const DFGPU_AGG_FINAL_PARTITIONED: i32 = 2;
pub const DFGPU_AGG_SINGLE: i32 = 3;
pub const DFGPU_AGG_SINGLE_PARTITIONED: i32 = 4;
// GroupsAccumulator implementations on the path
pub const DFGPU_SUM: i32 = 1;
pub const DFGPU_COUNT: i32 = 2;
pub const DFGPU_AVG: i32 = 3;
pub const DFGPU_MIN: i32 = 4;
pub const DFGPU_MAX: i32 = 5;
pub const DFGPU_BIT_AND: i32 = 6;
pub const DFGPU_BIT_OR: i32 = 7;
pub const DFGPU_BIT_XOR: i32 = 8;
pub const DFGPU_BOOL_AND: i32 = 9;
pub const DFGPU_BOOL_OR: i32 = 10;
// JoinType (common/src/join_type.rs:30-47)
pub const DFGPU_JOIN_INNER: i32 = 0;
pub const DFGPU_JOIN_LEFT: i32 = 1;
pub const DFGPU_JOIN_RIGHT: i32 = 2;
pub const DFGPU_JOIN_FULL: i32 = 3;
pub const DFGPU_JOIN_LEFT_SEMI: i32 = 4;
pub const DFGPU_JOIN_RIGHT_SEMI: i32 = 5;
pub const DFGPU_JOIN_LEFT_ANTI: i32 = 6;
pub const DFGPU_JOIN_RIGHT_ANTI: i32 = 7;
// PartitionMode (physical-plan/src/joins/mod.rs:39-47)
pub const DFGPU_PARTITIONED: i32 = 0;
pub const DFGPU_COLLECT_LEFT: i32 = 1;
// dfgpu_expr_node.kind
pub const DFGPU_EXPR_COLUMN: i32 = 0;
pub const DFGPU_EXPR_LITERAL: i32 = 1;
pub const DFGPU_EXPR_ADD: i32 = 2;
pub const DFGPU_EXPR_SUB: i32 = 3;
pub const DFGPU_EXPR_MUL: i32 = 4;
pub const DFGPU_NOALLOC_MAX_ROWS: i64 = 1 << 20;
pub const DFGPU_NOALLOC_MAX_PARTS: i32 = 256;

#[repr(C)]
pub struct dfgpu_options {
    pub device: i32,
    pub batch_size: i32,
    pub memory_limit: i64,
    pub load_factor: f64,
    pub force_hash_collisions: i32,
    pub rank: i32,
    pub world: i32,
    pub nccl_unique_id: *const c_void,
    pub chunk_rows: i64,
    pub coalesce_rows: i64,
}

#[repr(C)]
#[derive(Default, Clone, Copy, Debug)]
pub struct dfgpu_metrics {
    pub elapsed_compute_ns: i64,
    pub output_rows: i64,
    pub output_batches: i64,
    pub input_rows: i64,
    pub input_batches: i64,
    pub build_input_rows: i64,
    pub build_input_batches: i64,
    pub build_mem_used: i64,
    pub build_time_ns: i64,
    pub join_time_ns: i64,
    pub h2d_bytes: i64,
    pub d2h_bytes: i64,
    pub kernel_launches: i64,
    pub nvlink_bytes: i64,
    pub device_mem_used: i64,
}

#[repr(C)]
pub struct dfgpu_expr_node {
    pub kind: i32,
    pub column: i32,
    pub left: i32,
    pub right: i32,
    pub lit_lo: u64,
    pub lit_hi: u64,
    pub lit_precision: i32,
    pub lit_scale: i32,
}

#[repr(C)]
pub struct dfgpu_agg_desc {
    pub func: i32,
    pub name: *const c_char,
    pub arg_col: i32,
    pub filter_col: i32,
}

extern "C" {
    pub fn dfgpu_last_error() -> *const c_char;
    pub fn dfgpu_version() -> *const c_char;
    pub fn dfgpu_hash_contract() -> i32;
    pub fn dfgpu_options_default(opts: *mut dfgpu_options);
    pub fn dfgpu_ctx_create(opts: *const dfgpu_options, out: *mut *mut dfgpu_ctx) -> c_int;
    pub fn dfgpu_ctx_destroy(ctx: *mut dfgpu_ctx);
    pub fn dfgpu_ctx_synchronize(ctx: *mut dfgpu_ctx) -> c_int;
    pub fn dfgpu_ctx_stream(ctx: *mut dfgpu_ctx) -> *mut c_void;
    pub fn dfgpu_ctx_set_profiling(ctx: *mut dfgpu_ctx, enable: i32) -> c_int;
    pub fn dfgpu_ctx_profile_json(ctx: *mut dfgpu_ctx, buf: *mut c_char, cap: i64) -> c_int;
    pub fn dfgpu_ctx_kernel_launches(ctx: *const dfgpu_ctx) -> i64;
    pub fn dfgpu_host_alloc(ctx: *mut dfgpu_ctx, bytes: i64, out: *mut *mut c_void) -> c_int;
    pub fn dfgpu_host_free(ctx: *mut dfgpu_ctx, p: *mut c_void);
    pub fn dfgpu_batch_import(
        ctx: *mut dfgpu_ctx,
        array: *const FFI_ArrowArray,
        schema: *const FFI_ArrowSchema,
        buffers_on_device: i32,
        out: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_batch_export(
        ctx: *mut dfgpu_ctx,
        batch: *const dfgpu_batch,
        out_array: *mut FFI_ArrowArray,
        out_schema: *mut FFI_ArrowSchema,
    ) -> c_int;
    pub fn dfgpu_batch_num_rows(batch: *const dfgpu_batch) -> i64;
    pub fn dfgpu_batch_num_columns(batch: *const dfgpu_batch) -> i32;
    pub fn dfgpu_batch_schema(batch: *const dfgpu_batch, out_schema: *mut FFI_ArrowSchema) -> c_int;
    pub fn dfgpu_batch_concat(
        ctx: *mut dfgpu_ctx,
        batches: *const *mut dfgpu_batch,
        n: i32,
        out: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_batch_slice(
        ctx: *mut dfgpu_ctx,
        batch: *const dfgpu_batch,
        offset: i64,
        length: i64,
        out: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_batch_project(
        ctx: *mut dfgpu_ctx,
        batch: *const dfgpu_batch,
        cols: *const i32,
        n_cols: i32,
        out: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_coalescer_create(
        ctx: *mut dfgpu_ctx,
        schema: *const FFI_ArrowSchema,
        target_rows: i64,
        out: *mut *mut dfgpu_coalescer,
    ) -> c_int;
    pub fn dfgpu_coalescer_push(
        c: *mut dfgpu_coalescer,
        batch: *const FFI_ArrowArray,
        out: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_coalescer_finish(c: *mut dfgpu_coalescer, out: *mut *mut dfgpu_batch) -> c_int;
    pub fn dfgpu_coalescer_destroy(c: *mut dfgpu_coalescer);
    pub fn dfgpu_batch_eval_decimal(
        ctx: *mut dfgpu_ctx,
        batch: *const dfgpu_batch,
        nodes: *const dfgpu_expr_node,
        n_nodes: i32,
        name: *const c_char,
        out: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_batch_filter(
        ctx: *mut dfgpu_ctx,
        batch: *const dfgpu_batch,
        predicate_col: i32,
        out: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_batch_free(batch: *mut dfgpu_batch);
    pub fn dfgpu_create_hashes(
        ctx: *mut dfgpu_ctx,
        batch: *const dfgpu_batch,
        cols: *const i32,
        n_cols: i32,
        seed: u64,
        host_out: *mut u64,
    ) -> c_int;
    pub fn dfgpu_agg_create(
        ctx: *mut dfgpu_ctx,
        mode: i32,
        input_schema: *const FFI_ArrowSchema,
        group_cols: *const i32,
        n_group_cols: i32,
        aggs: *const dfgpu_agg_desc,
        n_aggs: i32,
        out: *mut *mut dfgpu_agg,
    ) -> c_int;
    pub fn dfgpu_agg_output_schema(agg: *const dfgpu_agg, out_schema: *mut FFI_ArrowSchema) -> c_int;
    pub fn dfgpu_agg_push(agg: *mut dfgpu_agg, batch: *const dfgpu_batch) -> c_int;
    pub fn dfgpu_agg_push_arrow(agg: *mut dfgpu_agg, batch: *const FFI_ArrowArray) -> c_int;
    pub fn dfgpu_agg_push_arrow_batches(agg: *mut dfgpu_agg, batches: *const *const FFI_ArrowArray, n: i32) -> c_int;
    pub fn dfgpu_agg_num_groups(agg: *mut dfgpu_agg, out: *mut i64) -> c_int;
    pub fn dfgpu_agg_exchange(agg: *mut dfgpu_agg) -> c_int;
    pub fn dfgpu_agg_exchange_virtual(aggs: *const *mut dfgpu_agg, n: i32) -> c_int;
    pub fn dfgpu_agg_emit(agg: *mut dfgpu_agg, out: *mut *mut dfgpu_batch) -> c_int;
    pub fn dfgpu_agg_emit_arrow(
        agg: *mut dfgpu_agg,
        out_array: *mut FFI_ArrowArray,
        out_schema: *mut FFI_ArrowSchema,
    ) -> c_int;
    pub fn dfgpu_agg_emit_first(agg: *mut dfgpu_agg, n: i64, out: *mut *mut dfgpu_batch) -> c_int;
    pub fn dfgpu_agg_metrics(agg: *const dfgpu_agg, out: *mut dfgpu_metrics) -> c_int;
    pub fn dfgpu_agg_destroy(agg: *mut dfgpu_agg);
    pub fn dfgpu_join_create(
        ctx: *mut dfgpu_ctx,
        join_type: i32,
        partition_mode: i32,
        null_equals_null: i32,
        build_schema: *const FFI_ArrowSchema,
        probe_schema: *const FFI_ArrowSchema,
        build_keys: *const i32,
        probe_keys: *const i32,
        n_keys: i32,
        out: *mut *mut dfgpu_join,
    ) -> c_int;
    pub fn dfgpu_join_output_schema(join: *const dfgpu_join, out_schema: *mut FFI_ArrowSchema) -> c_int;
    pub fn dfgpu_join_build_push(join: *mut dfgpu_join, batch: *const dfgpu_batch) -> c_int;
    pub fn dfgpu_join_build_finish(join: *mut dfgpu_join) -> c_int;
    pub fn dfgpu_join_probe(join: *mut dfgpu_join, probe: *const dfgpu_batch, out: *mut *mut dfgpu_batch) -> c_int;
    pub fn dfgpu_join_set_filter(join: *mut dfgpu_join, sides: *const i32, columns: *const i32, n: i32) -> c_int;
    pub fn dfgpu_join_probe_candidates(
        join: *mut dfgpu_join,
        probe: *const dfgpu_batch,
        out_intermediate: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_join_probe_commit(
        join: *mut dfgpu_join,
        filter_result: *const dfgpu_batch,
        column: i32,
        out: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_join_emit_unmatched(join: *mut dfgpu_join, out: *mut *mut dfgpu_batch) -> c_int;
    pub fn dfgpu_join_metrics(join: *const dfgpu_join, out: *mut dfgpu_metrics) -> c_int;
    pub fn dfgpu_join_destroy(join: *mut dfgpu_join);
    pub fn dfgpu_partition_hash(
        ctx: *mut dfgpu_ctx,
        batch: *const dfgpu_batch,
        key_cols: *const i32,
        n_keys: i32,
        n_parts: i32,
        out_parts: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_exchange_all_to_all(
        ctx: *mut dfgpu_ctx,
        parts: *const *mut dfgpu_batch,
        schema: *const FFI_ArrowSchema,
        out: *mut *mut dfgpu_batch,
    ) -> c_int;
    pub fn dfgpu_partition_i64_noalloc(
        stream: *mut c_void,
        dev_keys: *const i64,
        dev_validity: *const u8,
        n: i64,
        n_parts: i32,
        dev_perm_out: *mut i64,
        dev_offsets_out: *mut i64,
    ) -> c_int;
    pub fn dfgpu_nccl_get_unique_id(out128: *mut c_void) -> c_int;
    pub fn dfgpu_exchange_all_gather(
        ctx: *mut dfgpu_ctx,
        local: *const dfgpu_batch,
        schema: *const FFI_ArrowSchema,
        out: *mut *mut dfgpu_batch,
    ) -> c_int;
}

/// `dfgpu_status` -> `DataFusionError` (SURVEY 8b "Error convention"): the codes mirror the variants the reference's
/// operators raise on this path, the message is the library's thread-local last error.
pub fn check(status: c_int) -> datafusion_common::Result<()> {
    use datafusion_common::DataFusionError as E;
    if status == DFGPU_OK {
        return Ok(());
    }
    let msg = unsafe { std::ffi::CStr::from_ptr(dfgpu_last_error()) }.to_string_lossy().into_owned();
    Err(match status {
        DFGPU_INTERNAL => E::Internal(msg),
        DFGPU_PLAN => E::Plan(msg),
        DFGPU_EXECUTION => E::Execution(msg),
        DFGPU_RESOURCES_EXHAUSTED => E::ResourcesExhausted(msg),
        DFGPU_NOT_IMPLEMENTED => E::NotImplemented(msg),
        _ => E::External(msg.into()), // DFGPU_CUDA, DFGPU_NCCL
    })
}
