/*
 * dfgpu.h — C ABI of the B200-native hash-aggregation / hash-join / hash-repartition path.
 *
 * This is the drop-in boundary for DataFusion 36's hash-table operator path.  Every entry point
 * names the reference interface it replaces (paths relative to /root/reference/datafusion).
 * RecordBatches cross the boundary through the Arrow C Data Interface (include/arrow_c_data.h):
 * inputs are BORROWED for the duration of the call (the library copies them to HBM and never
 * calls `release`); outputs are library-allocated pinned host buffers whose `release` callback
 * frees them (standard producer semantics).
 *
 * Conventions
 *   - every function returns a dfgpu_status; on failure dfgpu_last_error() (thread-local) holds
 *     the message.  Codes mirror the DataFusionError variants used on this path
 *     (common/src/error.rs; SURVEY §8b "Error convention").
 *   - all handles are opaque; *_destroy / *_free accept NULL.
 *   - no exceptions cross the ABI, no torch / C++ types in signatures.
 *   - there is NO CPU fallback: without a CUDA device dfgpu_ctx_create fails with DFGPU_CUDA.
 */
#ifndef DFGPU_H
#define DFGPU_H

#include <stdint.h>
#include "arrow_c_data.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef int32_t dfgpu_status;
enum {
    DFGPU_OK = 0,
    DFGPU_INTERNAL = 1,            /* DataFusionError::Internal            */
    DFGPU_PLAN = 2,                /* DataFusionError::Plan                */
    DFGPU_EXECUTION = 3,           /* DataFusionError::Execution           */
    DFGPU_RESOURCES_EXHAUSTED = 4, /* DataFusionError::ResourcesExhausted  */
    DFGPU_NOT_IMPLEMENTED = 5,     /* DataFusionError::NotImplemented      */
    DFGPU_CUDA = 6,
    DFGPU_NCCL = 7
};

/* AggregateMode — physical-plan/src/aggregates/mod.rs:63-100 */
enum {
    DFGPU_AGG_PARTIAL = 0,
    DFGPU_AGG_FINAL = 1,
    DFGPU_AGG_FINAL_PARTITIONED = 2,
    DFGPU_AGG_SINGLE = 3,
    DFGPU_AGG_SINGLE_PARTITIONED = 4
};
/* the GroupsAccumulator implementations on the path */
enum { DFGPU_SUM = 1, DFGPU_COUNT = 2, DFGPU_AVG = 3, DFGPU_MIN = 4, DFGPU_MAX = 5,
       /* the remaining GroupsAccumulator implementations of the reference —
        * physical-expr/src/aggregate/bit_and_or_xor.rs:116-140, 281-301, 442-462 (integer types) and
        * bool_and_or.rs:137-148, 270-281 (Boolean; BooleanGroupsAccumulator, groups_accumulator/bool_op.rs) */
       DFGPU_BIT_AND = 6, DFGPU_BIT_OR = 7, DFGPU_BIT_XOR = 8, DFGPU_BOOL_AND = 9, DFGPU_BOOL_OR = 10 };
/* JoinType — common/src/join_type.rs:30-47 */
enum {
    DFGPU_JOIN_INNER = 0, DFGPU_JOIN_LEFT = 1, DFGPU_JOIN_RIGHT = 2, DFGPU_JOIN_FULL = 3,
    DFGPU_JOIN_LEFT_SEMI = 4, DFGPU_JOIN_RIGHT_SEMI = 5, DFGPU_JOIN_LEFT_ANTI = 6,
    DFGPU_JOIN_RIGHT_ANTI = 7
};
/* PartitionMode — physical-plan/src/joins/mod.rs:39-47 */
enum { DFGPU_PARTITIONED = 0, DFGPU_COLLECT_LEFT = 1 };

typedef struct dfgpu_ctx dfgpu_ctx;
typedef struct dfgpu_batch dfgpu_batch;
typedef struct dfgpu_agg dfgpu_agg;
typedef struct dfgpu_join dfgpu_join;

/* Options ~ the `gpu.*` ConfigOptions namespace (common/src/config.rs; SURVEY §5 "Config"). */
typedef struct dfgpu_options {
    int32_t device;                /* CUDA ordinal                                             */
    int32_t batch_size;            /* datafusion.execution.batch_size (config.rs:215), 8192    */
    int64_t memory_limit;          /* device bytes an operator may hold; 0 = unlimited
                                      (MemoryPool, execution/src/memory_pool/mod.rs:83-108)    */
    double load_factor;            /* hash-table load factor, default 0.5                      */
    int32_t force_hash_collisions; /* cargo feature force_hash_collisions (hash_utils.rs:307)  */
    int32_t rank, world;           /* this process' GPU rank / number of GPUs (1 = no NCCL)    */
    const void *nccl_unique_id;    /* 128-byte ncclUniqueId shared by all ranks, or NULL       */
    int64_t chunk_rows;            /* rows per internal launch wave, 0 = default (4 Mi)         */
    int64_t coalesce_rows;         /* dfgpu_agg_push_arrow packs host batches smaller than half of this into a pinned
                                      arena and pushes them as one (CoalesceBatchesExec, physical-plan/src/
                                      coalesce_batches.rs, moved to the PCIe boundary); 0 = default (1 Mi rows),
                                      < 0 = off.  Ignored under a memory_limit (per-batch accounting)          */
} dfgpu_options;

typedef struct dfgpu_metrics {     /* BaselineMetrics + BuildProbeJoinMetrics
                                      (physical-plan/src/metrics/baseline.rs:47-67,
                                       joins/utils.rs:1368-1426)                               */
    int64_t elapsed_compute_ns, output_rows, output_batches;
    int64_t input_rows, input_batches;
    int64_t build_input_rows, build_input_batches, build_mem_used, build_time_ns, join_time_ns;
    int64_t h2d_bytes, d2h_bytes, kernel_launches, nvlink_bytes, device_mem_used;
} dfgpu_metrics;

const char *dfgpu_last_error(void);
const char *dfgpu_version(void);
/* Identifies the row hash behind dfgpu_partition_hash / dfgpu_exchange_* / the join ("dfhash v1" = 1).  It is NOT ahash with
 * DataFusion's fixed seeds (repartition/mod.rs:115), so rows land in other partitions than BatchPartitioner would put them:
 * a plan must not pair a stream partitioned by this library with one partitioned by the CPU RepartitionExec (both inputs of
 * a Partitioned HashJoinExec, or Partial -> Final across an exchange).  The shim's rule asserts that both sides of such a
 * pair carry the same contract (rust/datafusion-gpu/src/rule.rs). */
int32_t dfgpu_hash_contract(void);

/* ---- context -------------------------------------------------------------------------- */
void dfgpu_options_default(dfgpu_options *opts);
dfgpu_status dfgpu_ctx_create(const dfgpu_options *opts, dfgpu_ctx **out);
void dfgpu_ctx_destroy(dfgpu_ctx *ctx);
dfgpu_status dfgpu_ctx_synchronize(dfgpu_ctx *ctx);
/* raw CUDA stream (cudaStream_t) every operator of this ctx launches on — for CUDA-event timing */
void *dfgpu_ctx_stream(dfgpu_ctx *ctx);
/* Per-kernel device timing (CUDA events recorded on the launching stream around every hot kernel;
 * the analogue of the reference's per-operator metrics::Time).  profile_json writes
 * [{"name","launches","ms","algo_bytes"},...] into buf and resets the records. */
dfgpu_status dfgpu_ctx_set_profiling(dfgpu_ctx *ctx, int32_t enable);
dfgpu_status dfgpu_ctx_profile_json(dfgpu_ctx *ctx, char *buf, int64_t cap);
/* number of kernels this library launched on ctx so far (bench.py "gpu_launches") */
int64_t dfgpu_ctx_kernel_launches(const dfgpu_ctx *ctx);
/* pinned host memory for callers that want PCIe-speed uploads (bench e2e leg) */
dfgpu_status dfgpu_host_alloc(dfgpu_ctx *ctx, int64_t bytes, void **out);
void dfgpu_host_free(dfgpu_ctx *ctx, void *p);

/* ---- RecordBatch in HBM (arrow RecordBatch; ownership rules of SURVEY §8b) -------------- */
/* `array` must be a struct array (one child per column) as produced by
 * `RecordBatch -> StructArray -> FFI` / pyarrow `RecordBatch._export_to_c`.
 * buffers_on_device != 0: the ArrowArray's buffer pointers are already device pointers on
 * opts.device (borrowed until the batch is freed; no copy) — the roofline benches use this. */
dfgpu_status dfgpu_batch_import(dfgpu_ctx *ctx, const struct ArrowArray *array,
                                const struct ArrowSchema *schema, int32_t buffers_on_device,
                                dfgpu_batch **out);
dfgpu_status dfgpu_batch_export(dfgpu_ctx *ctx, const dfgpu_batch *batch, struct ArrowArray *out_array,
                                struct ArrowSchema *out_schema);
int64_t dfgpu_batch_num_rows(const dfgpu_batch *batch);
int32_t dfgpu_batch_num_columns(const dfgpu_batch *batch);
dfgpu_status dfgpu_batch_schema(const dfgpu_batch *batch, struct ArrowSchema *out_schema);
/* compute::concat_batches (hash_join.rs:764) / CoalesceBatchesExec */
dfgpu_status dfgpu_batch_concat(dfgpu_ctx *ctx, dfgpu_batch *const *batches, int32_t n, dfgpu_batch **out);
/* zero-copy row slice (RecordBatch::slice; row_hash.rs:478-502 output chunking) */
dfgpu_status dfgpu_batch_slice(dfgpu_ctx *ctx, const dfgpu_batch *batch, int64_t offset, int64_t length,
                               dfgpu_batch **out);
/* zero-copy column selection (RecordBatch::project, as ProjectionExec does for plain column
 * expressions: physical-plan/src/projection.rs); columns share the input's device buffers */
dfgpu_status dfgpu_batch_project(dfgpu_ctx *ctx, const dfgpu_batch *batch, const int32_t *cols, int32_t n_cols,
                                 dfgpu_batch **out);
/* CoalesceBatchesExec (physical-plan/src/coalesce_batches.rs:205-262) at the PCIe boundary: host batches
 * are packed into a pinned arena and leave as ONE device batch once `target_rows` rows are buffered
 * (`buffered_rows >= target_batch_size`, :226); a batch of >= target_rows rows arriving at an empty buffer passes
 * through (:215-218); empty batches are discarded (:219).  push sets *out to the finished batch or NULL; finish
 * returns what is left (NULL when nothing is).  Inputs are borrowed for the duration of the call only.  Dictionary
 * columns -> NotImplemented (a dictionary per batch cannot be packed). */
typedef struct dfgpu_coalescer dfgpu_coalescer;
dfgpu_status dfgpu_coalescer_create(dfgpu_ctx *ctx, const struct ArrowSchema *schema, int64_t target_rows,
                                    dfgpu_coalescer **out);
dfgpu_status dfgpu_coalescer_push(dfgpu_coalescer *c, const struct ArrowArray *batch, dfgpu_batch **out);
dfgpu_status dfgpu_coalescer_finish(dfgpu_coalescer *c, dfgpu_batch **out);
void dfgpu_coalescer_destroy(dfgpu_coalescer *c);

/* ProjectionExec arithmetic: BinaryExpr {Plus, Minus, Multiply} over Decimal128 columns and literals,
 * physical-expr/src/expressions/binary.rs:260-275 -> arrow-rs add_wrapping / sub_wrapping / mul_wrapping, which for
 * decimals derive the result type (+/-: scale max(s1,s2), precision min(38, scale + max(p1-s1, p2-s2) + 1);
 * *: precision min(38, p1+p2+1), scale s1+s2) and compute checked in i128 on the valid slots (NULL if either side is).
 * nodes[] is the expression in post order (children before parents, the last node is the root); the planner's casts
 * are already applied (an Int64 literal arrives as Decimal128(20,0)).  The result batch holds the input columns
 * (shared, not copied) plus the root as one more Decimal128 column called `name`.
 * Execution("Arrow error: Compute error: Overflow ...") when an intermediate leaves i128. */
enum { DFGPU_EXPR_COLUMN = 0, DFGPU_EXPR_LITERAL = 1, DFGPU_EXPR_ADD = 2, DFGPU_EXPR_SUB = 3, DFGPU_EXPR_MUL = 4 };
typedef struct dfgpu_expr_node {
    int32_t kind;                 /* DFGPU_EXPR_*                                      */
    int32_t column;               /* COLUMN: column of the batch (Decimal128)          */
    int32_t left, right;          /* ADD / SUB / MUL: indices of earlier nodes         */
    uint64_t lit_lo, lit_hi;      /* LITERAL: unscaled value, two's complement         */
    int32_t lit_precision, lit_scale;
} dfgpu_expr_node;
dfgpu_status dfgpu_batch_eval_decimal(dfgpu_ctx *ctx, const dfgpu_batch *batch, const dfgpu_expr_node *nodes,
                                      int32_t n_nodes, const char *name, dfgpu_batch **out);

/* FilterExec: `batch_filter` -> arrow `filter_record_batch` (physical-plan/src/filter.rs:315-327).
 * predicate_col is the Boolean column the predicate PhysicalExpr evaluated to; the output holds every column of the
 * input (project afterwards) restricted to the rows where the predicate is true and not NULL, in input order. */
dfgpu_status dfgpu_batch_filter(dfgpu_ctx *ctx, const dfgpu_batch *batch, int32_t predicate_col, dfgpu_batch **out);
void dfgpu_batch_free(dfgpu_batch *batch);

/* ---- create_hashes (common/src/hash_utils.rs:357-416) — test hook; seed 0 is the fixed
 * RandomState::with_seeds(0,0,0,0) of joins (hash_join.rs:329) and repartition (repartition/mod.rs:115) */
dfgpu_status dfgpu_create_hashes(dfgpu_ctx *ctx, const dfgpu_batch *batch, const int32_t *cols, int32_t n_cols,
                                 uint64_t seed, uint64_t *host_out);

/* ---- AggregateExec / GroupedHashAggregateStream (aggregates/mod.rs:242-299, row_hash.rs) --- */
typedef struct dfgpu_agg_desc {
    int32_t func;       /* DFGPU_SUM ...                                                       */
    const char *name;   /* display name, e.g. "SUM(l_quantity)"; state fields become name[sum] */
    int32_t arg_col;    /* first-stage modes: input column of the argument, -1 = COUNT(*)/COUNT(1)
                           final modes: ignored — state columns are resolved by position,
                           group_cols + running offset (aggregates/mod.rs:1038-1067)           */
    int32_t filter_col; /* Boolean FILTER column or -1 (first-stage modes only)                */
} dfgpu_agg_desc;

/* input_schema: schema of the batches that will be pushed (raw rows, or the Partial output
 * schema in Final modes).  group_cols: indices of the group-by columns (Final modes: 0..n-1). */
dfgpu_status dfgpu_agg_create(dfgpu_ctx *ctx, int32_t mode, const struct ArrowSchema *input_schema,
                              const int32_t *group_cols, int32_t n_group_cols, const dfgpu_agg_desc *aggs,
                              int32_t n_aggs, dfgpu_agg **out);
/* create_schema (aggregates/mod.rs:749-787): keys ++ state fields (Partial) | result fields */
dfgpu_status dfgpu_agg_output_schema(const dfgpu_agg *agg, struct ArrowSchema *out_schema);
/* group_aggregate_batch (row_hash.rs:524-613): GroupValues::intern + update_batch/merge_batch */
dfgpu_status dfgpu_agg_push(dfgpu_agg *agg, const dfgpu_batch *batch);
dfgpu_status dfgpu_agg_push_arrow(dfgpu_agg *agg, const struct ArrowArray *batch);
/* n host batches in one call: the same result as n dfgpu_agg_push_arrow calls in that order.  The reference's stream
 * hands the operator one 8192-row RecordBatch at a time (row_hash.rs:423-470); a shim that has several ready (a
 * buffered channel, a CoalesceBatchesExec it replaced) passes them together and the library packs them into its
 * pinned arena with several host threads instead of one (one thread packs ~10 GB/s; PCIe 5 takes 55).  The arrays
 * are borrowed for the duration of the call. */
dfgpu_status dfgpu_agg_push_arrow_batches(dfgpu_agg *agg, const struct ArrowArray *const *batches, int32_t n);
dfgpu_status dfgpu_agg_num_groups(dfgpu_agg *agg, int64_t *out);
/* Sharded aggregation, fused: AggregateExec(Partial) -> RepartitionExec(Hash(group keys), world) ->
 * AggregateExec(FinalPartitioned) (core/src/physical_planner.rs:814-849) collapsed into ONE operator per
 * rank.  Collective over the ranks of the context: call it on every rank after the last push and before
 * emit.  The partial states accumulated so far are partitioned by the group-key hash, every rank receives
 * the partitions it owns over NCCL (one contiguous transfer per peer), and the following emit returns the
 * finished groups of the owned keys only - the union over the ranks is the result of the three-operator plan.
 * world == 1: no-op.  DFGPU_NOT_IMPLEMENTED (before any communication, identically on every rank) when the
 * operator does not run the two-level path; the caller then keeps the three-operator plan. */
dfgpu_status dfgpu_agg_exchange(dfgpu_agg *agg);
/* The same exchange between n operators of ONE process and context, aggs[r] playing rank r of a world of n ("virtual
 * ranks": how the reference runs its 4-way partitioned plans inside one process, joins/hash_join.rs:1505-1550 and the
 * target_partitions of core/tests/fuzz_cases/aggregate_fuzz.rs).  The runs move with device-to-device copies instead
 * of NVLink stores; everything else - partition order, offsets, piece tables, the merge of the owned partitions - is
 * the code dfgpu_agg_exchange runs.  Afterwards emit every operator: the union of the outputs is the result. */
dfgpu_status dfgpu_agg_exchange_virtual(dfgpu_agg *const *aggs, int32_t n);
/* emit(EmitTo::All) (row_hash.rs:626-662); resets the operator like the reference's
 * early emit (row_hash.rs:723-734) so it can be called mid-stream in Partial mode. */
dfgpu_status dfgpu_agg_emit(dfgpu_agg *agg, dfgpu_batch **out);
dfgpu_status dfgpu_agg_emit_arrow(dfgpu_agg *agg, struct ArrowArray *out_array, struct ArrowSchema *out_schema);
/* emit(EmitTo::First(n)) (expr/src/groups_accumulator.rs:24-58; the early emit of row_hash.rs:723-734, where n is a multiple
 * of batch_size): n of the groups accumulated so far leave the operator (fewer when fewer remain), the others stay queued for
 * the next call; rows pushed in between start new partial states.  Partial mode only - there a key may be emitted more than
 * once and the Final stage re-merges.  Group ids on the device are dense but not first-seen ordered, so which n groups come
 * out is the operator's choice; streaming emission over sorted input (aggregates/order/) is not offered
 * (DFGPU_NOT_IMPLEMENTED outside Partial mode) and such plans stay on the CPU.  dfgpu_agg_emit returns whatever is left. */
dfgpu_status dfgpu_agg_emit_first(dfgpu_agg *agg, int64_t n, dfgpu_batch **out);
dfgpu_status dfgpu_agg_metrics(const dfgpu_agg *agg, dfgpu_metrics *out);
void dfgpu_agg_destroy(dfgpu_agg *agg);

/* ---- HashJoinExec (joins/hash_join.rs:272-356, 505-674); left = build, right = probe ------ */
dfgpu_status dfgpu_join_create(dfgpu_ctx *ctx, int32_t join_type, int32_t partition_mode, int32_t null_equals_null,
                               const struct ArrowSchema *build_schema, const struct ArrowSchema *probe_schema,
                               const int32_t *build_keys, const int32_t *probe_keys, int32_t n_keys,
                               dfgpu_join **out);
/* build_join_schema (joins/utils.rs:657-729) */
dfgpu_status dfgpu_join_output_schema(const dfgpu_join *join, struct ArrowSchema *out_schema);
/* collect_left_input (hash_join.rs:678-768): push all build batches, then finish */
dfgpu_status dfgpu_join_build_push(dfgpu_join *join, const dfgpu_batch *batch);
dfgpu_status dfgpu_join_build_finish(dfgpu_join *join);
/* process_probe_batch (hash_join.rs:1238-1343).  After build_finish the table is read-only;
 * the visited bitmap is updated with atomics.  Output holds every row the reference emits
 * for this probe batch (all its <=batch_size sub-batches concatenated). */
dfgpu_status dfgpu_join_probe(dfgpu_join *join, const dfgpu_batch *probe, dfgpu_batch **out);
/* Residual JoinFilter (joins/utils.rs:1143-1176; JoinFilter = expression + column_indices + intermediate schema).  The
 * reference applies it to the candidate pairs of the equi-join BEFORE it marks build rows visited and BEFORE
 * adjust_indices_by_join_type (hash_join.rs:1259-1311), so for every join type but Inner it decides which rows count as
 * matched and cannot be applied to the joined output afterwards.  The predicate is a PhysicalExpr and is evaluated by the
 * host engine, hence three calls:
 *   set_filter        once, before the build side is pushed: filter.column_indices as (side, column) pairs, side 0 = left
 *                     / build, 1 = right / probe (JoinSide, common/src/join_type.rs:111-116)
 *   probe_candidates  instead of dfgpu_join_probe: returns the intermediate batch of those columns, one row per candidate
 *                     pair (what apply_join_filter_to_indices builds with build_batch_from_indices, utils.rs:1154-1162)
 *   probe_commit      the Boolean column the predicate evaluated to on that batch (NULL = false, arrow `filter`): filters
 *                     the pairs, marks visited build rows, applies the join type's fix-ups over the probe batch and returns
 *                     its output rows.  Exactly one commit per probe_candidates. */
dfgpu_status dfgpu_join_set_filter(dfgpu_join *join, const int32_t *sides, const int32_t *columns, int32_t n);
dfgpu_status dfgpu_join_probe_candidates(dfgpu_join *join, const dfgpu_batch *probe, dfgpu_batch **out_intermediate);
dfgpu_status dfgpu_join_probe_commit(dfgpu_join *join, const dfgpu_batch *filter_result, int32_t column, dfgpu_batch **out);
/* process_unmatched_build_batch (hash_join.rs:1348-1388); *out = NULL when the join type
 * produces nothing at the end (Inner, Right, RightSemi, RightAnti). */
dfgpu_status dfgpu_join_emit_unmatched(dfgpu_join *join, dfgpu_batch **out);
dfgpu_status dfgpu_join_metrics(const dfgpu_join *join, dfgpu_metrics *out);
void dfgpu_join_destroy(dfgpu_join *join);

/* ---- RepartitionExec(Hash) / BatchPartitioner (repartition/mod.rs:133-217, 442-580) -------- */
/* out_parts[n_parts]; entries for empty partitions are NULL (empty batches are never sent,
 * repartition/mod.rs:196-214).  Row order inside a partition follows input order. */
dfgpu_status dfgpu_partition_hash(dfgpu_ctx *ctx, const dfgpu_batch *batch, const int32_t *key_cols,
                                  int32_t n_keys, int32_t n_parts, dfgpu_batch **out_parts);
/* The exchange itself: parts[world] (NULL = nothing for that rank) -> concatenation of what
 * every rank sent to this rank.  NCCL all-to-all over NVLink; world == 1 degenerates to a move.
 * *out may be NULL when nothing was received.  `schema` describes the batches (needed when all
 * of this rank's parts are NULL). */
dfgpu_status dfgpu_exchange_all_to_all(dfgpu_ctx *ctx, dfgpu_batch *const *parts, const struct ArrowSchema *schema,
                                       dfgpu_batch **out);
/* Allocation-free partitioning.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * It is the partitioning step of dfgpu_partition_hash for ONE Int64 key column given as raw device pointers
 * (dev_validity: Arrow validity bitmap of the keys with bit offset 0, or NULL when there are no NULL keys): on
 * return dev_perm_out[n] holds the row numbers grouped by partition = hash % n_parts, input order kept inside a
 * partition, and dev_offsets_out[n_parts + 1] the start of every partition in it (the rows of partition p are
 * dev_perm_out[offsets[p] .. offsets[p + 1])) - the same rows in the same order as dfgpu_partition_hash's out_parts.
 * The scratch memory (partition ids, per-block histograms, scanned write positions; ~5 MiB) is a file-scope device
 * array that every GPU that loads the library holds until exit; the call itself performs no cudaMalloc / cudaFree /
 * pool allocation.  Limits: n <= DFGPU_NOALLOC_MAX_ROWS and n_parts <= DFGPU_NOALLOC_MAX_PARTS per call, otherwise
 * DFGPU_NOT_IMPLEMENTED (split the batch, or use dfgpu_partition_hash).  The work is enqueued on `stream`
 * (a cudaStream_t of the calling thread's current device; NULL = the default stream) and NOT synchronised: outputs
 * are valid once the stream reaches that point.  All calls on a device share the one scratch array: calls that
 * overlap on the GPU (two streams, two host threads) race - serialise them on one stream.  The first call on a
 * device resolves the array's device address (cudaGetSymbolAddress); make it outside a graph capture. */
#define DFGPU_NOALLOC_MAX_ROWS (1 << 20)
#define DFGPU_NOALLOC_MAX_PARTS 256
dfgpu_status dfgpu_partition_i64_noalloc(void *stream, const int64_t *dev_keys, const uint8_t *dev_validity, int64_t n,
                                         int32_t n_parts, int64_t *dev_perm_out, int64_t *dev_offsets_out);
/* 128-byte ncclUniqueId for dfgpu_options.nccl_unique_id: rank 0 creates it and shares it with
 * the other ranks out of band (the reference has no analogue: its exchange is in-process). */
dfgpu_status dfgpu_nccl_get_unique_id(void *out128);
/* CollectLeft build-side replication (hash_join.rs:594-606): all-gather of a batch */
dfgpu_status dfgpu_exchange_all_gather(dfgpu_ctx *ctx, const dfgpu_batch *local, const struct ArrowSchema *schema,
                                       dfgpu_batch **out);

#ifdef __cplusplus
}
#endif
#endif /* DFGPU_H */
