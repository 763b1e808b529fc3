"""Host-side mirror of the reference's operator interface for the hash-table path.

Names and argument meaning follow DataFusion 36 (paths relative to /root/reference/datafusion):
``AggregateExec`` / ``AggregateMode`` (physical-plan/src/aggregates/mod.rs:63-100, 242-299),
``HashJoinExec`` / ``PartitionMode`` (physical-plan/src/joins/hash_join.rs:272-356, joins/mod.rs:39-47),
``JoinType`` (common/src/join_type.rs:30-47), ``RepartitionExec(Partitioning::Hash)``
(physical-plan/src/repartition/mod.rs:295-580).  Every operator here is a thin driver over the
C ABI (include/dfgpu.h); all compute happens in libdfgpu's CUDA kernels.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from enum import IntEnum
from typing import Iterable, List, Optional, Sequence, Union

import numpy as np
import pyarrow as pa

from . import _native as N


class AggregateMode(IntEnum):
    Partial = 0
    Final = 1
    FinalPartitioned = 2
    Single = 3
    SinglePartitioned = 4


class AggFunc(IntEnum):
    Sum = 1
    Count = 2
    Avg = 3
    Min = 4
    Max = 5
    BitAnd = 6      # bit_and_or_xor.rs (integer arguments)
    BitOr = 7
    BitXor = 8
    BoolAnd = 9     # bool_and_or.rs (Boolean argument, BooleanGroupsAccumulator)
    BoolOr = 10


class JoinType(IntEnum):
    Inner = 0
    Left = 1
    Right = 2
    Full = 3
    LeftSemi = 4
    RightSemi = 5
    LeftAnti = 6
    RightAnti = 7


class PartitionMode(IntEnum):
    Partitioned = 0
    CollectLeft = 1


class GpuContext:
    """Owns one CUDA device/stream (+ NCCL communicator when world > 1) — the TaskContext analogue."""

    def __init__(self, device: int = 0, batch_size: int = 8192, memory_limit: int = 0, load_factor: float = 0.5,
                 force_hash_collisions: bool = False, rank: int = 0, world: int = 1, nccl_unique_id: Optional[bytes] = None,
                 chunk_rows: int = 0, coalesce_rows: int = 0):
        L = N.lib()
        o = N.Options()
        L.dfgpu_options_default(C.byref(o))
        o.device, o.batch_size, o.memory_limit, o.load_factor = device, batch_size, memory_limit, load_factor
        o.force_hash_collisions, o.rank, o.world, o.chunk_rows = int(force_hash_collisions), rank, world, chunk_rows
        o.coalesce_rows = coalesce_rows
        self._uid = None
        if nccl_unique_id is not None:
            self._uid = C.create_string_buffer(nccl_unique_id, 128)
            o.nccl_unique_id = C.cast(self._uid, C.c_void_p)
        h = C.c_void_p()
        N.check(L.dfgpu_ctx_create(C.byref(o), C.byref(h)))
        self.h = h
        self.batch_size = batch_size
        self.rank, self.world = rank, world

    def close(self):
        if getattr(self, "h", None):
            N.lib().dfgpu_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def synchronize(self):
        N.check(N.lib().dfgpu_ctx_synchronize(self.h))

    def set_profiling(self, enable: bool):
        N.check(N.lib().dfgpu_ctx_set_profiling(self.h, int(enable)))

    def profile(self) -> list:
        """Per-kernel CUDA-event times since the last call: [{name, launches, ms, algo_bytes}]."""
        import json
        buf = C.create_string_buffer(1 << 16)
        N.check(N.lib().dfgpu_ctx_profile_json(self.h, buf, len(buf)))
        return json.loads(buf.value.decode())

    @property
    def stream(self) -> int:
        return N.lib().dfgpu_ctx_stream(self.h)

    @property
    def kernel_launches(self) -> int:
        return N.lib().dfgpu_ctx_kernel_launches(self.h)

    def host_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        N.check(N.lib().dfgpu_host_alloc(self.h, nbytes, C.byref(p)))
        return p.value

    def host_free(self, ptr: int):
        N.lib().dfgpu_host_free(self.h, ptr)

    @staticmethod
    def nccl_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        N.check(N.lib().dfgpu_nccl_get_unique_id(buf))
        return buf.raw


class DeviceBatch:
    """A RecordBatch resident in HBM (dfgpu_batch)."""

    def __init__(self, ctx: GpuContext, handle, keep=None):
        self.ctx, self.h, self._keep = ctx, handle, keep

    @classmethod
    def from_arrow(cls, ctx: GpuContext, batch: pa.RecordBatch) -> "DeviceBatch":
        a, s = N.ArrowArray(), N.ArrowSchema()
        batch._export_to_c(C.addressof(a), C.addressof(s))
        h = C.c_void_p()
        try:
            N.check(N.lib().dfgpu_batch_import(ctx.h, C.addressof(a), C.addressof(s), 0, C.byref(h)))
        finally:
            # inputs are borrowed by the library: the producer (us) releases them
            if a.release:
                C.CFUNCTYPE(None, C.POINTER(N.ArrowArray))(a.release)(C.byref(a))
            N.release_schema(s)
        return cls(ctx, h)

    @classmethod
    def from_device(cls, ctx: GpuContext, schema: pa.Schema, columns: Sequence[dict], num_rows: int, keep=None) -> "DeviceBatch":
        """Wrap buffers that already live on the device.  columns[i] = {"buffers": [ptr|None,...],
        "null_count": int}; pointers are borrowed until the batch is freed (pass owners in keep)."""
        kids, hold = [], []
        for col in columns:
            bufs = (C.c_void_p * len(col["buffers"]))(*[b if b else None for b in col["buffers"]])
            a = N.ArrowArray()
            a.length, a.null_count, a.offset = num_rows, col.get("null_count", 0), 0
            a.n_buffers, a.buffers = len(col["buffers"]), C.cast(bufs, C.POINTER(C.c_void_p))
            hold += [bufs, a]
            kids.append(a)
        kp = (C.POINTER(N.ArrowArray) * len(kids))(*[C.pointer(k) for k in kids])
        top = N.ArrowArray()
        topbuf = (C.c_void_p * 1)(None)
        top.length, top.null_count, top.offset, top.n_buffers = num_rows, 0, 0, 1
        top.buffers = C.cast(topbuf, C.POINTER(C.c_void_p))
        top.n_children, top.children = len(kids), C.cast(kp, C.POINTER(C.POINTER(N.ArrowArray)))
        s = N.export_schema(schema)
        h = C.c_void_p()
        try:
            N.check(N.lib().dfgpu_batch_import(ctx.h, C.addressof(top), C.addressof(s), 1, C.byref(h)))
        finally:
            N.release_schema(s)
        return cls(ctx, h, keep=keep)

    def to_arrow(self) -> pa.RecordBatch:
        a, s = N.ArrowArray(), N.ArrowSchema()
        N.check(N.lib().dfgpu_batch_export(self.ctx.h, self.h, C.addressof(a), C.addressof(s)))
        return pa.RecordBatch._import_from_c(C.addressof(a), C.addressof(s))

    @property
    def num_rows(self) -> int:
        return N.lib().dfgpu_batch_num_rows(self.h)

    @property
    def schema(self) -> pa.Schema:
        s = N.ArrowSchema()
        N.check(N.lib().dfgpu_batch_schema(self.h, C.addressof(s)))
        return N.import_schema(s)

    def slice(self, offset: int, length: int) -> "DeviceBatch":
        h = C.c_void_p()
        N.check(N.lib().dfgpu_batch_slice(self.ctx.h, self.h, offset, length, C.byref(h)))
        return DeviceBatch(self.ctx, h, keep=self)

    def project(self, columns: Sequence[Union[int, str]]) -> "DeviceBatch":
        """RecordBatch::project: zero-copy column selection (what ProjectionExec does for plain columns)."""
        sch = self.schema
        idx = [_col_index(sch, c) for c in columns]
        h = C.c_void_p()
        N.check(N.lib().dfgpu_batch_project(self.ctx.h, self.h, (C.c_int32 * max(1, len(idx)))(*idx), len(idx), C.byref(h)))
        return DeviceBatch(self.ctx, h, keep=self)

    def eval_decimal(self, nodes: Sequence[tuple], name: str) -> "DeviceBatch":
        """ProjectionExec arithmetic (BinaryExpr Plus / Minus / Multiply over Decimal128, binary.rs:260-275): nodes in post
        order - ("col", column) | ("lit", unscaled int, precision, scale) | ("add" | "sub" | "mul", left node, right node) -
        the last node is the root.  Returns this batch plus the root as one more column `name`."""
        kinds = {"col": 0, "lit": 1, "add": 2, "sub": 3, "mul": 4}
        arr = (N.ExprNode * max(1, len(nodes)))()
        sch = self.schema
        for i, nd in enumerate(nodes):
            arr[i].kind = kinds[nd[0]]
            if nd[0] == "col":
                arr[i].column = _col_index(sch, nd[1])
            elif nd[0] == "lit":
                v = nd[1] & ((1 << 128) - 1)
                arr[i].lit_lo, arr[i].lit_hi = v & ((1 << 64) - 1), v >> 64
                arr[i].lit_precision, arr[i].lit_scale = nd[2], nd[3]
            else:
                arr[i].left, arr[i].right = nd[1], nd[2]
        h = C.c_void_p()
        N.check(N.lib().dfgpu_batch_eval_decimal(self.ctx.h, self.h, arr, len(nodes), name.encode(), C.byref(h)))
        return DeviceBatch(self.ctx, h, keep=self)

    def filter(self, predicate: Union[int, str]) -> "DeviceBatch":
        """FilterExec::batch_filter (filter.rs:315-327): rows where the Boolean column `predicate` is true and not NULL."""
        h = C.c_void_p()
        N.check(N.lib().dfgpu_batch_filter(self.ctx.h, self.h, _col_index(self.schema, predicate), C.byref(h)))
        return DeviceBatch(self.ctx, h)

    @staticmethod
    def concat(ctx: GpuContext, batches: Sequence["DeviceBatch"]) -> "DeviceBatch":
        arr = (C.c_void_p * len(batches))(*[b.h for b in batches])
        h = C.c_void_p()
        N.check(N.lib().dfgpu_batch_concat(ctx.h, arr, len(batches), C.byref(h)))
        return DeviceBatch(ctx, h, keep=list(batches))

    def free(self):
        if getattr(self, "h", None):
            N.lib().dfgpu_batch_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


BatchLike = Union[pa.RecordBatch, DeviceBatch]


def _dev(ctx: GpuContext, b: BatchLike) -> DeviceBatch:
    return b if isinstance(b, DeviceBatch) else DeviceBatch.from_arrow(ctx, b)


def _col_index(schema: pa.Schema, c) -> int:
    if c is None:
        return -1
    if isinstance(c, str):
        idx = schema.get_all_field_indices(c)
        if len(idx) != 1:
            raise N.DfGpuError(2, f"column '{c}' not found or ambiguous")
        return idx[0]
    return int(c)


def slice_batches(batch: pa.RecordBatch, batch_size: int) -> List[pa.RecordBatch]:
    """Output chunking of the ProducingOutput state (row_hash.rs:478-502)."""
    if batch.num_rows <= batch_size:
        return [batch]
    return [batch.slice(o, min(batch_size, batch.num_rows - o)) for o in range(0, batch.num_rows, batch_size)]


def create_hashes(ctx: GpuContext, batch: BatchLike, cols: Sequence[int], seed: int = 0) -> np.ndarray:
    """common/src/hash_utils.rs:357-416 (test hook)."""
    d = _dev(ctx, batch)
    out = np.zeros(d.num_rows, dtype=np.uint64)
    arr = (C.c_int32 * len(cols))(*cols)
    N.check(N.lib().dfgpu_create_hashes(ctx.h, d.h, arr, len(cols), seed, out.ctypes.data))
    return out


@dataclass
class AggregateExpr:
    """An AggregateExpr on the GroupsAccumulator path: SUM / COUNT / AVG / MIN / MAX."""
    func: AggFunc
    arg: Union[int, str, None]          # input column; None = COUNT(*) (arrives as COUNT(UInt8(1)))
    name: str                           # e.g. "SUM(l_quantity)"
    filter: Union[int, str, None] = None


class GpuAggregateExec:
    """Drop-in for AggregateExec on the hash path (GroupedHashAggregateStream, row_hash.rs)."""

    def __init__(self, ctx: GpuContext, mode: AggregateMode, input_schema: pa.Schema, group_by: Sequence[Union[int, str]],
                 aggr_expr: Sequence[AggregateExpr]):
        self.ctx, self.mode, self.input_schema = ctx, AggregateMode(mode), input_schema
        L = N.lib()
        gcols = [_col_index(input_schema, g) for g in group_by]
        garr = (C.c_int32 * max(1, len(gcols)))(*gcols)
        descs = (N.AggDesc * max(1, len(aggr_expr)))()
        self._names = []
        for i, e in enumerate(aggr_expr):
            nm = e.name.encode()
            self._names.append(nm)
            descs[i].func, descs[i].name = int(e.func), nm
            descs[i].arg_col = _col_index(input_schema, e.arg)
            descs[i].filter_col = _col_index(input_schema, e.filter)
        s = N.export_schema(input_schema)
        h = C.c_void_p()
        try:
            N.check(L.dfgpu_agg_create(ctx.h, int(self.mode), C.addressof(s), garr, len(gcols), descs, len(aggr_expr), C.byref(h)))
        finally:
            N.release_schema(s)
        self.h = h
        o = N.ArrowSchema()
        N.check(L.dfgpu_agg_output_schema(self.h, C.addressof(o)))
        self.schema = N.import_schema(o)

    def push(self, batch: BatchLike):
        """group_aggregate_batch (row_hash.rs:524-613)."""
        if isinstance(batch, DeviceBatch):
            N.check(N.lib().dfgpu_agg_push(self.h, batch.h))
        else:
            # host Arrow batch: the call the Rust shim makes (dfgpu_agg_push_arrow); small batches are packed into a
            # pinned arena by the library and cross PCIe as one (dfgpu_options.coalesce_rows)
            if batch.schema.types != self.input_schema.types:
                raise N.DfGpuError(1, f"Internal error: pushed batch schema {batch.schema.types} != operator input schema {self.input_schema.types}")
            a, s = N.ArrowArray(), N.ArrowSchema()
            batch._export_to_c(C.addressof(a), C.addressof(s))
            try:
                N.check(N.lib().dfgpu_agg_push_arrow(self.h, C.addressof(a)))
            finally:
                if a.release:
                    C.CFUNCTYPE(None, C.POINTER(N.ArrowArray))(a.release)(C.byref(a))
                N.release_schema(s)

    def export_batches(self, batches: Sequence[pa.RecordBatch]):
        """Arrow C Data Interface structs of host batches, ready for push_exported (what a Rust shim holds anyway: its
        RecordBatches are FFI-exportable without a copy).  Keep the returned object alive until push_exported returned."""
        arrs = []
        for b in batches:
            if b.schema.types != self.input_schema.types:
                raise N.DfGpuError(1, f"Internal error: pushed batch schema {b.schema.types} != operator input schema {self.input_schema.types}")
            a, s = N.ArrowArray(), N.ArrowSchema()
            b._export_to_c(C.addressof(a), C.addressof(s))
            N.release_schema(s)
            arrs.append(a)
        ptrs = (C.c_void_p * len(arrs))(*[C.addressof(a) for a in arrs])
        return arrs, ptrs

    def push_exported(self, exported):
        """dfgpu_agg_push_arrow_batches: all batches of export_batches() in one call."""
        arrs, ptrs = exported
        N.check(N.lib().dfgpu_agg_push_arrow_batches(self.h, ptrs, len(arrs)))

    @staticmethod
    def release_exported(exported):
        arrs, _ = exported
        for a in arrs:
            if a.release:
                C.CFUNCTYPE(None, C.POINTER(N.ArrowArray))(a.release)(C.byref(a))

    def push_many(self, batches: Sequence[pa.RecordBatch]):
        """n host batches in one ABI call (dfgpu_agg_push_arrow_batches): same result as push() on each, in order."""
        ex = self.export_batches(batches)
        try:
            self.push_exported(ex)
        finally:
            self.release_exported(ex)

    def exchange(self):
        """Collective: fused Partial -> RepartitionExec(Hash) -> FinalPartitioned (dfgpu_agg_exchange).  After it,
        emit() returns the finished groups of the keys this rank owns.  Raises DfGpuError(NotImplemented) when the
        operator cannot exchange at the record level (the caller keeps the three-operator plan)."""
        N.check(N.lib().dfgpu_agg_exchange(self.h))

    @staticmethod
    def exchange_virtual(ops: Sequence["GpuAggregateExec"]):
        """The same exchange between operators of one process, ops[r] playing rank r (dfgpu_agg_exchange_virtual): the
        reference's way of testing partitioned plans in one process (hash_join.rs:1505-1550)."""
        arr = (C.c_void_p * len(ops))(*[o.h for o in ops])
        N.check(N.lib().dfgpu_agg_exchange_virtual(arr, len(ops)))

    def num_groups(self) -> int:
        n = C.c_int64()
        N.check(N.lib().dfgpu_agg_num_groups(self.h, C.byref(n)))
        return n.value

    def emit_device(self) -> DeviceBatch:
        h = C.c_void_p()
        N.check(N.lib().dfgpu_agg_emit(self.h, C.byref(h)))
        return DeviceBatch(self.ctx, h)

    def emit_first(self, n: int) -> pa.RecordBatch:
        """emit(EmitTo::First(n)) (groups_accumulator.rs:24-58): n groups leave the operator, Partial mode only."""
        h = C.c_void_p()
        N.check(N.lib().dfgpu_agg_emit_first(self.h, n, C.byref(h)))
        return DeviceBatch(self.ctx, h).to_arrow()

    def emit(self) -> pa.RecordBatch:
        """emit(EmitTo::All) (row_hash.rs:626-662)."""
        return self.emit_device().to_arrow()

    def execute(self, batches: Iterable[BatchLike]) -> List[pa.RecordBatch]:
        """One partition's stream: consume the input, then produce batch_size-sized output batches."""
        for b in batches:
            self.push(b)
        out = self.emit()
        if out.num_rows == 0:
            return []            # empty input => empty output (row_hash.rs:632-634)
        return slice_batches(out, self.ctx.batch_size)

    def metrics(self) -> dict:
        m = N.Metrics()
        N.check(N.lib().dfgpu_agg_metrics(self.h, C.byref(m)))
        return m.as_dict()

    def close(self):
        if getattr(self, "h", None):
            N.lib().dfgpu_agg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class GpuHashJoinExec:
    """Drop-in for HashJoinExec: left = build side, right = probe side (hash_join.rs:404-407)."""

    def __init__(self, ctx: GpuContext, left_schema: pa.Schema, right_schema: pa.Schema, on: Sequence[tuple],
                 join_type: JoinType, partition_mode: PartitionMode = PartitionMode.CollectLeft, null_equals_null: bool = False):
        self.ctx, self.join_type = ctx, JoinType(join_type)
        self.left_schema, self.right_schema = left_schema, right_schema
        lk = [_col_index(left_schema, l) for l, _ in on]
        rk = [_col_index(right_schema, r) for _, r in on]
        ls, rs = N.export_schema(left_schema), N.export_schema(right_schema)
        h = C.c_void_p()
        try:
            N.check(N.lib().dfgpu_join_create(ctx.h, int(join_type), int(partition_mode), int(null_equals_null), C.addressof(ls),
                                              C.addressof(rs), (C.c_int32 * max(1, len(lk)))(*lk), (C.c_int32 * max(1, len(rk)))(*rk),
                                              len(lk), C.byref(h)))
        finally:
            N.release_schema(ls)
            N.release_schema(rs)
        self.h = h
        self._predicate = None
        o = N.ArrowSchema()
        N.check(N.lib().dfgpu_join_output_schema(self.h, C.addressof(o)))
        self.schema = N.import_schema(o)

    def set_filter(self, column_indices: Sequence[tuple], predicate):
        """Residual JoinFilter (joins/utils.rs:1143-1176): column_indices = [(side, column)] with side "left" | "right"; predicate is the
        host engine's PhysicalExpr: RecordBatch of the intermediate columns -> Boolean array.  Call before build()."""
        sides = (C.c_int32 * len(column_indices))(*[0 if s == "left" else 1 for s, _ in column_indices])
        cols = (C.c_int32 * len(column_indices))(*[int(i) for _, i in column_indices])
        N.check(N.lib().dfgpu_join_set_filter(self.h, sides, cols, len(column_indices)))
        self._predicate = predicate

    def build(self, batches: Iterable[BatchLike]):
        """collect_left_input (hash_join.rs:678-768)."""
        for b in batches:
            d = _dev(self.ctx, b)
            N.check(N.lib().dfgpu_join_build_push(self.h, d.h))
        N.check(N.lib().dfgpu_join_build_finish(self.h))

    def probe_device(self, batch: BatchLike) -> DeviceBatch:
        d = _dev(self.ctx, batch)
        h = C.c_void_p()
        if getattr(self, "_predicate", None) is not None:
            # candidate pairs -> intermediate batch -> the host evaluates the predicate -> commit (hash_join.rs:1259-1311)
            N.check(N.lib().dfgpu_join_probe_candidates(self.h, d.h, C.byref(h)))
            inter = DeviceBatch(self.ctx, h).to_arrow()
            mask = self._predicate(inter)
            mask = pa.array(mask) if not isinstance(mask, (pa.Array, pa.ChunkedArray)) else mask
            if isinstance(mask, pa.ChunkedArray):
                mask = mask.combine_chunks() if mask.num_chunks else pa.array([], pa.bool_())
            md = DeviceBatch.from_arrow(self.ctx, pa.record_batch([mask.cast(pa.bool_())], names=["filter"]))
            h2 = C.c_void_p()
            N.check(N.lib().dfgpu_join_probe_commit(self.h, md.h, 0, C.byref(h2)))
            return DeviceBatch(self.ctx, h2)
        N.check(N.lib().dfgpu_join_probe(self.h, d.h, C.byref(h)))
        return DeviceBatch(self.ctx, h)

    def probe(self, batch: BatchLike) -> pa.RecordBatch:
        """process_probe_batch (hash_join.rs:1238-1343)."""
        return self.probe_device(batch).to_arrow()

    def finish_device(self) -> Optional[DeviceBatch]:
        h = C.c_void_p()
        N.check(N.lib().dfgpu_join_emit_unmatched(self.h, C.byref(h)))
        return DeviceBatch(self.ctx, h) if h.value else None

    def finish(self) -> Optional[pa.RecordBatch]:
        """process_unmatched_build_batch (hash_join.rs:1348-1388)."""
        d = self.finish_device()
        return d.to_arrow() if d is not None else None

    def execute(self, left: Iterable[BatchLike], right: Iterable[BatchLike]) -> List[pa.RecordBatch]:
        self.build(left)
        out = [self.probe(b) for b in right]
        f = self.finish()
        if f is not None:
            out.append(f)
        res = []
        for b in out:
            res += slice_batches(b, self.ctx.batch_size)
        return res

    def metrics(self) -> dict:
        m = N.Metrics()
        N.check(N.lib().dfgpu_join_metrics(self.h, C.byref(m)))
        return m.as_dict()

    def close(self):
        if getattr(self, "h", None):
            N.lib().dfgpu_join_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class GpuCoalesceBatchesExec:
    """CoalesceBatchesExec (physical-plan/src/coalesce_batches.rs) at the PCIe boundary: host RecordBatches in, device
    batches of >= target_batch_size rows out (the last one may be smaller).  Same buffering rule as the reference:
    emit when buffered_rows >= target, pass a large batch through an empty buffer, drop empty batches."""

    def __init__(self, ctx: GpuContext, schema: pa.Schema, target_batch_size: int):
        self.ctx, self.schema, self.target_batch_size = ctx, schema, target_batch_size
        s = N.export_schema(schema)
        h = C.c_void_p()
        try:
            N.check(N.lib().dfgpu_coalescer_create(ctx.h, C.addressof(s), target_batch_size, C.byref(h)))
        finally:
            N.release_schema(s)
        self.h = h

    def push(self, batch: pa.RecordBatch) -> Optional[DeviceBatch]:
        if batch.schema.types != self.schema.types:
            raise N.DfGpuError(1, f"Internal error: pushed batch schema {batch.schema.types} != {self.schema.types}")
        a, s = N.ArrowArray(), N.ArrowSchema()
        batch._export_to_c(C.addressof(a), C.addressof(s))
        out = C.c_void_p()
        try:
            N.check(N.lib().dfgpu_coalescer_push(self.h, C.addressof(a), C.byref(out)))
        finally:
            if a.release:
                C.CFUNCTYPE(None, C.POINTER(N.ArrowArray))(a.release)(C.byref(a))
            N.release_schema(s)
        return DeviceBatch(self.ctx, out) if out.value else None

    def finish(self) -> Optional[DeviceBatch]:
        out = C.c_void_p()
        N.check(N.lib().dfgpu_coalescer_finish(self.h, C.byref(out)))
        return DeviceBatch(self.ctx, out) if out.value else None

    def execute(self, batches: Sequence[pa.RecordBatch]) -> List[DeviceBatch]:
        out = [d for d in (self.push(b) for b in batches) if d is not None]
        last = self.finish()
        return out + ([last] if last is not None else [])

    def close(self):
        if getattr(self, "h", None):
            N.lib().dfgpu_coalescer_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class GpuRepartitionExec:
    """RepartitionExec with Partitioning::Hash(exprs, n) (repartition/mod.rs:295-580)."""

    def __init__(self, ctx: GpuContext, schema: pa.Schema, hash_exprs: Sequence[Union[int, str]], num_partitions: int):
        self.ctx, self.schema, self.n = ctx, schema, num_partitions
        self.keys = [_col_index(schema, k) for k in hash_exprs]

    def partition_device(self, batch: BatchLike) -> List[Optional[DeviceBatch]]:
        d = _dev(self.ctx, batch)
        out = (C.c_void_p * self.n)()
        keys = (C.c_int32 * len(self.keys))(*self.keys)
        N.check(N.lib().dfgpu_partition_hash(self.ctx.h, d.h, keys, len(self.keys), self.n, out))
        return [DeviceBatch(self.ctx, C.c_void_p(p)) if p else None for p in out]

    def partition(self, batch: BatchLike) -> List[Optional[pa.RecordBatch]]:
        """BatchPartitioner::partition (repartition/mod.rs:133-217)."""
        return [p.to_arrow() if p is not None else None for p in self.partition_device(batch)]

    def exchange(self, parts: Sequence[Optional[DeviceBatch]]) -> Optional[DeviceBatch]:
        """The channel hop of RepartitionExec as an NCCL all-to-all: parts[r] goes to rank r."""
        arr = (C.c_void_p * len(parts))(*[p.h if p is not None else None for p in parts])
        s = N.export_schema(self.schema)
        h = C.c_void_p()
        try:
            N.check(N.lib().dfgpu_exchange_all_to_all(self.ctx.h, arr, C.addressof(s), C.byref(h)))
        finally:
            N.release_schema(s)
        return DeviceBatch(self.ctx, h) if h.value else None


def all_gather(ctx: GpuContext, schema: pa.Schema, local: Optional[DeviceBatch]) -> Optional[DeviceBatch]:
    """CollectLeft build-side replication (hash_join.rs:594-606)."""
    s = N.export_schema(schema)
    h = C.c_void_p()
    try:
        N.check(N.lib().dfgpu_exchange_all_gather(ctx.h, local.h if local is not None else None, C.addressof(s), C.byref(h)))
    finally:
        N.release_schema(s)
    return DeviceBatch(ctx, h) if h.value else None
