"""CPU ORACLE — test infrastructure, NOT product code.

ctypes + pyarrow driver around ``oracle/liboracle.so`` (``df_oracle.cpp``), the CPU restatement of
the reference's hash aggregation / hash join / hash repartition path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs import
this package; nothing in ``datafusion_upstream_b200`` does.

Output schemas follow the reference:
  * aggregate state / result fields — physical-plan/src/aggregates/mod.rs:749-787,
    physical-expr/src/expressions/mod.rs:97-99 (``name[sum]`` ...), sum.rs:110-116, count.rs:225-231,
    average.rs:121-134, min_max.rs:155-161;
  * result types — expr/src/type_coercion/aggregates.rs:397-498;
  * join schema — physical-plan/src/joins/utils.rs:657-729.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import pyarrow as pa

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

(OT_BOOL, OT_I8, OT_I16, OT_I32, OT_I64, OT_U8, OT_U16, OT_U32, OT_U64, OT_F32, OT_F64, OT_DATE32,
 OT_DATE64, OT_TS, OT_TIME32, OT_TIME64, OT_DEC128, OT_UTF8, OT_LARGE_UTF8, OT_BINARY,
 OT_LARGE_BINARY, OT_DICT) = range(1, 23)

AGG_FUNCS = {"sum": 1, "count": 2, "avg": 3, "min": 4, "max": 5,
             "bit_and": 6, "bit_or": 7, "bit_xor": 8, "bool_and": 9, "bool_or": 10}
MODES = {"partial": 0, "final": 1, "final_partitioned": 2, "single": 3, "single_partitioned": 4}
JOIN_TYPES = {"inner": 0, "left": 1, "right": 2, "full": 3, "left_semi": 4, "right_semi": 5,
              "left_anti": 6, "right_anti": 7}


def build(force: bool = False) -> str:
    """Compile liboracle.so with the committed Makefile (gcc only)."""
    src = os.path.join(_HERE, "df_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < max(
            os.path.getmtime(src), os.path.getmtime(os.path.join(_HERE, "df_oracle.h"))):
        subprocess.run(["make", "-s", "-C", _HERE], check=True)
    return _LIB_PATH


class _OCol(C.Structure):
    pass


_OCol._fields_ = [
    ("type", C.c_int32), ("index_type", C.c_int32), ("length", C.c_int64), ("offset", C.c_int64),
    ("validity", C.c_void_p), ("values", C.c_void_p), ("offsets", C.c_void_p),
    ("dict", C.POINTER(_OCol)), ("precision", C.c_int32), ("scale", C.c_int32),
]


class _OAggDesc(C.Structure):
    _fields_ = [("func", C.c_int32), ("input_type", C.c_int32), ("precision", C.c_int32),
                ("scale", C.c_int32)]


class _OColOut(C.Structure):
    _fields_ = [("type", C.c_int32), ("precision", C.c_int32), ("scale", C.c_int32),
                ("length", C.c_int64), ("null_count", C.c_int64), ("validity", C.c_void_p),
                ("values", C.c_void_p), ("offsets", C.c_void_p), ("data_bytes", C.c_int64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.oracle_last_error.restype = C.c_char_p
        L.oracle_agg_new.restype = C.c_void_p
        L.oracle_agg_new.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(_OAggDesc)]
        L.oracle_agg_free.argtypes = [C.c_void_p]
        L.oracle_agg_push.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_agg_num_groups.restype = C.c_int64
        L.oracle_agg_num_groups.argtypes = [C.c_void_p]
        L.oracle_agg_emit.argtypes = [C.c_void_p, C.c_int64]
        L.oracle_agg_out.argtypes = [C.c_void_p, C.c_int, C.POINTER(_OColOut)]
        L.oracle_agg_last_group_ids.restype = C.POINTER(C.c_uint32)
        L.oracle_agg_last_group_ids.argtypes = [C.c_void_p]
        L.oracle_create_hashes.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_int, C.c_void_p]
        L.oracle_join_new.restype = C.c_void_p
        L.oracle_join_new.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int]
        L.oracle_join_free.argtypes = [C.c_void_p]
        L.oracle_join_build_batch.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_join_build_finish.argtypes = [C.c_void_p]
        L.oracle_join_build_rows.restype = C.c_int64
        L.oracle_join_build_rows.argtypes = [C.c_void_p]
        L.oracle_join_probe.restype = C.c_int64
        L.oracle_join_probe.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]
        L.oracle_join_unmatched.restype = C.c_int64
        L.oracle_join_unmatched.argtypes = [C.c_void_p]
        L.oracle_join_set_filter.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_join_set_filter.restype = None
        L.oracle_join_out_build.restype = C.POINTER(C.c_int64)
        L.oracle_join_out_build.argtypes = [C.c_void_p]
        L.oracle_join_out_probe.restype = C.POINTER(C.c_int64)
        L.oracle_join_out_probe.argtypes = [C.c_void_p]
        L.oracle_partition_ids.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.oracle_bench_groupby_count_i64.restype = C.c_double
        L.oracle_bench_groupby_count_i64.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int64,
                                                     C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]
        L.oracle_bench_q1.restype = C.c_double
        L.oracle_bench_q1.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64,
                                      C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]
        L.oracle_bench_join_i64.restype = C.c_double
        L.oracle_bench_join_i64.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int,
                                            C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]
        _lib = L
    return _lib


class OracleError(RuntimeError):
    pass


def _err() -> str:
    return (lib().oracle_last_error() or b"").decode()


# --------------------------------------------------------------------------------------------
# pyarrow <-> ocol
# --------------------------------------------------------------------------------------------
def type_code(t: pa.DataType) -> int:
    if pa.types.is_boolean(t): return OT_BOOL
    if pa.types.is_int8(t): return OT_I8
    if pa.types.is_int16(t): return OT_I16
    if pa.types.is_int32(t): return OT_I32
    if pa.types.is_int64(t): return OT_I64
    if pa.types.is_uint8(t): return OT_U8
    if pa.types.is_uint16(t): return OT_U16
    if pa.types.is_uint32(t): return OT_U32
    if pa.types.is_uint64(t): return OT_U64
    if pa.types.is_float32(t): return OT_F32
    if pa.types.is_float64(t): return OT_F64
    if pa.types.is_date32(t): return OT_DATE32
    if pa.types.is_date64(t): return OT_DATE64
    if pa.types.is_timestamp(t) or pa.types.is_duration(t): return OT_TS
    if pa.types.is_time32(t): return OT_TIME32
    if pa.types.is_time64(t): return OT_TIME64
    if pa.types.is_decimal128(t): return OT_DEC128
    if pa.types.is_string(t): return OT_UTF8
    if pa.types.is_large_string(t): return OT_LARGE_UTF8
    if pa.types.is_binary(t): return OT_BINARY
    if pa.types.is_large_binary(t): return OT_LARGE_BINARY
    if pa.types.is_dictionary(t): return OT_DICT
    raise OracleError(f"NotImplemented: type {t}")


def _addr(buf) -> Optional[int]:
    return None if buf is None else buf.address


def to_ocol(arr, keep: list) -> _OCol:
    if isinstance(arr, pa.ChunkedArray):
        arr = arr.combine_chunks() if arr.num_chunks != 1 else arr.chunk(0)
    t = arr.type
    c = _OCol()
    c.type = type_code(t)
    c.length = len(arr)
    c.offset = arr.offset
    keep.append(arr)
    if c.type == OT_DICT:
        idx = arr.indices
        keep.append(idx)
        c.index_type = type_code(idx.type)
        b = idx.buffers()
        c.offset = idx.offset
        c.validity = _addr(b[0])
        c.values = _addr(b[1])
        d = to_ocol(arr.dictionary, keep)
        keep.append(d)
        c.dict = C.pointer(d)
        return c
    b = arr.buffers()
    c.validity = _addr(b[0]) if arr.null_count else None
    if c.type in (OT_UTF8, OT_LARGE_UTF8, OT_BINARY, OT_LARGE_BINARY):
        c.offsets = _addr(b[1])
        c.values = _addr(b[2]) if b[2] is not None else None
        if c.values is None:
            dummy = pa.py_buffer(b"\0")
            keep.append(dummy)
            c.values = dummy.address
    else:
        c.values = _addr(b[1])
    if pa.types.is_decimal128(t):
        c.precision, c.scale = t.precision, t.scale
    return c


def _col_array(cols: Sequence, keep: list):
    """Array of ocol* (None entries -> NULL)."""
    n = len(cols)
    arr = (C.c_void_p * max(n, 1))()
    for i, a in enumerate(cols):
        if a is None:
            arr[i] = None
        else:
            oc = to_ocol(a, keep)
            keep.append(oc)
            arr[i] = C.addressof(oc)
    keep.append(arr)
    return arr


def _out_to_arrow(o: _OColOut, want: pa.DataType) -> pa.Array:
    n = o.length
    validity = None
    if o.validity:
        validity = pa.py_buffer(C.string_at(o.validity, (n + 7) // 8))
    base = want.value_type if pa.types.is_dictionary(want) else want
    if o.type in (OT_UTF8, OT_LARGE_UTF8, OT_BINARY, OT_LARGE_BINARY):
        offs = pa.py_buffer(C.string_at(o.offsets, 8 * (n + 1)))
        data = pa.py_buffer(C.string_at(o.values, o.data_bytes) if o.data_bytes else b"")
        big = pa.large_string() if o.type in (OT_UTF8, OT_LARGE_UTF8) else pa.large_binary()
        arr = pa.Array.from_buffers(big, n, [validity, offs, data], null_count=o.null_count)
        arr = arr.cast(base)
    elif o.type == OT_BOOL:
        vals = pa.py_buffer(C.string_at(o.values, (n + 7) // 8) if n else b"")
        arr = pa.Array.from_buffers(pa.bool_(), n, [validity, vals], null_count=o.null_count)
    else:
        vals = pa.py_buffer(C.string_at(o.values, o.data_bytes) if o.data_bytes else b"")
        arr = pa.Array.from_buffers(base, n, [validity, vals], null_count=o.null_count)
    if pa.types.is_dictionary(want):
        # GroupValuesRows::emit casts keys back to the Dictionary type (row.rs:207-218)
        arr = arr.dictionary_encode().cast(want)
    return arr


# --------------------------------------------------------------------------------------------
# hashing / partitioning
# --------------------------------------------------------------------------------------------
def create_hashes(cols: Sequence[pa.Array], seed: int = 0, force_collisions: bool = False) -> np.ndarray:
    keep: list = []
    arr = _col_array(cols, keep)
    out = np.zeros(len(cols[0]), dtype=np.uint64)
    if lib().oracle_create_hashes(arr, len(cols), seed, int(force_collisions), out.ctypes.data) != 0:
        raise OracleError(_err())
    return out


def partition_ids(key_cols: Sequence[pa.Array], n_parts: int) -> np.ndarray:
    keep: list = []
    arr = _col_array(key_cols, keep)
    out = np.zeros(len(key_cols[0]), dtype=np.uint32)
    lib().oracle_partition_ids(arr, len(key_cols), n_parts, out.ctypes.data)
    return out


def filter_batch(batch: pa.RecordBatch, predicate: int) -> pa.RecordBatch:
    """FilterExec::batch_filter (physical-plan/src/filter.rs:315-327): arrow `filter_record_batch` keeps the rows whose
    predicate is true; a NULL predicate counts as false (arrow-rs `prep_null_mask_filter`).  Row order is preserved.
    Restated as an index list + take, row by row."""
    m = batch.column(predicate)
    if not pa.types.is_boolean(m.type):
        raise OracleError("Internal: filter predicate is not Boolean")
    keep = [i for i, v in enumerate(m.to_pylist()) if v is True]
    idx = pa.array(keep, pa.int64())
    return pa.record_batch([c.take(idx) for c in batch.columns], schema=batch.schema)


# Decimal128 arithmetic of ProjectionExec: BinaryExpr::evaluate (physical-expr/src/expressions/binary.rs:260-275) calls
# arrow-rs `add_wrapping` / `sub_wrapping` / `mul_wrapping` (arrow-arith 50, numeric.rs `decimal_op`; un-vendored).  For
# decimals those kernels derive the result type and compute CHECKED on the valid slots:
#   + / - : scale = max(s1, s2); precision = min(38, scale + max(p1 - s1, p2 - s2) + 1);
#           l * 10^(scale - s1) +/- r * 10^(scale - s2), every step checked in i128
#   *     : precision = min(38, p1 + p2 + 1); scale = s1 + s2 (error if > 38); l * r checked in i128
# Pinned by decimal.slt:229-360 (c1+1, c1+c5, c1-1, c1-c5, c1*20, c1*c5: types and values).
I128_MAX = (1 << 127) - 1
I128_MIN = -(1 << 127)


def decimal_result_type(op: str, lt: pa.DataType, rt: pa.DataType) -> pa.DataType:
    p1, s1, p2, s2 = lt.precision, lt.scale, rt.precision, rt.scale
    if op in ("add", "sub"):
        s = max(s1, s2)
        return pa.decimal128(min(38, s + max(p1 - s1, p2 - s2) + 1), s)
    if op == "mul":
        if s1 + s2 > 38:
            raise OracleError(f"Invalid argument error: Output scale of {lt} * {rt} would exceed max scale of 38")
        return pa.decimal128(min(38, p1 + p2 + 1), s1 + s2)
    raise OracleError(op)


def _chk(v: int, what: str) -> int:
    if v > I128_MAX or v < I128_MIN:
        raise OracleError(f"Arrow error: Compute error: Overflow happened on: {what}")
    return v


def eval_decimal(batch: pa.RecordBatch, nodes: Sequence[tuple]) -> pa.Array:
    """nodes (post-order, the last one is the root): ("col", index) | ("lit", unscaled int, precision, scale) |
    ("add" | "sub" | "mul", left node, right node).  Returns the Decimal128 array of the root."""
    n = batch.num_rows
    vals, types = [], []
    for nd in nodes:
        if nd[0] == "col":
            c = batch.column(nd[1])
            if not pa.types.is_decimal128(c.type):
                raise OracleError(f"NotImplemented: decimal arithmetic over {c.type}")
            s = c.type.scale
            vals.append([None if v is None else int(v.scaleb(s)) for v in c.to_pylist()])
            types.append(c.type)
        elif nd[0] == "lit":
            vals.append([nd[1]] * n)
            types.append(pa.decimal128(nd[2], nd[3]))
        else:
            op, li, ri = nd
            lt, rt = types[li], types[ri]
            t = decimal_result_type(op, lt, rt)
            out = []
            for a, b in zip(vals[li], vals[ri]):
                if a is None or b is None:
                    out.append(None)
                elif op == "mul":
                    out.append(_chk(a * b, f"{a} * {b}"))
                else:
                    la = _chk(a * 10 ** (t.scale - lt.scale), f"{a} * 10^{t.scale - lt.scale}")
                    rb = _chk(b * 10 ** (t.scale - rt.scale), f"{b} * 10^{t.scale - rt.scale}")
                    out.append(_chk(la + rb if op == "add" else la - rb, f"{la} {'+' if op == 'add' else '-'} {rb}"))
            vals.append(out)
            types.append(t)
    import decimal as _d
    ctx = _d.Context(prec=60)
    t = types[-1]
    return pa.array([None if v is None else ctx.create_decimal(v).scaleb(-t.scale, ctx) for v in vals[-1]], t)


def partition(batch: pa.RecordBatch, key_idx: Sequence[int], n_parts: int) -> List[Optional[pa.RecordBatch]]:
    """BatchPartitioner::partition (repartition/mod.rs:133-217): empty partitions are not produced."""
    ids = partition_ids([batch.column(i) for i in key_idx], n_parts)
    out: List[Optional[pa.RecordBatch]] = []
    for p in range(n_parts):
        idx = np.nonzero(ids == p)[0]
        out.append(batch.take(pa.array(idx, pa.int64())) if len(idx) else None)
    return out


# --------------------------------------------------------------------------------------------
# aggregation
# --------------------------------------------------------------------------------------------
@dataclass
class AggSpec:
    func: str                 # sum | count | avg | min | max
    name: str                 # display name, e.g. "AVG(b)"
    input_type: Optional[pa.DataType]  # raw argument type (None only for COUNT(*))


def _value_type(t: pa.DataType) -> pa.DataType:
    return t.value_type if pa.types.is_dictionary(t) else t


def sum_type(t: pa.DataType) -> pa.DataType:
    t = _value_type(t)
    if pa.types.is_decimal128(t): return pa.decimal128(min(38, t.precision + 10), t.scale)
    if pa.types.is_unsigned_integer(t): return pa.uint64()
    if pa.types.is_integer(t): return pa.int64()
    if pa.types.is_floating(t): return pa.float64()
    raise OracleError(f"NotImplemented: SUM({t})")


def avg_type(t: pa.DataType) -> pa.DataType:
    t = _value_type(t)
    if pa.types.is_decimal128(t): return pa.decimal128(min(38, t.precision + 4), min(38, t.scale + 4))
    return pa.float64()


def avg_sum_type(t: pa.DataType) -> pa.DataType:
    t = _value_type(t)
    return t if pa.types.is_decimal128(t) else pa.float64()


def state_fields(a: AggSpec):
    if a.func == "sum": return [(f"{a.name}[sum]", sum_type(a.input_type))]
    if a.func == "count": return [(f"{a.name}[count]", pa.int64())]
    if a.func == "avg": return [(f"{a.name}[count]", pa.uint64()), (f"{a.name}[sum]", avg_sum_type(a.input_type))]
    if a.func == "min": return [(f"{a.name}[min]", _value_type(a.input_type))]
    if a.func == "max": return [(f"{a.name}[max]", _value_type(a.input_type))]
    if a.func in ("bit_and", "bit_or", "bit_xor", "bool_and", "bool_or"):   # bit_and_or_xor.rs:96-102, bool_and_or.rs:117-123
        return [(f"{a.name}[{a.func}]", _value_type(a.input_type))]
    raise OracleError(a.func)


def result_field(a: AggSpec):
    if a.func == "sum": return (a.name, sum_type(a.input_type))
    if a.func == "count": return (a.name, pa.int64())
    if a.func == "avg": return (a.name, avg_type(a.input_type))
    return (a.name, _value_type(a.input_type))


class OracleAgg:
    """AggregateExec restatement.  ``push`` = group_aggregate_batch, ``emit`` = emit(EmitTo)."""

    def __init__(self, mode: str, group_names: Sequence[str], group_types: Sequence[pa.DataType],
                 aggs: Sequence[AggSpec]):
        self.mode = mode
        self.group_names = list(group_names)
        self.group_types = list(group_types)
        self.aggs = list(aggs)
        d = (_OAggDesc * max(1, len(aggs)))()
        for i, a in enumerate(aggs):
            d[i].func = AGG_FUNCS[a.func]
            if a.input_type is None:
                d[i].input_type = OT_I64
            else:
                vt = _value_type(a.input_type)
                d[i].input_type = type_code(vt)
                if pa.types.is_decimal128(vt):
                    d[i].precision, d[i].scale = vt.precision, vt.scale
        self.h = lib().oracle_agg_new(MODES[mode], len(group_types), len(aggs), d)
        if not self.h:
            raise OracleError(_err())

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_agg_free(self.h)
            self.h = None

    @property
    def first_stage(self) -> bool:
        return self.mode in ("partial", "single", "single_partitioned")

    def push(self, group_cols, args, filters=None):
        """args: first-stage: one array (or None = COUNT(*)) per aggregate;
        final modes: the flattened state columns."""
        keep: list = []
        g = _col_array(group_cols, keep)
        a = _col_array(args, keep)
        f = _col_array(filters, keep) if filters is not None else None
        if lib().oracle_agg_push(self.h, g, a, f) != 0:
            raise OracleError(_err())

    def last_group_ids(self, n: int) -> np.ndarray:
        p = lib().oracle_agg_last_group_ids(self.h)
        return np.ctypeslib.as_array(p, shape=(n,)).copy()

    def num_groups(self) -> int:
        return lib().oracle_agg_num_groups(self.h)

    def schema(self) -> pa.Schema:
        fields = [(n, t) for n, t in zip(self.group_names, self.group_types)]
        for a in self.aggs:
            fields += state_fields(a) if self.mode == "partial" else [result_field(a)]
        return fields

    def emit(self, first_n: Optional[int] = None) -> pa.RecordBatch:
        nc = lib().oracle_agg_emit(self.h, -1 if first_n is None else first_n)
        if nc < 0:
            raise OracleError(_err())
        fields = self.schema()
        assert nc == len(fields), (nc, fields)
        arrays = []
        for i, (_, t) in enumerate(fields):
            o = _OColOut()
            lib().oracle_agg_out(self.h, i, C.byref(o))
            arrays.append(_out_to_arrow(o, t))
        return pa.RecordBatch.from_arrays(arrays, names=[n for n, _ in fields])


# --------------------------------------------------------------------------------------------
# hash join
# --------------------------------------------------------------------------------------------
def _take_nullable(batch: pa.RecordBatch, idx: np.ndarray) -> List[pa.Array]:
    """build_batch_from_indices (joins/utils.rs:1180-1230): -1 -> NULL row."""
    out = []
    mask = idx < 0
    for col in batch.columns:
        if len(col) == 0 or mask.all():
            out.append(pa.nulls(len(idx), col.type))
        else:
            ind = pa.array(np.where(mask, 0, idx), pa.int64(), mask=mask)
            out.append(col.take(ind))
    return out


class OracleJoin:
    """HashJoinExec restatement (left = build, right = probe: hash_join.rs:404-407)."""

    def __init__(self, join_type: str, on: Sequence[tuple], null_equals_null: bool = False,
                 batch_size: int = 8192, force_collisions: bool = False):
        self.join_type = join_type
        self.on = list(on)
        self.h = lib().oracle_join_new(JOIN_TYPES[join_type], int(null_equals_null), len(on), batch_size,
                                       int(force_collisions))
        self._keep: list = []
        self.build_batches: List[pa.RecordBatch] = []
        self.build_cat: Optional[pa.RecordBatch] = None
        self.out_batches = 0
        self._filter_cb = None
        self._probe_batch = None

    def set_filter(self, column_indices: Sequence[tuple], predicate):
        """JoinFilter (joins/utils.rs:1143-1176): column_indices = [(side, index)] with side "left" | "right" selects the columns of
        the intermediate batch; predicate(intermediate RecordBatch) -> Boolean array; NULL counts as false (arrow `filter`)."""
        FN = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_int64, C.POINTER(C.c_uint8))

        def cb(_user, bi, pi, n, keep):
            b = np.ctypeslib.as_array(bi, shape=(n,)).copy()
            p = np.ctypeslib.as_array(pi, shape=(n,)).copy()
            cols = []
            for side, idx in column_indices:
                src, ix = (self.build_cat, b) if side == "left" else (self._probe_batch, p)
                cols.append(src.column(idx).take(pa.array(ix)))
            inter = pa.RecordBatch.from_arrays(cols, names=[f"c{i}" for i in range(len(cols))])
            mask = predicate(inter)
            m = np.asarray(pa.compute.fill_null(mask, False)) if len(mask) else np.zeros(0, bool)
            for i in range(n):
                keep[i] = 1 if m[i] else 0
        self._filter_cb = FN(cb)
        lib().oracle_join_set_filter(self.h, self._filter_cb, None)

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_join_free(self.h)
            self.h = None

    def build(self, batches: Sequence[pa.RecordBatch], schema: Optional[pa.Schema] = None):
        for b in batches:
            self.build_batches.append(b)
            arr = _col_array([b.column(l) for l, _ in self.on], self._keep)
            lib().oracle_join_build_batch(self.h, arr)
        lib().oracle_join_build_finish(self.h)
        sch = schema or self.build_batches[0].schema
        # concat_batches(batches.rev()) — hash_join.rs:764
        rev = list(reversed(self.build_batches))
        self.build_cat = pa.Table.from_batches(rev, schema=sch).combine_chunks().to_batches()[0] \
            if sum(b.num_rows for b in rev) else pa.RecordBatch.from_pylist([], schema=sch)
        self.build_schema = sch

    def _materialize(self, n: int, probe: Optional[pa.RecordBatch], probe_schema: pa.Schema) -> pa.RecordBatch:
        bi = np.ctypeslib.as_array(lib().oracle_join_out_build(self.h), shape=(n,)).copy() if n else np.zeros(0, np.int64)
        pi = np.ctypeslib.as_array(lib().oracle_join_out_probe(self.h), shape=(n,)).copy() if n else np.zeros(0, np.int64)
        jt = self.join_type
        cols, names = [], []
        if jt not in ("right_semi", "right_anti"):
            cols += _take_nullable(self.build_cat, bi)
            names += self.build_schema.names
        if jt not in ("left_semi", "left_anti"):
            if probe is None:
                probe = pa.RecordBatch.from_pylist([], schema=probe_schema)
            cols += _take_nullable(probe, pi)
            names += probe_schema.names
        return pa.RecordBatch.from_arrays(cols, names=names)

    def probe(self, batch: pa.RecordBatch) -> pa.RecordBatch:
        keep: list = []
        arr = _col_array([batch.column(r) for _, r in self.on], keep)
        nb = C.c_int64(0)
        self._probe_batch = batch
        n = lib().oracle_join_probe(self.h, arr, C.byref(nb))
        self.out_batches += nb.value
        self.probe_schema = batch.schema
        return self._materialize(n, batch, batch.schema)

    def unmatched(self, probe_schema: pa.Schema) -> Optional[pa.RecordBatch]:
        n = lib().oracle_join_unmatched(self.h)
        if n < 0:
            return None
        return self._materialize(n, None, probe_schema)


def hash_join(join_type: str, left: Sequence[pa.RecordBatch], right: Sequence[pa.RecordBatch], on,
              null_equals_null=False, batch_size=8192, force_collisions=False,
              left_schema=None, right_schema=None, filter=None) -> pa.Table:
    """One HashJoinExec partition: collect build, stream probe batches, then unmatched rows.
    filter = (column_indices, predicate): the residual JoinFilter, see OracleJoin.set_filter."""
    j = OracleJoin(join_type, on, null_equals_null, batch_size, force_collisions)
    if filter is not None:
        j.set_filter(*filter)
    left_schema = left_schema or left[0].schema
    right_schema = right_schema or right[0].schema
    j.build(left, left_schema)
    outs = [j.probe(b) for b in right]
    u = j.unmatched(right_schema)
    if u is not None:
        outs.append(u)
    if not outs:
        outs = [j._materialize(0, None, right_schema)]
    hash_join.last_out_batches = j.out_batches
    return pa.Table.from_batches(outs)


# --------------------------------------------------------------------------------------------
# CPU baseline drivers (bench.py)
# --------------------------------------------------------------------------------------------
def bench_groupby_count_i64(keys: np.ndarray, threads: int, batch: int = 8192):
    g, c = C.c_int64(0), C.c_uint64(0)
    keys = np.ascontiguousarray(keys, dtype=np.int64)
    s = lib().oracle_bench_groupby_count_i64(keys.ctypes.data, len(keys), threads, batch, C.byref(g), C.byref(c))
    return s, g.value, c.value


def bench_q1(flag: np.ndarray, status: np.ndarray, dec_cols: Sequence[np.ndarray], threads: int, batch: int = 8192):
    ptrs = (C.c_void_p * 5)(*[d.ctypes.data for d in dec_cols])
    g, c = C.c_int64(0), C.c_uint64(0)
    s = lib().oracle_bench_q1(flag.ctypes.data, status.ctypes.data, ptrs, len(flag), threads, batch, C.byref(g), C.byref(c))
    return s, g.value, c.value


def bench_join_i64(build: np.ndarray, probe: np.ndarray, threads: int, batch: int = 8192):
    r, c = C.c_int64(0), C.c_uint64(0)
    s = lib().oracle_bench_join_i64(build.ctypes.data, len(build), probe.ctypes.data, len(probe), threads, batch,
                                    C.byref(r), C.byref(c))
    return s, r.value, c.value
