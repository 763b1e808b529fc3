"""Drivers shared by the -m gpu parity tests: run the same operator on the GPU (through the C ABI)
and on the CPU oracle, on the same Arrow inputs."""
import numpy as np
import pyarrow as pa

import oracle
from oracle import AggSpec, OracleAgg

FUNCS = {"sum": 1, "count": 2, "avg": 3, "min": 4, "max": 5, "bit_and": 6, "bit_or": 7, "bit_xor": 8, "bool_and": 9, "bool_or": 10}


def gpu_ctx(**kw):
    import datafusion_upstream_b200 as df
    return df.GpuContext(device=0, **kw)


def gpu_agg(ctx, mode, batches, group_by, specs, schema=None):
    """specs: list of (func, arg column name | None, display name, filter column name | None)."""
    import datafusion_upstream_b200 as df
    schema = schema or batches[0].schema
    modes = {"partial": 0, "final": 1, "final_partitioned": 2, "single": 3, "single_partitioned": 4}
    if mode.startswith("final"):   # state columns are resolved by position (aggregates/mod.rs:1038-1067)
        specs = [(f, None, n, None) for f, _, n, _ in specs]
    exprs = [df.AggregateExpr(df.AggFunc(FUNCS[f]), a, n, flt) for f, a, n, flt in specs]
    op = df.GpuAggregateExec(ctx, df.AggregateMode(modes[mode]), schema, group_by, exprs)
    for b in batches:
        op.push(b)
    out = op.emit()
    op.close()
    return out


def oracle_agg(mode, batches, group_by, specs, schema=None):
    schema = schema or batches[0].schema
    gidx = [schema.get_field_index(g) if isinstance(g, str) else g for g in group_by]
    first = mode in ("partial", "single", "single_partitioned")
    aggs = []
    if first:
        for f, a, n, _ in specs:
            aggs.append(AggSpec(f, n, None if a is None else schema.field(a).type))
    else:
        # Final modes: input types are the state types; recover the raw type for schema naming
        pos = len(gidx)
        for f, a, n, _ in specs:
            if f == "count":
                aggs.append(AggSpec(f, n, None)); pos += 1
            elif f == "avg":
                st = schema.field(pos + 1).type
                aggs.append(AggSpec(f, n, st)); pos += 2
            else:
                st = schema.field(pos).type
                aggs.append(_FinalSpec(f, n, st)); pos += 1
    o = OracleAgg(mode, [schema.field(i).name for i in gidx], [schema.field(i).type for i in gidx], aggs)
    if not first:
        # SUM result type in Final mode is the state type itself (not widened again)
        o.schema = lambda o=o: _final_schema(o)
    for b in batches:
        if b.num_rows == 0:
            continue
        gcols = [b.column(i) for i in gidx]
        if first:
            args = [None if a is None else b.column(a) for _, a, _, _ in specs]
            filts = [None if flt is None else b.column(flt) for _, _, _, flt in specs]
            o.push(gcols, args, filts if any(f is not None for f in filts) else None)
        else:
            o.push(gcols, [b.column(i) for i in range(len(gidx), b.num_columns)])
    return o.emit()


class _FinalSpec(AggSpec):
    """AggSpec whose input_type is already the state type (sum type must not be widened again)."""
    final = True


def _final_schema(o):
    fields = [(n, t) for n, t in zip(o.group_names, o.group_types)]
    for a in o.aggs:
        if a.func == "sum" and getattr(a, "final", False):
            fields.append((a.name, a.input_type))
        else:
            fields.append(oracle.result_field(a))
    return fields


def random_keys(rng, n, kind, card, null_frac=0.0):
    """Key arrays of many Arrow types with `card` distinct values."""
    base = rng.integers(0, card, n)
    if kind == "int64":
        with np.errstate(over="ignore"):
            vals = (base.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)).view(np.int64) - (base % 3 == 0) * 7
        arr = pa.array(vals, pa.int64())
    elif kind == "int64_small":
        arr = pa.array(base.astype(np.int64) - 1, pa.int64())   # includes -1 == the U64 EMPTY bit pattern
    elif kind == "uint64":
        arr = pa.array(base.astype(np.uint64) + np.uint64(2**63), pa.uint64())
    elif kind == "int32":
        arr = pa.array((base - card // 2).astype(np.int32), pa.int32())
    elif kind == "int16":
        arr = pa.array((base % 30000).astype(np.int16), pa.int16())
    elif kind == "int8":
        arr = pa.array((base % 100).astype(np.int8), pa.int8())
    elif kind == "uint8":
        arr = pa.array((base % 200).astype(np.uint8), pa.uint8())
    elif kind == "bool":
        arr = pa.array(base % 2 == 0, pa.bool_())
    elif kind == "float64":
        arr = pa.array(np.where(base % 11 == 0, -0.0, base * 0.5), pa.float64())
    elif kind == "float32":
        arr = pa.array((base * 0.25).astype(np.float32), pa.float32())
    elif kind == "date32":
        arr = pa.array(base.astype(np.int32) + 8000, pa.int32()).cast(pa.date32())
    elif kind == "timestamp":
        arr = pa.array(base.astype(np.int64) * 1000, pa.int64()).cast(pa.timestamp("us"))
    elif kind == "decimal":
        import decimal
        arr = pa.array([decimal.Decimal(int(v) - 5).scaleb(-2) for v in base], pa.decimal128(20, 2))
    elif kind in ("utf8", "large_utf8", "binary", "dict_utf8", "dict_utf8_i8"):
        words = ["", "a", "b", "ab", "abcdefgh", "abcdefghi", "the quick brown fox", "x" * 40]
        strs = [words[v % len(words)] + (str(v // len(words)) if v >= len(words) else "") for v in range(card)]
        vals = [strs[v] for v in base]
        if kind == "utf8":
            arr = pa.array(vals, pa.string())
        elif kind == "large_utf8":
            arr = pa.array(vals, pa.large_string())
        elif kind == "binary":
            arr = pa.array([v.encode() for v in vals], pa.binary())
        elif kind == "dict_utf8":
            arr = pa.array(vals, pa.string()).dictionary_encode()
        else:
            arr = pa.DictionaryArray.from_arrays(pa.array((base % min(card, 100)).astype(np.int8), pa.int8()),
                                                 pa.array(strs[:min(card, 100)], pa.string()))
    else:
        raise ValueError(kind)
    if null_frac > 0:
        mask = rng.random(n) < null_frac
        if pa.types.is_dictionary(arr.type):
            idx = arr.indices.to_numpy(zero_copy_only=False)
            arr = pa.DictionaryArray.from_arrays(pa.array(idx, arr.type.index_type, mask=mask), arr.dictionary)
        else:
            arr = pa.array(arr.to_pylist(), arr.type, mask=mask) if not pa.types.is_decimal(arr.type) else \
                pa.array([None if m else v for v, m in zip(arr.to_pylist(), mask)], arr.type)
    return arr


def random_values(rng, n, kind, null_frac=0.0):
    import decimal
    if kind == "int64":
        arr = pa.array(rng.integers(-10**12, 10**12, n), pa.int64())
    elif kind == "int32":
        arr = pa.array(rng.integers(-10**6, 10**6, n).astype(np.int32), pa.int32())
    elif kind == "int8":
        arr = pa.array(rng.integers(-128, 128, n).astype(np.int8), pa.int8())
    elif kind == "uint32":
        arr = pa.array(rng.integers(0, 2**32, n).astype(np.uint32), pa.uint32())
    elif kind == "uint64":
        arr = pa.array(rng.integers(0, 2**63, n).astype(np.uint64), pa.uint64())
    elif kind == "float64":
        arr = pa.array(rng.normal(0, 1000, n), pa.float64())
    elif kind == "float32":
        arr = pa.array(rng.normal(0, 10, n).astype(np.float32), pa.float32())
    elif kind == "decimal":
        arr = pa.array([decimal.Decimal(int(v)).scaleb(-2) for v in rng.integers(-10**9, 10**9, n)], pa.decimal128(15, 2))
    elif kind == "decimal_big":
        arr = pa.array([decimal.Decimal(int(v) * 10**18 + 12345).scaleb(-4) for v in rng.integers(-10**12, 10**12, n)], pa.decimal128(38, 4))
    elif kind == "date32":
        arr = pa.array(rng.integers(0, 20000, n).astype(np.int32), pa.int32()).cast(pa.date32())
    else:
        raise ValueError(kind)
    if null_frac > 0:
        mask = rng.random(n) < null_frac
        arr = pa.array([None if m else v for v, m in zip(arr.to_pylist(), mask)], arr.type)
    return arr
