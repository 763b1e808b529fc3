"""-m gpu parity tests for grouped hash aggregation: CUDA path (through the C ABI) vs the CPU oracle
and vs the reference's golden vectors.  Bit-exact for integers / Decimal128 / keys; Float64 sum/avg
within 1e-9 relative (north_star)."""
import decimal
import json
import os

import numpy as np
import pyarrow as pa
import pytest

from gpu_helpers import FUNCS, gpu_agg, gpu_ctx, oracle_agg, random_keys, random_values
from util import assert_rows_equal, assert_tables_equal, rows_of

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
AGGS = json.load(open(os.path.join(G, "aggregates.json")))


@pytest.fixture(scope="module")
def ctx():
    c = gpu_ctx()
    yield c
    c.close()


def some_data():
    return [pa.record_batch([pa.array(b["a"], pa.uint32()), pa.array(b["b"], pa.float64())], names=["a", "b"])
            for b in AGGS["some_data"]["batches"]]


def test_golden_check_aggregates(ctx):
    g = AGGS["check_aggregates"]
    specs = [("avg", "b", "AVG(b)", None)]
    st = gpu_agg(ctx, "partial", some_data(), ["a"], specs)
    assert st.schema.names == g["partial_no_spill"]["header"]
    assert_rows_equal(rows_of(st), [(int(a), int(c), float(s)) for a, c, s in g["partial_no_spill"]["rows"]])
    out = gpu_agg(ctx, "final", [st], ["a"], specs)
    assert out.schema.names == g["final"]["header"]
    assert_rows_equal(rows_of(out), [(int(a), float(v)) for a, v in g["final"]["rows"]], rel=0.0)


def test_golden_grouping_sets_null_keys(ctx):
    g = AGGS["check_grouping_sets"]
    batches = []
    for b in some_data():
        for null_a, null_b in g["groups"]:
            a = pa.nulls(b.num_rows, pa.uint32()) if null_a else b.column(0)
            bb = pa.nulls(b.num_rows, pa.float64()) if null_b else b.column(1)
            batches.append(pa.record_batch([a, bb], names=["a", "b"]))
    specs = [("count", None, "COUNT(1)", None)]
    conv = lambda r: (None if r[0] is None else int(r[0]), None if r[1] is None else float(r[1]), int(r[2]))
    st = gpu_agg(ctx, "partial", batches, ["a", "b"], specs)
    assert st.schema.names == g["partial_no_spill"]["header"]
    assert_rows_equal(rows_of(st), [conv(r) for r in g["partial_no_spill"]["rows"]])
    out = gpu_agg(ctx, "final", [st], ["a", "b"], specs)
    assert_rows_equal(rows_of(out), [conv(r) for r in g["final"]["rows"]])


@pytest.mark.parametrize("name", ["avg_decimal", "avg_decimal_with_nulls"])
def test_golden_avg_decimal(ctx, name):
    g = AGGS[name]
    t = pa.decimal128(*g["input_type"])
    arr = pa.array([None if v is None else decimal.Decimal(v) for v in g["input"]], t)
    b = pa.record_batch([pa.array([0] * len(arr), pa.int8()), arr], names=["k", "v"])
    out = gpu_agg(ctx, "single", [b], ["k"], [("avg", "v", "avg", None)])
    assert out.schema.field(1).type == pa.decimal128(*g["expected_type"])
    assert out.column(1)[0].as_py() == decimal.Decimal(g["expected_unscaled"]).scaleb(-g["expected_type"][1])


def test_golden_decimal_slt(ctx):
    d = AGGS["decimal_slt"]
    t = pa.decimal128(*d["c1_type"])
    c1 = pa.array([decimal.Decimal(v) for v in d["c1"]], t)
    c4 = pa.array(d["c4"], pa.bool_())
    b = pa.record_batch([pa.array([0] * len(c1), pa.int8()), c1, c4], names=["k", "c1", "c4"])
    out = gpu_agg(ctx, "single", [b], ["k"], [("sum", "c1", "s", None), ("avg", "c1", "a", None)])
    assert out.schema.field(1).type == pa.decimal128(20, 6) and out.schema.field(2).type == pa.decimal128(14, 10)
    assert out.column(1)[0].as_py() == decimal.Decimal(d["sum_c1"][0][-1])
    assert out.column(2)[0].as_py() == decimal.Decimal(d["avg_c1"][0][-1])     # truncation, not rounding
    out = gpu_agg(ctx, "single", [b], ["c1"], [("count", None, "c", None)])
    assert_rows_equal(rows_of(out), [(decimal.Decimal(v), int(c)) for c, v in d["count_by_c1"]])
    out = gpu_agg(ctx, "single", [b], ["c1", "c4"], [("count", None, "c", None)])
    assert_rows_equal(rows_of(out), [(decimal.Decimal(v), bb == "true", int(c)) for c, v, bb in d["count_by_c1_c4"]])


KEY_SETS = [
    ["int64"], ["int64_small"], ["uint64"], ["int32"], ["int16", "int8"], ["bool", "float64"], ["float32"], ["date32"],
    ["timestamp"], ["decimal"], ["utf8"], ["large_utf8"], ["binary"], ["dict_utf8"], ["dict_utf8_i8"],
    ["utf8", "int32"], ["utf8", "utf8"], ["int64", "date32", "int32"], ["dict_utf8", "int16"], ["uint8", "uint8"],
    ["utf8", "decimal", "bool"],
]


def make_table(rng, n, key_kinds, card, key_nulls, val_nulls):
    cols, names = [], []
    for i, k in enumerate(key_kinds):
        cols.append(random_keys(rng, n, k, card, key_nulls))
        names.append(f"k{i}")
    for nm, kind in [("i", "int64"), ("f", "float64"), ("d", "decimal"), ("s", "int32"), ("u", "uint32"), ("g", "float32"),
                     ("big", "decimal_big"), ("dt", "date32"), ("b8", "int8")]:
        cols.append(random_values(rng, n, kind, val_nulls))
        names.append(nm)
    cols.append(pa.array(rng.random(n) < 0.7, pa.bool_(), mask=(rng.random(n) < 0.1) if val_nulls else None))
    names.append("flt")
    return pa.record_batch(cols, names=names)


SPECS = [
    ("count", None, "COUNT(*)", None), ("count", "i", "COUNT(i)", None), ("sum", "i", "SUM(i)", None),
    ("sum", "f", "SUM(f)", None), ("sum", "d", "SUM(d)", None), ("sum", "u", "SUM(u)", None), ("sum", "g", "SUM(g)", None),
    ("sum", "big", "SUM(big)", None), ("avg", "f", "AVG(f)", None), ("avg", "d", "AVG(d)", None), ("avg", "s", "AVG(s)", None),
    ("min", "i", "MIN(i)", None), ("max", "i", "MAX(i)", None), ("min", "f", "MIN(f)", None), ("max", "g", "MAX(g)", None),
    ("min", "d", "MIN(d)", None), ("max", "big", "MAX(big)", None), ("min", "dt", "MIN(dt)", None), ("max", "b8", "MAX(b8)", None),
    ("max", "u", "MAX(u)", None), ("sum", "i", "SUM(i) FILTER", "flt"), ("count", "f", "COUNT(f) FILTER", "flt"),
    ("avg", "d", "AVG(d) FILTER", "flt"), ("min", "s", "MIN(s) FILTER", "flt"),
]


@pytest.mark.parametrize("keys", KEY_SETS, ids=["+".join(k) for k in KEY_SETS])
@pytest.mark.parametrize("nulls", [False, True], ids=["nonull", "nulls"])
def test_single_mode_vs_oracle(ctx, keys, nulls):
    rng = np.random.default_rng(sum(map(ord, "".join(keys))) * 2 + nulls)
    batches = [make_table(rng, n, keys, 37, 0.1 if nulls else 0.0, 0.15 if nulls else 0.0) for n in (1000, 1, 0, 4097)]
    gk = [f"k{i}" for i in range(len(keys))]
    got = gpu_agg(ctx, "single", batches, gk, SPECS)
    want = oracle_agg("single", batches, gk, SPECS)
    assert_tables_equal(got, want, key_cols=list(range(len(keys))), what=str(keys))


@pytest.mark.parametrize("keys", [["int64"], ["utf8", "int32"], ["dict_utf8"], ["int16", "int8"]], ids=lambda k: "+".join(k))
def test_partial_then_final_vs_oracle(ctx, keys):
    """Partial per input partition -> (exchange) -> Final merge: Partial emits a key once per partition,
    Final must re-merge duplicates."""
    rng = np.random.default_rng(11)
    gk = [f"k{i}" for i in range(len(keys))]
    parts = [[make_table(rng, n, keys, 53, 0.05, 0.1) for n in (700, 300)] for _ in range(3)]
    g_states = [gpu_agg(ctx, "partial", p, gk, SPECS) for p in parts]
    o_states = [oracle_agg("partial", p, gk, SPECS) for p in parts]
    for gs, os_ in zip(g_states, o_states):
        assert_tables_equal(gs, os_, key_cols=list(range(len(keys))), what="partial state " + str(keys))
    final_specs = [(f, None, n, None) for f, _, n, _ in SPECS]
    got = gpu_agg(ctx, "final_partitioned", g_states, list(range(len(keys))), final_specs)
    want = oracle_agg("final_partitioned", o_states, list(range(len(keys))), final_specs)
    assert_tables_equal(got, want, key_cols=list(range(len(keys))), what="final " + str(keys))
    # and equals Single over all the data
    flat = [b for p in parts for b in p]
    single = oracle_agg("single", flat, gk, SPECS)
    assert_rows_equal(rows_of(got), rows_of(single), key_cols=list(range(len(keys))))


def test_high_cardinality_growth_and_chunks():
    """ClickBench shape (config C2) at a size the oracle finishes in seconds; small chunks force several
    table growths, rehashes and pending-id resolution."""
    ctx = gpu_ctx(chunk_rows=50_000)
    rng = np.random.default_rng(5)
    n = 1_000_000
    keys = (rng.zipf(1.2, n) % 300_000).astype(np.int64) * 7919 - 1
    b = pa.record_batch([pa.array(keys), pa.array(rng.integers(0, 100, n), pa.int64())], names=["UserID", "v"])
    specs = [("count", None, "COUNT(*)", None), ("sum", "v", "SUM(v)", None), ("max", "v", "MAX(v)", None)]
    got = gpu_agg(ctx, "single", [b.slice(0, 600_000), b.slice(600_000)], ["UserID"], specs)
    want = oracle_agg("single", [b], ["UserID"], specs)
    assert_tables_equal(got, want, key_cols=[0])
    # checksum property at the same size: counts add up to the row count
    assert sum(got.column(1).to_pylist()) == n
    ctx.close()


def test_q1_shape_private_path(ctx):
    """TPC-H Q1 shape (config C1): 2 Utf8 keys, 4 groups, Decimal128 SUM x4 / AVG x3 / COUNT(*)."""
    rng = np.random.default_rng(1)
    n = 200_000
    pick = rng.choice(4, n, p=[0.25, 0.006, 0.493, 0.251])
    flag = pa.array(np.array(["A", "N", "N", "R"])[pick])
    status = pa.array(np.array(["F", "F", "O", "F"])[pick])
    dec = lambda lo, hi, p, s: pa.array([decimal.Decimal(int(v)).scaleb(-s) for v in rng.integers(lo, hi, n)], pa.decimal128(p, s))
    b = pa.record_batch([flag, status, dec(100, 5001, 15, 2), dec(90000, 10495001, 15, 2), dec(0, 11, 15, 2),
                         dec(10**9, 10**11, 38, 4), dec(10**11, 10**13, 38, 6)],
                        names=["l_returnflag", "l_linestatus", "qty", "price", "disc", "disc_price", "charge"])
    specs = [("sum", "qty", "sum_qty", None), ("sum", "price", "sum_base_price", None), ("sum", "disc_price", "sum_disc_price", None),
             ("sum", "charge", "sum_charge", None), ("avg", "qty", "avg_qty", None), ("avg", "price", "avg_price", None),
             ("avg", "disc", "avg_disc", None), ("count", None, "count_order", None)]
    got = gpu_agg(ctx, "single", [b], ["l_returnflag", "l_linestatus"], specs)
    want = oracle_agg("single", [b], ["l_returnflag", "l_linestatus"], specs)
    assert got.num_rows == 4
    assert_tables_equal(got, want, key_cols=[0, 1])
    # partial -> final over 3 partitions gives the same answer
    states = [gpu_agg(ctx, "partial", [b.slice(o, 70_000)], ["l_returnflag", "l_linestatus"], specs) for o in (0, 70_000, 140_000)]
    fin = gpu_agg(ctx, "final", states, [0, 1], [(f, None, nm, None) for f, _, nm, _ in specs])
    assert_tables_equal(fin, want, key_cols=[0, 1])


def test_force_hash_collisions():
    """cargo feature force_hash_collisions: every key collides, equality is load-bearing."""
    ctx = gpu_ctx(force_hash_collisions=True)
    rng = np.random.default_rng(3)
    for keys in (["int64"], ["utf8"], ["utf8", "int32"], ["int64", "date32", "int32"]):
        b = make_table(rng, 600, keys, 41, 0.1, 0.1)
        gk = [f"k{i}" for i in range(len(keys))]
        assert_tables_equal(gpu_agg(ctx, "single", [b], gk, SPECS[:6]), oracle_agg("single", [b], gk, SPECS[:6]),
                            key_cols=list(range(len(keys))))
    ctx.close()


def test_float_minmax_quirks_and_empty(ctx):
    import sys
    k = pa.array([1, 1, 2, 3, 3, 4], pa.int64())
    v = pa.array([float("nan"), float("nan"), float("-inf"), 1.5, None, None], pa.float64())
    b = pa.record_batch([k, v], names=["k", "v"])
    got = gpu_agg(ctx, "single", [b], ["k"], [("max", "v", "mx", None), ("min", "v", "mn", None), ("count", "v", "c", None)])
    fm = sys.float_info.max
    assert_rows_equal(rows_of(got), [(1, -fm, fm, 2), (2, -fm, float("-inf"), 1), (3, 1.5, 1.5, 1), (4, None, None, 0)])
    empty = gpu_agg(ctx, "single", [b.slice(0, 0)], ["k"], [("max", "v", "mx", None)])
    assert empty.num_rows == 0 and empty.schema.names == ["k", "mx"]


def test_decimal_avg_overflow_error(ctx):
    import datafusion_upstream_b200 as df
    t = pa.decimal128(38, 0)
    b = pa.record_batch([pa.array([1, 1], pa.int8()), pa.array([decimal.Decimal(10**37), decimal.Decimal(10**37)], t)], names=["k", "v"])
    with pytest.raises(df.DfGpuError) as e:
        gpu_agg(ctx, "single", [b], ["k"], [("avg", "v", "a", None)])
    assert e.value.kind == "Execution" and "Arithmetic Overflow in AvgAccumulator" in str(e.value)


def test_early_emit_under_memory_limit():
    """Partial mode + memory pressure => early emit (row_hash.rs:723-734); Final re-merges duplicates."""
    ctx = gpu_ctx(memory_limit=1 << 20, batch_size=64, chunk_rows=20_000)
    rng = np.random.default_rng(9)
    n = 200_000
    b = pa.record_batch([pa.array(rng.integers(0, 100_000, n), pa.int64()), pa.array(rng.integers(0, 9, n), pa.int64())], names=["k", "v"])
    specs = [("sum", "v", "s", None), ("count", None, "c", None)]
    st = gpu_agg(ctx, "partial", [b], ["k"], specs)
    assert st.num_rows > len(set(b.column(0).to_pylist()))          # some keys were emitted more than once
    ctx2 = gpu_ctx()
    fin = gpu_agg(ctx2, "final", [st], [0], [(f, None, nm, None) for f, _, nm, _ in specs])
    assert_tables_equal(fin, oracle_agg("single", [b], ["k"], specs), key_cols=[0])
    import datafusion_upstream_b200 as df
    with pytest.raises(df.DfGpuError) as e:
        gpu_agg(ctx, "single", [b], ["k"], specs)
    assert e.value.kind == "ResourcesExhausted" and "GroupedHashAggregateStream" in str(e.value)
    ctx.close(); ctx2.close()


def test_early_emit_releases_memory():
    """A high-cardinality spike followed by low-cardinality batches under a memory limit: after the early emit the operator must be
    back under its limit (the reference clear_shrink()s, row_hash.rs:680, 708-712), so the small batches - fewer than batch_size
    groups, which cannot trigger another early emit - go through instead of raising ResourcesExhausted."""
    ctx = gpu_ctx(memory_limit=1 << 20, batch_size=64, chunk_rows=20_000)
    rng = np.random.default_rng(10)
    spike = pa.record_batch([pa.array(rng.integers(0, 100_000, 60_000), pa.int64()), pa.array(rng.integers(0, 9, 60_000), pa.int64())], names=["k", "v"])
    small = [pa.record_batch([pa.array(rng.integers(0, 20, 5_000), pa.int64()), pa.array(rng.integers(0, 9, 5_000), pa.int64())], names=["k", "v"]) for _ in range(4)]
    specs = [("sum", "v", "s", None), ("count", None, "c", None)]
    st = gpu_agg(ctx, "partial", [spike] + small, ["k"], specs)
    ctx2 = gpu_ctx()
    fin = gpu_agg(ctx2, "final", [st], [0], [(f, None, nm, None) for f, _, nm, _ in specs])
    assert_tables_equal(fin, oracle_agg("single", [spike] + small, ["k"], specs), key_cols=[0])
    ctx.close(); ctx2.close()


@pytest.mark.parametrize("keys", [["int64"], ["utf8", "int32"]], ids=lambda k: "+".join(k))
def test_emit_first_n(ctx, keys):
    """emit(EmitTo::First(n)) (groups_accumulator.rs:24-58) in Partial mode: n groups at a time, pushes in between, emit() for the
    rest; the union of everything emitted, re-merged by Final, is the Single result.  Outside Partial mode: NotImplemented."""
    import datafusion_upstream_b200 as df
    rng = np.random.default_rng(21)
    specs = [("sum", "i", "s", None), ("count", None, "c", None), ("avg", "d", "a", None), ("min", "f", "m", None)]
    bs = [make_table(rng, n, keys, 700, 0.05, 0.1) for n in (5000, 3000, 4000)]
    gk = [f"k{i}" for i in range(len(keys))]
    exprs = [df.AggregateExpr(df.AggFunc[f.capitalize()], a, nm) for f, a, nm, _ in specs]
    op = df.GpuAggregateExec(ctx, df.AggregateMode.Partial, bs[0].schema, gk, exprs)
    parts = []
    op.push(bs[0])
    parts.append(op.emit_first(100)); assert parts[-1].num_rows == 100
    parts.append(op.emit_first(64)); assert parts[-1].num_rows == 64
    op.push(bs[1])                                   # new partial states next to the queued remainder
    parts.append(op.emit_first(10_000))              # more than the queue holds: what is left of it
    assert 0 < parts[-1].num_rows < 10_000
    op.push(bs[2])
    parts.append(op.emit_first(5))
    parts.append(op.emit())
    op.close()
    assert all(p.schema == parts[0].schema for p in parts)
    fin = gpu_agg(ctx, "final", parts, list(range(len(keys))), [(f, None, nm, None) for f, _, nm, _ in specs])
    assert_tables_equal(fin, oracle_agg("single", bs, gk, specs), key_cols=list(range(len(keys))))
    single = df.GpuAggregateExec(ctx, df.AggregateMode.Single, bs[0].schema, gk, exprs)
    single.push(bs[0])
    with pytest.raises(df.DfGpuError) as e:
        single.emit_first(10)
    assert e.value.kind == "NotImplemented"
    single.close()


def test_not_implemented_and_plan_errors(ctx):
    import datafusion_upstream_b200 as df
    b = pa.record_batch([pa.array([1], pa.int64()), pa.array(["x"])], names=["k", "s"])
    with pytest.raises(df.DfGpuError) as e:
        gpu_agg(ctx, "single", [b], ["k"], [("sum", "s", "s", None)])
    assert e.value.kind == "NotImplemented"
    with pytest.raises(df.DfGpuError) as e:
        gpu_agg(ctx, "single", [b], [7], [("count", None, "c", None)])
    assert e.value.kind == "Plan"


# ---- BIT_AND / BIT_OR / BIT_XOR (bit_and_or_xor.rs) and BOOL_AND / BOOL_OR (bool_and_or.rs) ----
BIT_SPECS = [
    ("bit_and", "i", "BIT_AND(i)", None), ("bit_or", "i", "BIT_OR(i)", None), ("bit_xor", "i", "BIT_XOR(i)", None),
    ("bit_and", "s", "BIT_AND(s)", None), ("bit_or", "u", "BIT_OR(u)", None), ("bit_xor", "b8", "BIT_XOR(b8)", None),
    ("bit_and", "u", "BIT_AND(u) FILTER", "flt"), ("bit_xor", "s", "BIT_XOR(s) FILTER", "flt"),
    ("bool_and", "flt", "BOOL_AND(flt)", None), ("bool_or", "flt", "BOOL_OR(flt)", None), ("count", None, "COUNT(*)", None),
    ("sum", "i", "SUM(i)", None),
]


def test_golden_bit_aggregate_slt(ctx):
    """aggregate.slt:1768-1813: grouped bit_and / bit_or / bit_xor over SMALLINT columns, `GROUP BY tag`."""
    g = AGGS["bit_aggregate_slt"]
    cols = list(zip(*g["rows"]))
    b = pa.record_batch([pa.array(list(cols[0]), pa.int16()), pa.array(list(cols[1]), pa.int16()), pa.array(list(cols[2]), pa.int16()),
                         pa.array(list(cols[3]), pa.string())], names=g["columns"])
    specs = [(f, n, f"{f}({n})", None) for f in ("bit_and", "bit_or", "bit_xor") for n in ("c1", "c2", "c3")]
    want = [tuple([r[-1]] + r[:-1]) for r in g["expected"]]
    out = gpu_agg(ctx, "single", [b], ["tag"], specs)
    assert [f.type for f in out.schema][1:] == [pa.int16()] * 9
    assert_rows_equal(rows_of(out), want)
    st = [gpu_agg(ctx, "partial", [b.slice(0, 2)], ["tag"], specs), gpu_agg(ctx, "partial", [b.slice(2)], ["tag"], specs)]
    assert st[0].schema.names[1] == "bit_and(c1)[bit_and]" and st[0].schema.names[9] == "bit_xor(c3)[bit_xor]"
    out = gpu_agg(ctx, "final", st, [0], [(f, None, n, None) for f, _, n, _ in specs])
    assert_rows_equal(rows_of(out), want)


@pytest.mark.parametrize("case", AGGS["bit_unit_tests"], ids=[c["name"] for c in AGGS["bit_unit_tests"]])
def test_golden_bit_unit_vectors(ctx, case):
    t = pa.uint32() if case["unsigned"] else pa.int32()      # bit_and_or_xor.rs:708-787
    v = pa.array(case["input"], t)
    b = pa.record_batch([pa.array([0] * len(v), pa.int8()), v], names=["k", "v"])
    out = gpu_agg(ctx, "single", [b], ["k"], [(case["op"], "v", "r", None)])
    assert out.schema.field(1).type == t
    assert out.column(1).to_pylist() == [case["expected"]]


@pytest.mark.parametrize("keys,card", [(["int64"], 37), (["utf8", "int32"], 37), (["int64"], 5000), (["dict_utf8"], 400)],
                         ids=["int64-private", "utf8+int32", "int64-global", "dict"])
@pytest.mark.parametrize("nulls", [False, True], ids=["nonull", "nulls"])
def test_bit_bool_ops_vs_oracle(ctx, keys, card, nulls):
    rng = np.random.default_rng(77 + card + nulls)
    batches = [make_table(rng, n, keys, card, 0.1 if nulls else 0.0, 0.2 if nulls else 0.0) for n in (3000, 1, 0, 9001)]
    gk = [f"k{i}" for i in range(len(keys))]
    got = gpu_agg(ctx, "single", batches, gk, BIT_SPECS)
    want = oracle_agg("single", batches, gk, BIT_SPECS)
    assert_tables_equal(got, want, key_cols=list(range(len(keys))), what="bit/bool single " + str(keys))
    # Partial per partition -> FinalPartitioned
    parts = [batches[:1], batches[1:]]
    g_states = [gpu_agg(ctx, "partial", p, gk, BIT_SPECS) for p in parts]
    o_states = [oracle_agg("partial", p, gk, BIT_SPECS) for p in parts]
    for gs, os_ in zip(g_states, o_states):
        assert_tables_equal(gs, os_, key_cols=list(range(len(keys))), what="bit/bool partial state")
    fs = [(f, None, n, None) for f, _, n, _ in BIT_SPECS]
    got = gpu_agg(ctx, "final_partitioned", g_states, list(range(len(keys))), fs)
    assert_rows_equal(rows_of(got), rows_of(want), key_cols=list(range(len(keys))))


def test_bit_ops_reject_unsupported_types(ctx):
    b = pa.record_batch([pa.array([1, 2], pa.int32()), pa.array([1.0, 2.0]), pa.array([1, 2], pa.int32())], names=["k", "f", "i"])
    with pytest.raises(Exception, match="not implemented|NotImplemented"):
        gpu_agg(ctx, "single", [b], ["k"], [("bit_and", "f", "x", None)])       # downcast_integer! -> not_impl_err
    with pytest.raises(Exception, match="not implemented|NotImplemented"):
        gpu_agg(ctx, "single", [b], ["k"], [("bool_or", "i", "x", None)])       # bool_and_or.rs:142-146


@pytest.mark.parametrize("keys", [["utf8", "int32"], ["int64"], ["large_utf8"], ["bool", "int16"]], ids=lambda k: "+".join(k))
@pytest.mark.parametrize("target", [3000, 64, -1], ids=["coalesce3000", "coalesce64", "off"])
def test_host_batch_coalescing(keys, target):
    """dfgpu_agg_push_arrow packs small host batches into a pinned arena (CoalesceBatchesExec at the PCIe boundary,
    dfgpu_options.coalesce_rows): slices with arbitrary offsets (validity / Boolean bit offsets, Utf8 offset rebasing),
    flushes in the middle of the stream, batches large enough to bypass the arena, empty batches."""
    ctx = gpu_ctx(coalesce_rows=target)
    rng = np.random.default_rng(len(keys) * 1000 + target + 5)
    whole = make_table(rng, 20_000, keys, 211, 0.1, 0.2)
    lens = [1, 7, 0, 500, 1499, 13, 1501, 64, 8, 8, 3, 2900, 29, 6000, 1, 1, 300]
    pieces, off = [], 0
    while off < whole.num_rows:
        for ln in lens:
            ln = min(ln, whole.num_rows - off)
            pieces.append(whole.slice(off, ln))
            off += ln
            if off >= whole.num_rows:
                break
    specs = SPECS + BIT_SPECS[:10]
    gk = [f"k{i}" for i in range(len(keys))]
    got = gpu_agg(ctx, "single", pieces, gk, specs)
    want = oracle_agg("single", [whole], gk, specs)
    assert_tables_equal(got, want, key_cols=list(range(len(keys))), what=f"coalescing {keys} target={target}")
    # num_groups() in the middle of the stream sees the rows still waiting in the arena
    import datafusion_upstream_b200 as df
    op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, whole.schema, gk, [df.AggregateExpr(df.AggFunc.Count, None, "c")])
    for p in pieces[:9]:
        op.push(p)
    seen = sum(p.num_rows for p in pieces[:9])
    ref = oracle_agg("single", [whole.slice(0, seen)], gk, [("count", None, "c", None)])
    assert op.num_groups() == ref.num_rows
    for p in pieces[9:]:
        op.push(p)
    out = op.emit()
    assert sum(out.column(len(keys)).to_pylist()) == whole.num_rows
    assert op.metrics()["input_batches"] == len(pieces) and op.metrics()["input_rows"] == whole.num_rows
    op.close()
    ctx.close()


@pytest.mark.parametrize("nulls", [0.0, 0.1], ids=["no_nulls_threaded_copy", "nulls_serial_copy"])
@pytest.mark.parametrize("target", [4096, 1 << 20, -1], ids=["coalesce4096", "default", "off"])
def test_push_many_equals_push(target, nulls):
    """dfgpu_agg_push_arrow_batches: n host batches in one call give what n pushes give - runs of small batches are packed
    into the arena together (several host threads when all columns are fixed-width and free of nulls), batches large enough
    to bypass the arena and empty batches in between keep their place in the order."""
    import datafusion_upstream_b200 as df
    ctx = gpu_ctx(coalesce_rows=target)
    rng = np.random.default_rng(4242 + int(nulls * 10))
    n = 300_000
    k = rng.integers(0, 5000, n)
    v = rng.integers(-1000, 1000, n)
    whole = pa.record_batch([pa.array(k, pa.int64(), mask=(rng.random(n) < nulls) if nulls else None), pa.array(v, pa.int32(), mask=(rng.random(n) < nulls) if nulls else None),
                             pa.array(rng.random(n), pa.float64())], names=["k", "v", "f"])
    lens = [8192] * 12 + [0, 100_000, 1, 8192, 77] + [8192] * 30
    pieces, off = [], 0
    for ln in lens:
        ln = min(ln, n - off)
        pieces.append(whole.slice(off, ln))
        off += ln
    if off < n:
        pieces.append(whole.slice(off))
    exprs = [df.AggregateExpr(df.AggFunc.Count, None, "c"), df.AggregateExpr(df.AggFunc.Sum, "v", "s"), df.AggregateExpr(df.AggFunc.Max, "f", "m"),
             df.AggregateExpr(df.AggFunc.Count, "v", "cv")]
    outs = []
    for many in (False, True):
        op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, whole.schema, ["k"], exprs)
        if many:
            op.push_many(pieces[:20])
            op.push_many(pieces[20:])
            op.push_many([])
        else:
            for p in pieces:
                op.push(p)
        outs.append(op.emit())
        m = op.metrics()
        assert m["input_rows"] == n and m["input_batches"] == len(pieces), m
        op.close()
    assert_tables_equal(outs[1], outs[0], key_cols=[0], what="push_many vs push")
    want = oracle_agg("single", [whole], ["k"], [("count", None, "c", None), ("sum", "v", "s", None), ("max", "f", "m", None), ("count", "v", "cv", None)])
    assert_tables_equal(outs[1], want, key_cols=[0], what="push_many vs oracle")
    ctx.close()


@pytest.mark.parametrize("shape", ["count_star_int64", "mixed_nullable", "q1_like"])
def test_pipelined_host_push(monkeypatch, shape):
    """A large host batch of fixed-width columns is copied chunk by chunk on a second stream while the kernels of the previous
    chunk run (DFGPU_PIPELINED_PUSH lowers the size at which that starts): same result as one import + push, validity bitmaps
    with bit offsets per chunk, metrics count one batch."""
    import datafusion_upstream_b200 as df
    monkeypatch.setenv("DFGPU_PIPELINED_PUSH", "50000")
    rng = np.random.default_rng(99)
    n = 333_337
    if shape == "count_star_int64":
        monkeypatch.setenv("DFGPU_FA_PART", "1")
        b = pa.record_batch([pa.array(rng.integers(0, 90_000, n), pa.int64())], names=["k"])
        gk, specs = ["k"], [("count", None, "c", None)]
    elif shape == "mixed_nullable":
        b = pa.record_batch([pa.array(rng.integers(0, 3000, n), pa.int32(), mask=rng.random(n) < 0.02), pa.array(rng.integers(-99, 99, n), pa.int64(), mask=rng.random(n) < 0.1),
                             pa.array(rng.random(n), pa.float64())], names=["k", "v", "f"])
        gk, specs = ["k"], [("count", None, "c", None), ("sum", "v", "s", None), ("avg", "f", "a", None), ("count", "v", "cv", None)]
    else:
        b = pa.record_batch([pa.array(rng.integers(65, 68, n).astype(np.uint8)), pa.array(rng.integers(70, 72, n).astype(np.uint8)),
                             pa.array(rng.integers(0, 10**6, n), pa.int64()).cast(pa.decimal128(25, 2))], names=["f", "s", "d"])
        gk, specs = ["f", "s"], [("sum", "d", "sd", None), ("avg", "d", "ad", None), ("count", None, "c", None)]
    ctx = gpu_ctx()
    try:
        exprs = [df.AggregateExpr(df.AggFunc(FUNCS[f]), a_, nm, fl) for f, a_, nm, fl in specs]
        op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, b.schema, gk, exprs)
        op.push(b.slice(3))            # (a slice: value and validity buffers start at an offset)
        op.push(b.slice(0, 3))
        got = op.emit()
        m = op.metrics()
        assert m["input_rows"] == n and m["input_batches"] == 2, m
        op.close()
        assert_tables_equal(got, oracle_agg("single", [b], gk, specs), key_cols=list(range(len(gk))), what=shape)
    finally:
        ctx.close()


PA_T = {"int32": pa.int32(), "uint32": pa.uint32(), "int64": pa.int64(), "uint64": pa.uint64(), "float32": pa.float32(),
        "float64": pa.float64()}


@pytest.mark.parametrize("case", AGGS["simple_unit_tests"], ids=[c["name"] for c in AGGS["simple_unit_tests"]])
def test_golden_simple_unit_vectors(ctx, case):
    """sum.rs:359-400, average.rs:631-673, count.rs:343-361, min_max.rs:1285-1406: known answers incl. all-NULL inputs."""
    t = PA_T[case["input_type"]]
    v = pa.array(case["input"], t)
    b = pa.record_batch([pa.array([0] * len(v), pa.int8()), v], names=["k", "v"])
    out = gpu_agg(ctx, "single", [b], ["k"], [(case["func"], "v", "r", None)])
    assert out.schema.field(1).type == PA_T[case["expected_type"]], case["source"]
    assert out.column(1).to_pylist() == [case["expected"]], case["source"]
