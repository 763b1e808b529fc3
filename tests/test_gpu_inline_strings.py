"""Utf8 / Binary key columns whose values are short: a string of <= 3 bytes is its own id inside the packed key image (the inline
rule of ArrowBytesMap, physical-expr/src/binary_map.rs:381-393, at the width of a 4-byte field), so the two-level path reads such
columns in place - as plain fixed-width columns when every value has the same length of 1 or 2 bytes (TPC-H Q1's l_returnflag /
l_linestatus, BASELINE configs[0]) - and no interner runs.  Every case against the oracle; the kernel profile says which path ran."""
import numpy as np
import pyarrow as pa
import pytest

from gpu_helpers import FUNCS, gpu_agg, gpu_ctx, oracle_agg, random_values
from util import assert_tables_equal

pytestmark = pytest.mark.gpu

SPECS = [("count", None, "c", None), ("sum", "v", "s", None), ("avg", "d", "a", None), ("min", "v", "mn", None)]


@pytest.fixture(scope="module")
def ctx():
    c = gpu_ctx()
    yield c
    c.close()


def run(ctx, batches, gb, specs=SPECS):
    ctx.set_profiling(True)
    try:
        got = gpu_agg(ctx, "single", batches, gb, specs)
        names = {r["name"] for r in ctx.profile()}
    finally:
        ctx.set_profiling(False)
    assert_tables_equal(got, oracle_agg("single", batches, gb, specs), key_cols=list(range(len(gb))))
    return names


def values(rng, n):
    return [random_values(rng, n, "int64", 0.1), random_values(rng, n, "decimal", 0.1)]


@pytest.mark.parametrize("typ", [pa.string(), pa.large_string(), pa.binary()], ids=["utf8", "large_utf8", "binary"])
def test_q1_shape_one_byte_keys_read_in_place(ctx, typ):
    rng = np.random.default_rng(1)
    n = 200_000
    pick = rng.choice(4, n, p=[0.25, 0.006, 0.493, 0.251])
    conv = (lambda x: x.encode()) if typ == pa.binary() else (lambda x: x)
    flag = pa.array([conv("ANNR"[i]) for i in pick], typ)
    status = pa.array([conv("FFOF"[i]) for i in pick], typ)
    b = pa.record_batch([flag, status] + values(rng, n), names=["f", "s", "v", "d"])
    names = run(ctx, [b, b.slice(1234, 50_001)], ["f", "s"])     # (a slice: offsets do not start at 0)
    assert "k_fa_l1_reg" in names and "k_intern_rows<var>" not in names, names


def test_short_strings_of_varying_length_with_nulls(ctx):
    rng = np.random.default_rng(2)
    n = 120_000
    words = ["", "a", "b", "ab", "abc", "zzz", "é", "xy"]        # "é" is two bytes
    k = pa.array([words[i] for i in rng.integers(0, len(words), n)], pa.string(), mask=rng.random(n) < 0.05)
    r = pa.array(rng.integers(0, 40, n).astype(np.int16), pa.int16())     # 2 + 4 bytes + a null bit: a 64-bit image
    b = pa.record_batch([r, k] + values(rng, n), names=["r", "k", "v", "d"])
    names = run(ctx, [b], ["r", "k"])
    assert "k_intern_rows<var>" not in names and ("k_fa_l1" in names or "k_fa_l1_reg" in names), names
    names = run(ctx, [b], ["k"])                                 # a lone string key packs the same way
    assert "k_intern_rows<var>" not in names, names


def test_two_byte_constant_length_odd_and_even_alignment(ctx):
    rng = np.random.default_rng(3)
    n = 50_000
    codes = ["AA", "AB", "BA", "ZZ", "qq"]
    k = pa.array([codes[i] for i in rng.integers(0, 5, n)], pa.string())
    pad = pa.array(["x"] + [codes[i] for i in rng.integers(0, 5, n)], pa.string()).slice(1)   # data starts at an odd byte
    for col in (k, pad):
        b = pa.record_batch([col] + values(rng, n), names=["k", "v", "d"])
        run(ctx, [b], ["k"])


def test_batches_with_and_without_long_strings_share_ids(ctx):
    """Batch 1 is read in place, batch 2 has long strings and goes through the interner: "a" must be the same group in both."""
    rng = np.random.default_rng(4)
    short = pa.array([["a", "b", "", "abc"][i] for i in rng.integers(0, 4, 30_000)], pa.string())
    mixed = pa.array([["a", "b", "the quick brown fox", "abcd", "", "abc"][i] for i in rng.integers(0, 6, 30_000)], pa.string())
    bs = [pa.record_batch([c, pa.array(rng.integers(0, 3, 30_000).astype(np.int16))] + values(rng, 30_000), names=["k", "j", "v", "d"]) for c in (short, mixed, short)]
    names = run(ctx, bs, ["k", "j"])
    assert "k_intern_rows<var>" in names
    names = run(ctx, bs[:1], ["k", "j"])
    assert "k_intern_rows<var>" not in names


def test_partial_final_and_table_path_agree(ctx, monkeypatch):
    rng = np.random.default_rng(5)
    n = 40_000
    b = pa.record_batch([pa.array([["N", "O", "F", ""][i] for i in rng.integers(0, 4, n)], pa.string(), mask=rng.random(n) < 0.1)] + values(rng, n),
                        names=["k", "v", "d"])
    st = gpu_agg(ctx, "partial", [b], ["k"], SPECS)
    assert_tables_equal(st, oracle_agg("partial", [b], ["k"], SPECS), key_cols=[0])
    fin = [(f, None, nm, None) for f, _, nm, _ in SPECS]
    assert_tables_equal(gpu_agg(ctx, "final", [st], [0], fin), oracle_agg("single", [b], ["k"], SPECS), key_cols=[0])
    monkeypatch.setenv("DFGPU_DISABLE_TWO_LEVEL", "1")
    c2 = gpu_ctx()
    try:
        assert_tables_equal(gpu_agg(c2, "single", [b], ["k"], SPECS), oracle_agg("single", [b], ["k"], SPECS), key_cols=[0])
        b2 = pa.record_batch([b.column(0), pa.array(rng.integers(0, 9, n))] + values(rng, n), names=["k", "j", "v", "d"])   # packed image on the table path
        assert_tables_equal(gpu_agg(c2, "single", [b2], ["k", "j"], SPECS), oracle_agg("single", [b2], ["k", "j"], SPECS), key_cols=[0, 1])
    finally:
        c2.close()


def test_short_string_keys_cross_the_record_exchange(ctx):
    """Inline ids mean the same on every rank, so the fused sharded aggregation takes such keys (virtual ranks); one rank with a long
    string makes everybody refuse, before anything moves."""
    import datafusion_upstream_b200 as df
    rng = np.random.default_rng(6)
    exprs = [df.AggregateExpr(df.AggFunc(FUNCS[f]), a, nm, flt) for f, a, nm, flt in SPECS]

    def shard(long_one=False):
        n = 30_000
        w = ["A", "N", "R", "", "OK"] + (["a long string"] if long_one else [])
        return pa.record_batch([pa.array([w[i] for i in rng.integers(0, len(w), n)], pa.string()), pa.array(rng.integers(0, 5, n).astype(np.int8))] + values(rng, n),
                               names=["k", "j", "v", "d"])
    shards = [shard() for _ in range(4)]
    ops = [df.GpuAggregateExec(ctx, df.AggregateMode.Single, shards[0].schema, ["k", "j"], exprs) for _ in shards]
    for op, s in zip(ops, shards):
        op.push(s)
    df.GpuAggregateExec.exchange_virtual(ops)
    outs = [op.emit() for op in ops]
    got = pa.Table.from_batches([o for o in outs if o.num_rows])
    assert_tables_equal(got, oracle_agg("single", shards, ["k", "j"], SPECS), key_cols=[0, 1])
    shards[2] = shard(long_one=True)
    for op, s in zip(ops, shards):
        op.push(s)
    with pytest.raises(df.DfGpuError) as e:
        df.GpuAggregateExec.exchange_virtual(ops)
    assert e.value.kind == "NotImplemented"
    for op in ops:
        op.close()
