"""Parity of the two-level aggregation path (fastagg.h: shared-memory / register pre-aggregation,
radix partition, per-partition merge) with the oracle, on every key shape it accepts, in every
level-1 mode, with overflow rounds and spills forced, and against the table path on the same input.

The library picks the path itself (64-bit key images, no MIN/MAX over Decimal128, no memory limit);
the environment hooks below only shrink its tables so the rare branches run on small inputs.
"""
import numpy as np
import pyarrow as pa
import pytest

from gpu_helpers import gpu_agg, gpu_ctx, oracle_agg, random_keys
from test_gpu_agg import SPECS, make_table
from util import assert_rows_equal, assert_tables_equal, rows_of

pytestmark = pytest.mark.gpu

# everything the two-level path supports: MIN/MAX over Decimal128 (128-bit CAS state) stays on the table path
# and the operator takes the path only with <= 24 state ops (FA_MAX_OPS; AVG = 2 ops): AVG(s) is left out to stay within it
FAST_SPECS = [s for s in SPECS if not (s[0] in ("min", "max") and s[1] in ("d", "big")) and s[:2] != ("avg", "s")]
assert sum(2 if s[0] == "avg" else 1 for s in FAST_SPECS) <= 24


def ran_two_level(c, fn):
    """Run fn() with kernel profiling on and report whether any two-level kernel (k_fa_l1 / k_fa_l1_reg / k_fa_l2) ran."""
    c.set_profiling(True)
    try:
        out = fn()
        names = {r["name"] for r in c.profile()}
    finally:
        c.set_profiling(False)
    return out, any(n.startswith("k_fa_l") for n in names), sorted(names)
FAST_KEYS = [["int64"], ["int64_small"], ["uint64"], ["int32"], ["int16", "int8"], ["float32"], ["date32"], ["timestamp"],
             ["uint8", "uint8"], ["int32", "int16", "bool"], ["bool"], ["float64"]]

HOOKS = {
    "default": {},
    "tiny_tables": {"DFGPU_FA_S2MAX": "64", "DFGPU_FA_S1": "32"},    # spills at level 1, overflow rounds at level 2
    "no_register_mode": {"DFGPU_FA_NOREG": "1"},
    "always_partition": {"DFGPU_FA_SINGLE_MAX": "0"},                 # small record sets through hist -> scan -> scatter -> level 2 too
    "table_path": {"DFGPU_DISABLE_TWO_LEVEL": "1"},                   # same specs through GroupTable + k_accumulate
}


@pytest.fixture(scope="module")
def ctx():
    c = gpu_ctx()
    yield c
    c.close()


@pytest.mark.parametrize("hook", list(HOOKS), ids=list(HOOKS))
@pytest.mark.parametrize("card", [3, 37, 5000], ids=lambda c: f"card{c}")
@pytest.mark.parametrize("keys", FAST_KEYS, ids=["+".join(k) for k in FAST_KEYS])
def test_single_vs_oracle(ctx, monkeypatch, keys, card, hook):
    for k, v in HOOKS[hook].items():
        monkeypatch.setenv(k, v)
    rng = np.random.default_rng(sum(map(ord, "".join(keys))) + card)
    nulls = card != 37
    batches = [make_table(rng, n, keys, card, 0.1 if nulls else 0.0, 0.15 if nulls else 0.0) for n in (3000, 1, 0, 9001)]
    gk = [f"k{i}" for i in range(len(keys))]
    got, fast, names = ran_two_level(ctx, lambda: gpu_agg(ctx, "single", batches, gk, FAST_SPECS))
    assert fast == (hook != "table_path"), f"path taken: {names}"      # the library picks the path; make sure it is the one under test
    want = oracle_agg("single", batches, gk, FAST_SPECS)
    assert_tables_equal(got, want, key_cols=list(range(len(keys))), what=f"{keys} card={card} {hook}")


@pytest.mark.parametrize("hook", ["default", "tiny_tables"])
@pytest.mark.parametrize("keys", [["int64_small"], ["int16", "int8"], ["uint8", "uint8"]], ids=lambda k: "+".join(k))
def test_partial_final_vs_oracle(ctx, monkeypatch, keys, hook):
    for k, v in HOOKS[hook].items():
        monkeypatch.setenv(k, v)
    rng = np.random.default_rng(23)
    gk = [f"k{i}" for i in range(len(keys))]
    parts = [[make_table(rng, n, keys, 211, 0.05, 0.1) for n in (1500, 700)] for _ in range(3)]
    g_states = [gpu_agg(ctx, "partial", p, gk, FAST_SPECS) for p in parts]
    o_states = [oracle_agg("partial", p, gk, FAST_SPECS) for p in parts]
    for gs, os_ in zip(g_states, o_states):
        assert_tables_equal(gs, os_, key_cols=list(range(len(keys))), what="partial state")
    fin = [(f, None, n, None) for f, _, n, _ in FAST_SPECS]
    got = gpu_agg(ctx, "final_partitioned", g_states, list(range(len(keys))), fin)
    want = oracle_agg("final_partitioned", o_states, list(range(len(keys))), fin)
    assert_tables_equal(got, want, key_cols=list(range(len(keys))), what="final")
    single = oracle_agg("single", [b for p in parts for b in p], gk, FAST_SPECS)
    assert_rows_equal(rows_of(got), rows_of(single), key_cols=list(range(len(keys))))


def test_side_groups_null_and_all_ones_key(ctx):
    """The NULL key and the key whose image is 0xFFFF_FFFF_FFFF_FFFF (-1) live outside the hash tables."""
    rng = np.random.default_rng(3)
    n = 50_000
    raw = rng.integers(-2, 3, n)                      # -2..2: includes -1
    keys = pa.array(raw, pa.int64(), mask=rng.random(n) < 0.2)
    v = pa.array(rng.integers(-1000, 1000, n), pa.int64(), mask=rng.random(n) < 0.1)
    b = pa.record_batch([keys, v], names=["k", "v"])
    specs = [("count", None, "c", None), ("count", "v", "cv", None), ("sum", "v", "s", None), ("min", "v", "mn", None),
             ("max", "v", "mx", None), ("avg", "v", "a", None)]
    for mode_batches in ([b], [b.slice(0, 7), b.slice(7, 20_000), b.slice(20_007)]):
        got = gpu_agg(ctx, "single", mode_batches, ["k"], specs)
        want = oracle_agg("single", [b], ["k"], specs)
        assert got.num_rows == 6
        assert_tables_equal(got, want, key_cols=[0])
    # only side groups present
    only = pa.record_batch([pa.array([-1, None, -1, None], pa.int64()), pa.array([1, 2, 3, 4], pa.int64())], names=["k", "v"])
    assert_tables_equal(gpu_agg(ctx, "single", [only], ["k"], specs), oracle_agg("single", [only], ["k"], specs), key_cols=[0])


@pytest.mark.parametrize("s2max", ["16", "64", "256"])
def test_overflow_rounds_never_duplicate_a_key(monkeypatch, s2max):
    """Many rows per key and level-2 tables far smaller than the key set: every partition goes through overflow
    rounds while thousands of lanes insert the same keys concurrently.  A key must come out of exactly one round
    (an occupancy-counter criterion once let one lane hand a key to the overflow round while another inserted it:
    two output rows for one key)."""
    monkeypatch.setenv("DFGPU_FA_S2MAX", s2max)
    monkeypatch.setenv("DFGPU_FA_S1", "32")
    monkeypatch.setenv("DFGPU_FA_SINGLE_MAX", "0")
    c = gpu_ctx()
    try:
        rng = np.random.default_rng(int(s2max))
        n, card = 600_000, 1500
        # all keys in few level-2 partitions: sequential small integers hash everywhere, so take keys whose top hash
        # bits collide rarely but use few distinct keys with ~400 rows each
        k = pa.array(rng.integers(0, card, n).astype(np.int64) * 7919 - 3, pa.int64())
        v = pa.array(rng.integers(-50, 50, n), pa.int64())
        b = pa.record_batch([k, v], names=["k", "v"])
        specs = [("count", None, "c", None), ("sum", "v", "s", None), ("max", "v", "mx", None)]
        for _ in range(3):
            got = gpu_agg(c, "single", [b], ["k"], specs)
            assert got.num_rows == card
            assert len(set(got.column(0).to_pylist())) == card
            assert sum(got.column(1).to_pylist()) == n
        assert_tables_equal(got, oracle_agg("single", [b], ["k"], specs), key_cols=[0])
    finally:
        c.close()


def test_register_mode_q1_shape_and_late_keys(ctx):
    """Q1 shape with fixed-width keys (register mode), value nulls, a filter, and keys that first
    appear after the sample the device used to choose the mode (exact per-row slow path)."""
    import decimal
    rng = np.random.default_rng(9)
    n = 120_000
    pick = rng.choice(4, n, p=[0.25, 0.006, 0.493, 0.251])
    flag = np.array([65, 78, 78, 82], np.uint8)[pick]
    status = np.array([70, 70, 79, 70], np.uint8)[pick]
    flag[100_000:100_050] = 90          # a 5th and 6th group long after the probe sample
    status[110_000:110_010] = 91
    dec = lambda lo, hi, p, s, nf: pa.array([decimal.Decimal(int(x)).scaleb(-s) for x in rng.integers(lo, hi, n)], pa.decimal128(p, s),
                                            mask=(rng.random(n) < nf) if nf else None)
    b = pa.record_batch([pa.array(flag), pa.array(status, pa.uint8(), mask=np.arange(n) == 115_000), dec(100, 5001, 15, 2, 0),
                         dec(90000, 10495001, 15, 2, 0.05),
                         pa.array([decimal.Decimal(int(a) * 10**13 + int(c)).scaleb(-6) for a, c in
                                   zip(rng.integers(-(10**17), 10**17, n), rng.integers(0, 10**13, n))], pa.decimal128(38, 6)),
                         pa.array(rng.random(n) < 0.5),
                         pa.array(rng.normal(size=n))],
                        names=["f", "s", "qty", "price", "charge", "flt", "x"])
    specs = [("sum", "qty", "sum_qty", None), ("avg", "qty", "avg_qty", None), ("sum", "price", "sum_price", None),
             ("avg", "price", "avg_price", None), ("sum", "charge", "sum_charge", None), ("count", None, "cnt", None),
             ("count", "price", "cnt_price", None), ("sum", "qty", "sum_qty_f", "flt"), ("min", "x", "min_x", None),
             ("avg", "x", "avg_x", None)]
    got = gpu_agg(ctx, "single", [b], ["f", "s"], specs)
    want = oracle_agg("single", [b], ["f", "s"], specs)
    assert got.num_rows == want.num_rows and got.num_rows > 4
    assert_tables_equal(got, want, key_cols=[0, 1])


@pytest.mark.parametrize("general", ["0", "1"], ids=["one_byte_key_form", "general_key_phase"])
@pytest.mark.parametrize("shape", ["u8_u8", "u8", "utf8_utf8", "sliced_unaligned"])
def test_register_mode_one_byte_keys(ctx, monkeypatch, shape, general):
    """Register mode maps one-byte key columns to groups four rows at a time (one 32-bit load per column, 16-bit codes): row
    counts that are not a multiple of four, groups that first appear late (claimed from inside a quad), a slice whose key
    buffers are not 4-byte aligned (takes the general form), one and two key columns, one-byte strings read in place."""
    monkeypatch.setenv("DFGPU_FA_NOK8", general)
    rng = np.random.default_rng(31)
    n = 262_147
    pick = rng.choice(4, n, p=[0.25, 0.006, 0.493, 0.251])
    f = np.array([65, 78, 78, 82], np.uint8)[pick]
    st = np.array([70, 70, 79, 70], np.uint8)[pick]
    f[200_001:200_007] = 90                 # a 5th group long after the probe sample: slow rows
    f[-1] = 66
    v = pa.array(rng.integers(-10**6, 10**6, n), pa.int64())
    if shape == "utf8_utf8":
        keys = [pa.array([chr(c) for c in f]), pa.array([chr(c) for c in st])]
    else:
        keys = [pa.array(f), pa.array(st)]
    if shape == "u8":
        b = pa.record_batch([keys[0], v], names=["f", "v"])
        gk = ["f"]
    else:
        b = pa.record_batch(keys + [v], names=["f", "s", "v"])
        gk = ["f", "s"]
    if shape == "sliced_unaligned":
        b = b.slice(1)
    specs = [("count", None, "c", None), ("sum", "v", "sv", None), ("avg", "v", "av", None), ("min", "v", "mn", None)]
    ctx.set_profiling(True)
    got = gpu_agg(ctx, "single", [b], gk, specs)
    names = {x["name"] for x in ctx.profile()}
    ctx.set_profiling(False)
    assert "k_fa_l1_reg" in names, names
    assert_tables_equal(got, oracle_agg("single", [b], gk, specs), key_cols=list(range(len(gk))))


def test_num_groups_mid_stream_then_more_input(ctx):
    """num_groups() compacts the record store; the operator keeps accepting input afterwards."""
    import datafusion_upstream_b200 as df
    rng = np.random.default_rng(4)
    mk = lambda n: pa.record_batch([random_keys(rng, n, "int64_small", 1000, 0.05), pa.array(rng.integers(0, 9, n), pa.int64())], names=["k", "v"])
    b1, b2 = mk(20_000), mk(30_000)
    op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, b1.schema, ["k"],
                             [df.AggregateExpr(df.AggFunc.Count, None, "c"), df.AggregateExpr(df.AggFunc.Sum, "v", "s"),
                              df.AggregateExpr(df.AggFunc.Avg, "v", "a")])
    op.push(b1)
    g1 = op.num_groups()
    assert g1 == len(set(b1.column(0).to_pylist()))
    op.push(b2)
    assert op.num_groups() == len(set(b1.column(0).to_pylist()) | set(b2.column(0).to_pylist()))
    got = op.emit()
    op.close()
    want = oracle_agg("single", [b1, b2], ["k"], [("count", None, "c", None), ("sum", "v", "s", None), ("avg", "v", "a", None)])
    assert_tables_equal(got, want, key_cols=[0])


@pytest.mark.parametrize("dist", ["zipf", "uniform", "all_distinct"])
def test_clickbench_shape_properties_at_scale(ctx, dist):
    """Config C2 at 20 M rows (too big for the oracle in seconds): size-independent properties —
    counts add up to the row count, keys are distinct, group count equals numpy's, and a checksum of
    (key * count) matches the checksum of the input keys."""
    rng = np.random.default_rng(8)
    n = 20_000_000
    if dist == "zipf":
        r = np.exp(rng.random(n) * np.log(8_000_000)).astype(np.uint64)
    elif dist == "uniform":
        r = rng.integers(0, 3_000_000, n).astype(np.uint64)
    else:
        r = np.arange(n, dtype=np.uint64)
    with np.errstate(over="ignore"):
        keys = (r * np.uint64(0x9E3779B97F4A7C15)).view(np.int64)
    b = pa.record_batch([pa.array(keys)], names=["UserID"])
    got = gpu_agg(ctx, "single", [b], ["UserID"], [("count", None, "COUNT(*)", None)])
    k = got.column(0).to_numpy()
    c = got.column(1).to_numpy()
    assert c.sum() == n
    assert len(np.unique(k)) == len(k) == len(np.unique(keys))
    with np.errstate(over="ignore"):
        assert (k.view(np.uint64) * c.astype(np.uint64)).sum() == keys.view(np.uint64).sum()


# ---- Utf8 / Binary key columns as 4-byte string-id fields of the key image; NOT NULL columns carry no null bit ----
def _with_nullability(batch, not_null):
    """Re-declare the listed columns NOT NULL (they hold no nulls)."""
    fields = [pa.field(f.name, f.type, nullable=f.name not in not_null) for f in batch.schema]
    return pa.record_batch(batch.columns, schema=pa.schema(fields))


STRING_KEYS = [
    (["int32", "utf8"], ["k0"]),             # ClickBench (RegionID NOT NULL, SearchPhrase): 4 + 4 bytes, no null bit
    (["utf8", "utf8"], []),                  # two string ids, NULL encoded inside each field
    (["int16", "large_utf8", "int8"], ["k0", "k2"]),
    (["binary", "int32"], ["k1"]),
    (["int32", "int16"], ["k0", "k1"]),      # no strings: NOT NULL only changes the image layout
    (["utf8"], []),                          # a lone string key: packed image = its string id
    (["large_utf8"], ["k0"]),
    (["binary"], []),
    (["dict_utf8"], []),                     # Dictionary(Int32, Utf8): the dictionary is interned once per batch
    (["dict_utf8", "int32"], ["k1"]),
]


@pytest.mark.parametrize("hook", ["default", "tiny_tables", "always_partition", "table_path"])
@pytest.mark.parametrize("card", [5, 3000], ids=lambda c: f"card{c}")
@pytest.mark.parametrize("keys,not_null", STRING_KEYS, ids=["+".join(k) for k, _ in STRING_KEYS])
def test_string_id_fields_vs_oracle(monkeypatch, keys, not_null, card, hook):
    for k, v in HOOKS[hook].items():
        monkeypatch.setenv(k, v)
    c = gpu_ctx()
    try:
        rng = np.random.default_rng(sum(map(ord, "".join(keys))) + card + 1)
        batches = []
        for n in (4000, 1, 0, 12_001):
            cols, names = [], []
            for i, k in enumerate(keys):
                cols.append(random_keys(rng, n, k, card, 0.0 if f"k{i}" in not_null else 0.1))
                names.append(f"k{i}")
            t = make_table(rng, n, ["int8"], 3, 0.0, 0.15)           # value / filter columns
            b = pa.record_batch(cols + t.columns[1:], names=names + t.schema.names[1:])
            batches.append(_with_nullability(b, not_null))
        gk = [f"k{i}" for i in range(len(keys))]
        got, fast, names = ran_two_level(c, lambda: gpu_agg(c, "single", batches, gk, FAST_SPECS))
        assert fast == (hook != "table_path"), f"path taken: {names}"
        want = oracle_agg("single", batches, gk, FAST_SPECS)
        assert_tables_equal(got, want, key_cols=list(range(len(keys))), what=f"{keys} card={card} {hook}")
        assert [f.nullable for f in got.schema][:len(keys)] == [f"k{i}" not in not_null for i in range(len(keys))]
        # Partial per partition -> FinalPartitioned (state batches keep the key fields' nullability)
        parts = [batches[:1], batches[1:]]
        states = [gpu_agg(c, "partial", p, gk, FAST_SPECS) for p in parts]
        fin = [(f, None, n, None) for f, _, n, _ in FAST_SPECS]
        got2 = gpu_agg(c, "final_partitioned", states, list(range(len(keys))), fin)
        assert_rows_equal(rows_of(got2), rows_of(want), key_cols=list(range(len(keys))))
    finally:
        c.close()


def test_string_keys_refuse_record_exchange():
    """String ids are local to a rank's interner: the fused record exchange must refuse (the caller keeps the
    three-operator plan), before any communication."""
    import datafusion_upstream_b200 as df
    c = gpu_ctx()
    b = pa.record_batch([pa.array([1, 2], pa.int32()), pa.array(["a", "b"])], names=["r", "s"])
    op = df.GpuAggregateExec(c, df.AggregateMode.Single, b.schema, ["r", "s"], [df.AggregateExpr(df.AggFunc.Count, None, "c")])
    op.push(b)
    op.exchange()       # world == 1: nothing to exchange, no error
    assert op.emit().num_rows == 2
    op.close()
    c.close()
