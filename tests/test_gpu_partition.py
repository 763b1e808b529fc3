"""-m gpu parity tests for create_hashes and the hash partitioner / exchange (RepartitionExec)."""
import numpy as np
import pyarrow as pa
import pytest

import oracle
from gpu_helpers import gpu_agg, gpu_ctx, oracle_agg, random_keys, random_values
from util import assert_rows_equal, assert_tables_equal, rows_of

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = gpu_ctx()
    yield c
    c.close()


KEYS = [["int64"], ["int32"], ["int8", "uint8"], ["float64"], ["float32"], ["bool"], ["date32"], ["timestamp"], ["decimal"],
        ["utf8"], ["large_utf8"], ["binary"], ["dict_utf8"], ["utf8", "int64"], ["dict_utf8", "decimal", "bool"]]


@pytest.mark.parametrize("keys", KEYS, ids=lambda k: "+".join(k))
@pytest.mark.parametrize("nulls", [0.0, 0.2])
def test_create_hashes_bit_exact(ctx, keys, nulls):
    """The device hash equals the oracle's dfhash v1 bit for bit (needed so partition ownership agrees
    on every rank and for both join sides)."""
    import datafusion_upstream_b200 as df
    rng = np.random.default_rng(len(keys) * 100 + int(nulls * 10))
    cols = [random_keys(rng, 3000, k, 500, nulls) for k in keys]
    b = pa.record_batch(cols, names=[f"c{i}" for i in range(len(cols))])
    got = df.create_hashes(ctx, b, list(range(len(cols))), seed=0)
    want = oracle.create_hashes(cols, seed=0)
    assert (got == want).all()
    got7 = df.create_hashes(ctx, b.slice(100, 1000), list(range(len(cols))), seed=7)   # sliced (offset) input
    assert (got7 == oracle.create_hashes([c.slice(100, 1000) for c in cols], seed=7)).all()


def test_hash_properties(ctx):
    import datafusion_upstream_b200 as df
    s = pa.array(["foo", None, "bar", "foo", "", None])
    b = pa.record_batch([s, s.dictionary_encode()], names=["s", "d"])
    h, hd = df.create_hashes(ctx, b, [0]), df.create_hashes(ctx, b, [1])
    assert (h == hd).all() and h[0] == h[3] and h[1] == 0 and h[5] == 0 and h[0] != h[2]   # hash_utils.rs:520-562


@pytest.mark.parametrize("keys", [["int64"], ["utf8", "int32"], ["dict_utf8"], ["decimal"]], ids=lambda k: "+".join(k))
@pytest.mark.parametrize("nparts", [1, 4, 7, 64])
def test_partition_matches_oracle(ctx, keys, nparts):
    import datafusion_upstream_b200 as df
    rng = np.random.default_rng(nparts)
    n = 20_000
    cols = [random_keys(rng, n, k, 300, 0.1) for k in keys]
    cols += [random_values(rng, n, "decimal", 0.1), pa.array([None if i % 5 == 0 else "v" * (i % 9) for i in range(n)]),
             pa.array(rng.random(n) < 0.5, pa.bool_()), pa.array(np.arange(n), pa.int64())]
    b = pa.record_batch(cols, names=[f"c{i}" for i in range(len(cols))])
    kidx = list(range(len(keys)))
    got = df.GpuRepartitionExec(ctx, b.schema, kidx, nparts).partition(b)
    want = oracle.partition(b, kidx, nparts)
    assert sum(p.num_rows for p in got if p is not None) == n          # repartition/mod.rs:1010-1031
    for g, w in zip(got, want):
        assert (g is None) == (w is None)                               # empty partitions are not produced
        if g is not None:
            assert_tables_equal(g, w, ordered=True)                     # input order preserved inside a partition


def test_partial_exchange_final_virtual_ranks(ctx):
    """N logical shards on one device (SURVEY §8e "test without 8 GPUs"): Partial per shard ->
    hash partition by owner -> concat per owner -> FinalPartitioned == Single over everything."""
    import datafusion_upstream_b200 as df
    rng = np.random.default_rng(12)
    R = 4
    specs = [("count", None, "c", None), ("sum", "v", "s", None), ("avg", "d", "a", None), ("min", "v", "m", None)]
    shards = [pa.record_batch([random_keys(rng, 5000, "int64", 900, 0.05), random_keys(rng, 5000, "utf8", 5, 0.05),
                               random_values(rng, 5000, "int64", 0.1), random_values(rng, 5000, "decimal", 0.1)],
                              names=["k", "s", "v", "d"]) for _ in range(R)]
    inbox = [[] for _ in range(R)]
    for sh in shards:
        op = df.GpuAggregateExec(ctx, df.AggregateMode.Partial, sh.schema, ["k", "s"],
                                 [df.AggregateExpr(df.AggFunc({"count": 2, "sum": 1, "avg": 3, "min": 4}[f]), a, n) for f, a, n, _ in specs])
        op.push(sh)
        st = op.emit_device()
        parts = df.GpuRepartitionExec(ctx, op.schema, [0, 1], R).partition_device(st)
        for r, p in enumerate(parts):
            if p is not None:
                inbox[r].append(p)
        state_schema = op.schema
    finals = []
    for r in range(R):
        if not inbox[r]:
            continue
        merged = df.DeviceBatch.concat(ctx, inbox[r])
        fin = df.GpuAggregateExec(ctx, df.AggregateMode.FinalPartitioned, state_schema, [0, 1],
                                  [df.AggregateExpr(df.AggFunc({"count": 2, "sum": 1, "avg": 3, "min": 4}[f]), None, n) for f, _, n, _ in specs])
        fin.push(merged)
        finals.append(fin.emit())
    want = oracle_agg("single", shards, ["k", "s"], specs)
    assert_rows_equal(rows_of(finals), rows_of(want), key_cols=[0, 1])


# ---- FilterExec::batch_filter on the device ----
@pytest.mark.parametrize("n,p_true,p_null", [(0, 0.5, 0.1), (1, 1.0, 0.0), (31, 0.5, 0.2), (4097, 0.02, 0.0), (100_003, 0.9, 0.05),
                                              (50_000, 0.0, 0.0), (50_000, 1.0, 0.0)])
def test_filter_vs_oracle(n, p_true, p_null):
    import datafusion_upstream_b200 as df
    rng = np.random.default_rng(n + 3)
    ctx = gpu_ctx()
    cols = [random_values(rng, n, "int64", 0.1), random_keys(rng, n, "utf8", 50, 0.1), random_values(rng, n, "decimal", 0.0),
            pa.array(rng.random(n) < 0.5, pa.bool_()), random_keys(rng, n, "dict_utf8", 20, 0.1),
            pa.array(rng.random(n) < p_true, pa.bool_(), mask=(rng.random(n) < p_null) if p_null else None)]
    b = pa.record_batch(cols, names=["i", "s", "d", "flag", "dict", "pred"])
    # a sliced input: non-zero bit offsets on the predicate and the validity bitmaps
    for batch in ([b] if n < 10 else [b, b.slice(5, n - 7)]):
        got = df.DeviceBatch.from_arrow(ctx, batch).filter("pred").to_arrow()
        want = oracle.filter_batch(batch, 5)
        assert got.schema.types == want.schema.types
        assert_rows_equal(rows_of(got), rows_of(want), ordered=True, what=f"filter n={n}")   # input order is preserved
    ctx.close()


def test_filter_rejects_non_boolean():
    import datafusion_upstream_b200 as df
    ctx = gpu_ctx()
    b = pa.record_batch([pa.array([1, 2, 3], pa.int64())], names=["i"])
    with pytest.raises(Exception, match="Boolean"):
        df.DeviceBatch.from_arrow(ctx, b).filter("i")
    ctx.close()


# ---- CoalesceBatchesExec at the PCIe boundary (dfgpu_coalescer_*) ----
def test_coalesce_batches_golden():
    """coalesce_batches.rs:298-316 (test_concat_batches): 10 batches x 8 rows, target 21 -> 24, 24, 24, 8 rows."""
    import datafusion_upstream_b200 as df
    ctx = gpu_ctx()
    schema = pa.schema([pa.field("c0", pa.uint32(), nullable=False)])
    b = pa.record_batch([pa.array([1, 2, 3, 4, 5, 6, 7, 8], pa.uint32())], schema=schema)
    op = df.GpuCoalesceBatchesExec(ctx, schema, 21)
    outs = [d.to_arrow() for d in op.execute([b] * 10)]
    assert [o.num_rows for o in outs] == [24, 24, 24, 8]
    assert all(o.column(0).to_pylist() == [1, 2, 3, 4, 5, 6, 7, 8] * (o.num_rows // 8) for o in outs)
    op.close()
    ctx.close()


def test_coalesce_batches_vs_concat():
    """Every column kind the arena packs, slices with odd offsets, pass-through of a large batch, empty batches dropped;
    the concatenation of the outputs equals the concatenation of the inputs, row for row and in order."""
    import datafusion_upstream_b200 as df
    rng = np.random.default_rng(17)
    ctx = gpu_ctx()
    n = 30_000
    whole = pa.record_batch([random_values(rng, n, "int64", 0.1), random_keys(rng, n, "utf8", 500, 0.1), random_values(rng, n, "decimal", 0.05),
                             pa.array(rng.random(n) < 0.5, pa.bool_(), mask=rng.random(n) < 0.1), random_keys(rng, n, "large_utf8", 90, 0.0),
                             random_values(rng, n, "int8", 0.0)], names=["i", "s", "d", "b", "ls", "t"])
    lens = [5, 0, 1000, 17, 4000, 3, 3, 8192, 1, 2500, 9000, 11]
    pieces, off = [], 0
    while off < n:
        for ln in lens:
            ln = min(ln, n - off)
            pieces.append(whole.slice(off, ln))
            off += ln
            if off >= n:
                break
    op = df.GpuCoalesceBatchesExec(ctx, whole.schema, 4096)
    outs = [d.to_arrow() for d in op.execute(pieces)]
    assert all(o.num_rows >= 4096 for o in outs[:-1]) and sum(o.num_rows for o in outs) == n
    assert_rows_equal(rows_of(outs), rows_of(whole), ordered=True, what="coalesce")
    op.close()
    with pytest.raises(Exception, match="not implemented|NotImplemented"):
        df.GpuCoalesceBatchesExec(ctx, pa.schema([("d", pa.dictionary(pa.int32(), pa.string()))]), 100)
    ctx.close()


# ---- ProjectionExec arithmetic over Decimal128 (dfgpu_batch_eval_decimal) ----
def test_decimal_arithmetic_golden_slt():
    """decimal.slt:205-362: c1+1, c1+c5, c1-1, c1-c5, c1*20, c1*c5 - result types and values."""
    import datafusion_upstream_b200 as df
    from test_oracle_golden import AGGS, DEC_ARITH, decimal_simple
    import decimal
    ctx = gpu_ctx()
    dev = df.DeviceBatch.from_arrow(ctx, decimal_simple())
    for name, nodes in DEC_ARITH.items():
        g = AGGS["decimal_slt"]["arithmetic"][name]
        out = dev.eval_decimal(nodes, name).to_arrow()
        assert out.num_columns == 3 and out.schema.names[2] == name
        assert out.column(2).type == pa.decimal128(*g["type"])
        assert sorted(out.column(2).to_pylist()) == sorted(decimal.Decimal(v) for v in g["values_rowsort"])
    ctx.close()


def test_decimal_arithmetic_vs_oracle_q1_projections():
    """l_extendedprice * (1 - l_discount) and that * (1 + l_tax) (TPC-H Q1's projections) on random Decimal128(15,2) columns
    with nulls, negative values and slices; bit-exact against the oracle, types Decimal128(38,4) / (38,6)."""
    import datafusion_upstream_b200 as df
    import decimal
    rng = np.random.default_rng(99)
    ctx = gpu_ctx()
    n = 50_001
    d152 = pa.decimal128(15, 2)
    def col(lo, hi, nulls):
        return pa.array([decimal.Decimal(int(x)).scaleb(-2) for x in rng.integers(lo, hi, n)], d152, mask=(rng.random(n) < nulls) if nulls else None)
    b = pa.record_batch([col(-10_000_000, 10_495_001, 0.1), col(0, 11, 0.05), col(0, 9, 0.0), pa.array(rng.integers(0, 9, n), pa.int32())],
                        names=["price", "disc", "tax", "other"])
    disc_price = [("col", 0), ("lit", 1, 20, 0), ("col", 1), ("sub", 1, 2), ("mul", 0, 3)]
    charge = disc_price + [("lit", 1, 20, 0), ("col", 2), ("add", 5, 6), ("mul", 4, 7)]
    for batch in (b, b.slice(13, 40_000)):
        dev = df.DeviceBatch.from_arrow(ctx, batch)
        for nodes, t in ((disc_price, pa.decimal128(38, 4)), (charge, pa.decimal128(38, 6))):
            got = dev.eval_decimal(nodes, "x").to_arrow()
            want = oracle.eval_decimal(batch, nodes)
            assert got.column(4).type == t == want.type
            assert got.column(4).to_pylist() == want.to_pylist()
            assert got.column(0).to_pylist() == batch.column(0).to_pylist()       # the input columns ride along untouched
    ctx.close()


def test_decimal_arithmetic_errors():
    import datafusion_upstream_b200 as df
    import decimal
    ctx = gpu_ctx()
    big = pa.record_batch([pa.array([decimal.Decimal(10) ** 37, None, decimal.Decimal(3)], pa.decimal128(38, 0)),
                           pa.array([decimal.Decimal("0.1")] * 3, pa.decimal128(38, 20)), pa.array([1, 2, 3], pa.int64())], names=["x", "s", "i"])
    dev = df.DeviceBatch.from_arrow(ctx, big)
    with pytest.raises(Exception, match="Overflow"):
        dev.eval_decimal([("col", 0), ("col", 0), ("mul", 0, 1)], "y")            # 10^74 leaves i128 (arrow: mul_checked)
    ok = dev.slice(1, 2).eval_decimal([("col", 0), ("col", 0), ("mul", 0, 1)], "y").to_arrow()
    assert ok.column(3).to_pylist() == [None, decimal.Decimal(9)]                 # NULL slots are not computed
    with pytest.raises(Exception, match="max scale"):
        dev.eval_decimal([("col", 1), ("col", 1), ("mul", 0, 1)], "y")
    with pytest.raises(Exception, match="not implemented|NotImplemented"):
        dev.eval_decimal([("col", 2), ("col", 0), ("add", 0, 1)], "y")
    with pytest.raises(Exception, match="post order|precede"):
        dev.eval_decimal([("col", 0), ("add", 0, 2), ("col", 0)], "y")
    ctx.close()
