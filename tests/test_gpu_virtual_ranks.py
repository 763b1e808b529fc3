"""The fused sharded aggregation (Partial -> record exchange -> FinalPartitioned in one operator per rank, dfgpu_agg_exchange)
between VIRTUAL ranks: N operators of one process on one GPU, exchanged by dfgpu_agg_exchange_virtual (SURVEY 8e "test without
8 GPUs"; the reference tests its partitioned plans the same way, joins/hash_join.rs:1505-1550).  Everything but the transport is
the code the real ranks run: partition order, offsets, piece tables, k_fa_l2 over FaSegs pieces (incl. the 1024-thread variant
for more than two ranks), k_hp_merge over key and record pieces, hot keys and side groups changing owner."""
import numpy as np
import pyarrow as pa
import pytest

from gpu_helpers import FUNCS, gpu_ctx, oracle_agg, random_keys, random_values
from util import assert_tables_equal

pytestmark = pytest.mark.gpu

COUNT = [("count", None, "COUNT(*)", None)]
MIXED = [("count", None, "c", None), ("sum", "v", "s", None), ("avg", "d", "a", None), ("min", "v", "mn", None), ("max", "f", "mx", None),
         ("count", "v", "cv", None)]


def run_virtual(ctx, shards, group_by, specs, kernels=None):
    """One Single-mode operator per shard, virtual exchange, emit each: returns the per-rank outputs."""
    import datafusion_upstream_b200 as df
    exprs = [df.AggregateExpr(df.AggFunc(FUNCS[f]), a, n, flt) for f, a, n, flt in specs]
    ops = [df.GpuAggregateExec(ctx, df.AggregateMode.Single, shards[0][0].schema, group_by, exprs) for _ in shards]
    for op, batches in zip(ops, shards):
        for b in batches:
            op.push(b)
    ctx.set_profiling(True)
    try:
        df.GpuAggregateExec.exchange_virtual(ops)
        outs = [op.emit() for op in ops]
        names = {r["name"] for r in ctx.profile()}
    finally:
        ctx.set_profiling(False)
    for op in ops:
        op.close()
    if kernels is not None:
        kernels |= names
    return outs


def check_union(outs, shards, group_by, specs, nkeys):
    want = oracle_agg("single", [b for s in shards for b in s], group_by, specs)
    got = pa.Table.from_batches([o for o in outs if o.num_rows] or [outs[0]])
    assert_tables_equal(got, want, key_cols=list(range(nkeys)))   # a key emitted by two ranks would show up as an extra row


def zipf_keys(rng, n, universe):
    r = np.exp(rng.random(n) * np.log(universe)).astype(np.int64)
    with np.errstate(over="ignore"):
        return (r.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)).view(np.int64)


@pytest.mark.parametrize("hook", [{}, {"DFGPU_FA_S2MAX": "64", "DFGPU_FA_S1": "32"}], ids=["default", "tiny_tables"])
@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_general_accumulators_over_record_pieces(monkeypatch, world, hook):
    """SUM / AVG(Decimal128) / MIN / MAX(Float64) / COUNT(x): partial-state records, level 2 over one piece per source rank."""
    for k, v in hook.items():
        monkeypatch.setenv(k, v)
    ctx = gpu_ctx()
    try:
        rng = np.random.default_rng(world)
        shards = []
        for r in range(world):
            n = [30_000, 0, 7, 12_345, 20_000, 1, 9_999, 15_000][r]
            if n == 0:
                n = 1 if world == 2 else 0
            bs = []
            for m in ([n] if n < 10_000 else [n // 2, n - n // 2]):
                bs.append(pa.record_batch([random_keys(rng, m, "int64_small", 4000, 0.03), random_values(rng, m, "int64", 0.1),
                                           random_values(rng, m, "decimal", 0.1), random_values(rng, m, "float64", 0.05)], names=["k", "v", "d", "f"]))
            shards.append(bs)
        names = set()
        outs = run_virtual(ctx, shards, ["k"], MIXED, names)
        assert "k_fa_l2" in names and "k_hp_merge" not in names, names
        check_union(outs, shards, ["k"], MIXED, 1)
        # ownership is by hash range: no key comes out of two ranks
        allk = np.concatenate([o.column(0).to_numpy(zero_copy_only=False) for o in outs if o.num_rows])
        assert len(allk) == len(set(allk.tolist()))
    finally:
        ctx.close()


@pytest.mark.parametrize("model", ["pull", "pull_remote_loop", "push"])
@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("modes", ["all_partition", "all_cache", "mixed"])
def test_count_star_key_and_record_pieces(monkeypatch, world, modes, model):
    """COUNT(*): ranks on the partition path send bare keys + hot keys, ranks on the cache path send records; the owner merges both
    kinds of pieces (k_hp_merge).  NULL and all-ones keys (side groups) end up on rank 0.  model = pull: the owner's merge reads the
    other ranks' streams where they lie (the default between real ranks: NVLink loads); push: runs are copied into the owner's
    window first (the fallback, and the path of operators with other accumulators)."""
    import datafusion_upstream_b200 as df
    monkeypatch.setenv("DFGPU_EXCHANGE", "push" if model == "push" else "pull")
    if model == "pull_remote_loop":   # the merge loop that real ranks use for pieces in another GPU's memory (4 keys per lane in flight, no prefetch)
        monkeypatch.setenv("DFGPU_FA_REMOTE_LOOP", "1")
    ctx = gpu_ctx()
    try:
        rng = np.random.default_rng(100 + world)
        exprs = [df.AggregateExpr(df.AggFunc.Count, None, "COUNT(*)")]
        shards, ops = [], []
        for r in range(world):
            n = 60_000 + 5_000 * r
            raw = zipf_keys(rng, n, 80_000)
            raw[rng.random(n) < 0.01] = -1
            b = pa.record_batch([pa.array(raw, pa.int64(), mask=rng.random(n) < 0.02)], names=["k"])
            part = modes == "all_partition" or (modes == "mixed" and r % 2 == 0)
            monkeypatch.setenv("DFGPU_FA_PART", "1" if part else "0")    # read when the operator is created
            op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, b.schema, ["k"], exprs)
            halves = [b.slice(0, n // 3), b.slice(n // 3)] if r % 3 == 1 else [b]      # two key pieces on some ranks
            for h in halves:
                op.push(h)
            shards.append(halves)
            ops.append(op)
        ctx.set_profiling(True)
        df.GpuAggregateExec.exchange_virtual(ops)
        outs = [op.emit() for op in ops]
        names = {x["name"] for x in ctx.profile()}
        ctx.set_profiling(False)
        for op in ops:
            op.close()
        assert "k_hp_merge" in names, names
        assert ("k_hp_part1" in names) == (modes != "all_cache")
        assert ("stage:record_exchange_peer_push" in names) == (model == "push"), names
        check_union(outs, shards, ["k"], COUNT, 1)
        null_rows = [i for i, o in enumerate(outs) if o.num_rows and o.column(0).null_count]
        assert null_rows == [0]
    finally:
        ctx.close()


@pytest.mark.parametrize("groups", [4, 200, 300], ids=["q1_like_4_groups", "200_groups", "300_groups_regular_path"])
@pytest.mark.parametrize("world", [2, 5, 8])
def test_small_states_ride_in_the_meta_blocks(world, groups):
    """Ranks whose partial states compact to <= 256 records put them into their meta block; when every rank did, the all-gather
    was the exchange (no window, push, barrier or plan).  One rank is empty, one holds NULL keys (side group -> rank 0); with 300
    groups per rank the regular path runs instead, after the same compaction."""
    ctx = gpu_ctx()
    try:
        rng = np.random.default_rng(world * 1000 + groups)
        shards = []
        for r in range(world):
            n = 0 if r == 1 else 5_000 + 100 * r
            k = rng.integers(0, groups, n).astype(np.int64) * 7919 - 3
            b = pa.record_batch([pa.array(k, pa.int64(), mask=(rng.random(n) < 0.01) if r == 0 else None), random_values(rng, n, "int64", 0.1),
                                 random_values(rng, n, "decimal", 0.1), random_values(rng, n, "float64", 0.05)], names=["k", "v", "d", "f"])
            shards.append([b])
        names = set()
        outs = run_virtual(ctx, shards, ["k"], MIXED, names)
        assert ("stage:record_exchange_small" in names) == (groups <= 200), names
        check_union(outs, shards, ["k"], MIXED, 1)
        allk = [x for o in outs if o.num_rows for x in o.column(0).to_pylist()]
        assert len(allk) == len(set(allk))
        assert all(o.column(0).null_count == 0 for o in outs[1:] if o.num_rows)   # the NULL group comes out of rank 0
    finally:
        ctx.close()


def test_more_key_pieces_than_streams_fold_into_records(monkeypatch):
    """A rank that pushed more than two batches through the partition path folds them into records before the exchange."""
    monkeypatch.setenv("DFGPU_FA_PART", "1")
    ctx = gpu_ctx(coalesce_rows=-1)
    try:
        rng = np.random.default_rng(77)
        shards = [[pa.record_batch([pa.array(zipf_keys(rng, 20_000, 30_000), pa.int64())], names=["k"]) for _ in range(5 if r == 0 else 1)] for r in range(3)]
        outs = run_virtual(ctx, shards, ["k"], COUNT)
        check_union(outs, shards, ["k"], COUNT, 1)
    finally:
        ctx.close()


def test_exchange_twice_on_the_same_operators(monkeypatch):
    """push -> exchange -> emit, then again on the same operators: the exchange state is reset by emit."""
    import datafusion_upstream_b200 as df
    monkeypatch.setenv("DFGPU_FA_PART", "1")
    ctx = gpu_ctx()
    try:
        rng = np.random.default_rng(5)
        exprs = [df.AggregateExpr(df.AggFunc.Count, None, "COUNT(*)")]
        schema = pa.schema([("k", pa.int64())])
        ops = [df.GpuAggregateExec(ctx, df.AggregateMode.Single, schema, ["k"], exprs) for _ in range(4)]
        for rnd in range(2):
            shards = [[pa.record_batch([pa.array(zipf_keys(rng, 40_000, 50_000), pa.int64())], schema=schema)] for _ in ops]
            for op, s in zip(ops, shards):
                op.push(s[0])
            df.GpuAggregateExec.exchange_virtual(ops)
            check_union([op.emit() for op in ops], shards, ["k"], COUNT, 1)
        for op in ops:
            op.close()
    finally:
        ctx.close()
