"""Two-GPU parity of the sharded paths (needs >= 2 GPUs; skipped otherwise): the fused operator
(`dfgpu_agg_exchange`: level-1 records partitioned by key hash, exchanged over NCCL, merged by their owner) and the
three-operator plan it replaces - Partial on each rank's row range
-> `dfgpu_partition_hash` -> `dfgpu_exchange_all_to_all` (NCCL) -> FinalPartitioned, against the oracle's Single
aggregation of the whole table.  Also the join exchange: both sides repartitioned by key owner, per-rank
Partitioned joins, union equals the oracle's join."""
import os
import socket
import sys

import numpy as np
import pyarrow as pa
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _table(seed, n):
    rng = np.random.default_rng(seed)
    k1 = pa.array(rng.integers(-50, 5000, n), pa.int64(), mask=rng.random(n) < 0.02)
    k2 = pa.array(rng.integers(0, 3, n).astype(np.int16), pa.int16())
    v = pa.array(rng.integers(-10**6, 10**6, n), pa.int64(), mask=rng.random(n) < 0.1)
    f = pa.array(rng.normal(size=n))
    return pa.record_batch([k1, k2, v, f], names=["k1", "k2", "v", "f"])


SPECS = [("count", None, "COUNT(*)", None), ("sum", "v", "SUM(v)", None), ("avg", "f", "AVG(f)", None), ("min", "v", "MIN(v)", None),
         ("max", "f", "MAX(f)", None)]


def _worker(rank, world, port, n_rows, out_q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import datafusion_upstream_b200 as df
        from datafusion_upstream_b200.dist import bootstrap_nccl_id, shard_range, sharded_aggregate
        uid = bootstrap_nccl_id(dist, df.GpuContext.nccl_unique_id)
        ctx = df.GpuContext(device=rank, rank=rank, world=world, nccl_unique_id=uid)
        full = _table(11, n_rows)
        lo, hi = shard_range(n_rows, rank, world)
        mine = full.slice(lo, hi - lo)
        exprs = [df.AggregateExpr(df.AggFunc[f.capitalize()], a, nm) for f, a, nm, _ in SPECS]
        out = sharded_aggregate(ctx, full.schema, ["k1", "k2"], exprs, [mine.slice(0, 1000), mine.slice(1000)], world)
        # the fused operator (dfgpu_agg_exchange): a nullable Int64 key with NULLs and -1 (side groups), a 3-group key
        # (register mode) and the two-column key above (not a 64-bit image: falls back to the three-operator plan)
        from datafusion_upstream_b200.dist import fused_sharded_aggregate
        fused = [fused_sharded_aggregate(ctx, full.schema, gb, exprs, [mine.slice(0, 1000), mine.slice(1000)], world)
                 for gb in (["k1"], ["k2"], ["k1", "k2"])]
        # join: repartition both sides by key owner, join locally
        left = pa.record_batch([pa.array(np.arange(3000, dtype=np.int64) % 1500), pa.array(np.arange(3000, dtype=np.int32))], names=["a", "x"])
        right = pa.record_batch([pa.array((np.arange(9000, dtype=np.int64) * 7) % 2000), pa.array(np.arange(9000, dtype=np.int32))], names=["b", "y"])
        llo, lhi = shard_range(left.num_rows, rank, world)
        rlo, rhi = shard_range(right.num_rows, rank, world)
        lp = df.GpuRepartitionExec(ctx, left.schema, ["a"], world)
        rpx = df.GpuRepartitionExec(ctx, right.schema, ["b"], world)
        lmine = lp.exchange(lp.partition_device(left.slice(llo, lhi - llo)))
        rmine = rpx.exchange(rpx.partition_device(right.slice(rlo, rhi - rlo)))
        j = df.GpuHashJoinExec(ctx, left.schema, right.schema, [("a", "b")], df.JoinType.Inner, df.PartitionMode.Partitioned)
        jout = j.execute([lmine] if lmine is not None else [], [rmine] if rmine is not None else [])
        jrows = sum(b.num_rows for b in jout)
        jsum = int(sum(int(np.sum(b.column(1).to_numpy().astype(np.int64) * 31 + b.column(3).to_numpy())) for b in jout))
        import io

        def ipc(b):
            sink = io.BytesIO()
            with pa.ipc.new_stream(sink, b.schema) as w:
                w.write_batch(b)
            return sink.getvalue()
        res = [None] * world
        dist.all_gather_object(res, (ipc(out), jrows, jsum, [ipc(f) for f in fused]))
        if rank == 0:
            out_q.put(res)
        ctx.synchronize()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(900)
@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_gpu_sharded_aggregate_and_join(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (run with gpurun --gpus {world}); the same exchange runs between virtual ranks in test_gpu_virtual_ranks.py")
    import torch.multiprocessing as mp
    import oracle
    from gpu_helpers import oracle_agg
    from util import assert_rows_equal, rows_of
    n_rows = 200_003
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_worker, args=(r, world, port, n_rows, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=500)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    rows = []
    for buf, _, _, _ in res:
        for b in pa.ipc.open_stream(buf).read_all().to_batches():
            rows += rows_of(b)
    want = oracle_agg("single", [_table(11, n_rows)], ["k1", "k2"], SPECS)
    assert_rows_equal(rows, rows_of(want), key_cols=[0, 1])
    for i, gb in enumerate((["k1"], ["k2"], ["k1", "k2"])):
        rows, per_rank = [], []
        for r in res:
            got = [x for b in pa.ipc.open_stream(r[3][i]).read_all().to_batches() for x in rows_of(b)]
            per_rank.append(len(got))
            rows += got
        want = oracle_agg("single", [_table(11, n_rows)], gb, SPECS)
        assert_rows_equal(rows, rows_of(want), key_cols=list(range(len(gb))), what=f"fused {gb}")
        if gb == ["k1"]:
            assert min(per_rank) > 0.3 * len(rows) / world   # ownership by hash range: both ranks hold a share
    # join: row count and a checksum over (x, y) equal the oracle's
    left = pa.table({"a": np.arange(3000, dtype=np.int64) % 1500, "x": np.arange(3000, dtype=np.int32)})
    right = pa.table({"b": (np.arange(9000, dtype=np.int64) * 7) % 2000, "y": np.arange(9000, dtype=np.int32)})
    oj = oracle.hash_join("inner", left.to_batches(), right.to_batches(), [("a", "b")])
    assert sum(r[1] for r in res) == oj.num_rows
    assert sum(r[2] for r in res) == int(np.sum(oj.column(1).to_numpy().astype(np.int64) * 31 + oj.column(3).to_numpy()))


def _big_keys(rank, n):
    """Zipf(1.0) ranks over a shared universe, spread over 64 bits: every rank sees the hot keys, the tail is mostly its own."""
    rng = np.random.default_rng(1000 + rank)
    r = np.exp(rng.random(n) * np.log(40_000_000)).astype(np.int64)
    with np.errstate(over="ignore"):
        return (r.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)).view(np.int64)


def _worker_big(rank, world, port, n_rows, s2max, out_q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    if s2max:
        os.environ["DFGPU_FA_S2MAX"] = s2max
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import datafusion_upstream_b200 as df
        from datafusion_upstream_b200.dist import bootstrap_nccl_id
        uid = bootstrap_nccl_id(dist, df.GpuContext.nccl_unique_id)
        ctx = df.GpuContext(device=rank, rank=rank, world=world, nccl_unique_id=uid)
        k = _big_keys(rank, n_rows)
        b = pa.record_batch([pa.array(k, pa.int64())], names=["k"])
        res = []
        count, count_sum = [("count", None, "c")], [("count", None, "c"), ("sum", "k", "s")]   # partition path (key pieces) / cache path (records)
        # count-only operators: the owner's merge reads the peers' key pieces over NVLink (pull; twice, the second time in two pushes
        # so that the send arena holds two pieces and has to grow), runs copied into the owner's window (push), NCCL send / recv
        for specs, model, pushes in ((count, "pull", 1), (count, "pull", 2), (count, "push", 1), (count_sum, "pull", 1), (count, "nccl", 1)):
            os.environ["DFGPU_EXCHANGE"] = model
            op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, b.schema, ["k"], [df.AggregateExpr(df.AggFunc[f.capitalize()], a, nm) for f, a, nm in specs])
            if pushes == 1:
                op.push(b)
            else:
                cut = (2 * n_rows) // 3 + 1000 * rank
                op.push(b.slice(0, cut))
                op.push(b.slice(cut))
            op.exchange()
            out = op.emit()
            op.close()
            gk, gc = out.column(0).to_numpy(), out.column(1).to_numpy()
            with np.errstate(over="ignore"):
                cks = int((gk.view(np.uint64) * np.uint64(0xBF58476D1CE4E5B9) * gc.astype(np.uint64)).sum(dtype=np.uint64))
            res.append((len(gk), int(gc.sum()), cks, len(gk) - len(np.unique(gk))))
        allr = [None] * world
        dist.all_gather_object(allr, res)
        if rank == 0:
            out_q.put(allr)
        ctx.synchronize()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(1500)
@pytest.mark.parametrize("s2max", [None, "256"], ids=["default", "tiny_level2_tables"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_gpu_count_star_20m_rows_per_rank(world, s2max):
    """20 M rows per rank, ~10 M distinct keys per rank from one universe: groups, sum of counts and a key-weighted checksum of the
    union against numpy on the same keys; no key may come out twice (tiny level-2 tables force the overflow rounds that once let a
    key through two rounds at N = 8)."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (run with gpurun --gpus {world})")
    import torch.multiprocessing as mp
    n_rows = 20_000_000
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_worker_big, args=(r, world, port, n_rows, s2max, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=1200)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    allk = np.concatenate([_big_keys(r, n_rows) for r in range(world)])
    uk, uc = np.unique(allk, return_counts=True)
    with np.errstate(over="ignore"):
        want_cks = int((uk.view(np.uint64) * np.uint64(0xBF58476D1CE4E5B9) * uc.astype(np.uint64)).sum(dtype=np.uint64))
    for variant in range(5):
        G = sum(r[variant][0] for r in res)
        rows = sum(r[variant][1] for r in res)
        cks = sum(r[variant][2] for r in res) % (1 << 64)
        dup = sum(r[variant][3] for r in res)
        assert (G, rows, cks, dup) == (len(uk), n_rows * world, want_cks, 0), (variant, G, len(uk), rows, dup)
