"""world_size-2 `gloo` tests (CPU) of the host logic of the sharded path: row-range sharding, the
ncclUniqueId bootstrap, and the Partial -> hash exchange -> FinalPartitioned protocol (ownership = hash % N,
every key merged on exactly one rank).  The device is stood in for by the oracle here; the same protocol on
real GPUs is `tests/test_gpu_dist.py` / `bench.py --gpus N`."""
import io
import os
import socket
import sys

import numpy as np
import pyarrow as pa
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _table(seed, n):
    rng = np.random.default_rng(seed)
    keys = pa.array(rng.integers(-50, 400, n), pa.int64(), mask=rng.random(n) < 0.03)
    v = pa.array(rng.integers(-10**6, 10**6, n), pa.int64(), mask=rng.random(n) < 0.1)
    f = pa.array(rng.normal(size=n))
    return pa.record_batch([keys, v, f], names=["k", "v", "f"])


SPECS = [("count", None, "COUNT(*)", None), ("sum", "v", "SUM(v)", None), ("avg", "f", "AVG(f)", None), ("min", "v", "MIN(v)", None),
         ("max", "f", "MAX(f)", None)]


def _ipc(batch):
    sink = io.BytesIO()
    with pa.ipc.new_stream(sink, batch.schema) as w:
        w.write_batch(batch)
    return sink.getvalue()


def _unipc(buf):
    return pa.ipc.open_stream(buf).read_all().combine_chunks().to_batches()


def _worker(rank, world, port, n_rows, out_q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from gpu_helpers import oracle_agg
        from datafusion_upstream_b200.dist import bootstrap_nccl_id, shard_range
        # 1. bootstrap: every rank ends up with rank 0's id
        uid = bootstrap_nccl_id(dist, lambda: bytes(range(128)))
        assert uid == bytes(range(128))
        # 2. range sharding
        full = _table(7, n_rows)
        lo, hi = shard_range(n_rows, rank, world)
        mine = full.slice(lo, hi - lo)
        # 3. Partial on the shard, partition by owner, exchange, FinalPartitioned on the owned keys
        st = oracle_agg("partial", [mine], ["k"], SPECS)
        parts = oracle.partition(st, [0], world)          # hash % world, empty partitions are None
        send = [None if p is None else _ipc(p) for p in parts]
        recv = [None] * world
        gathered = [None] * world
        dist.all_gather_object(gathered, send)            # gathered[src][dst]
        recv = [gathered[src][rank] for src in range(world)]
        batches = [b for buf in recv if buf is not None for b in _unipc(buf)]
        fin_specs = [(f, None, nm, None) for f, _, nm, _ in SPECS]
        if batches:
            fin = oracle_agg("final_partitioned", batches, [0], fin_specs, schema=st.schema)
            owned = set(fin.column(0).to_pylist())
        else:
            fin, owned = None, set()
        res = [None] * world
        dist.all_gather_object(res, (None if fin is None else _ipc(fin), sorted(owned, key=lambda x: (x is None, x))))
        if rank == 0:
            out_q.put(res)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_sharded_partial_exchange_final_world2():
    import torch.multiprocessing as mp
    from gpu_helpers import oracle_agg
    from util import assert_rows_equal, rows_of
    world, n_rows = 2, 20_001
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_rows, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=200)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # every key is owned by exactly one rank
    owned = [set(o) for _, o in res]
    assert not (owned[0] & owned[1])
    rows = []
    for buf, _ in res:
        if buf is not None:
            for b in _unipc(buf):
                rows += rows_of(b)
    want = oracle_agg("single", [_table(7, n_rows)], ["k"], SPECS)
    assert_rows_equal(rows, rows_of(want), key_cols=[0])


def test_shard_range_covers_and_balances():
    from datafusion_upstream_b200.dist import shard_range
    for n in (0, 1, 7, 100, 600_037_902):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)
