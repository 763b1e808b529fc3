"""-m gpu test of dfgpu_partition_i64_noalloc, the allocation-free variant of dfgpu_partition_hash's
partitioning step (include/dfgpu.h).  It has no counterpart in the reference; the checker is the existing entry
point on the same rows.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu


def _existing(ctx, keys: pa.Array, nparts: int):
    """dfgpu_partition_hash on (k, rowid): the rowid columns of the parts, concatenated, are the permutation."""
    import datafusion_upstream_b200 as df
    n = len(keys)
    b = pa.record_batch([keys, pa.array(np.arange(n), pa.int64())], names=["k", "rowid"])
    parts = df.GpuRepartitionExec(ctx, b.schema, [0], nparts).partition(b)
    sizes = [0 if p is None else p.num_rows for p in parts]
    perm = np.concatenate([p.column(1).to_numpy() for p in parts if p is not None]) if n else np.zeros(0, np.int64)
    return perm, np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)


def _noalloc(keys_np: np.ndarray, valid_np, nparts: int, check_no_alloc: bool):
    import torch
    from datafusion_upstream_b200 import _native as N
    L = N.lib()
    n = len(keys_np)
    dev = torch.device("cuda:0")
    k = torch.from_numpy(keys_np).to(dev)
    v = None if valid_np is None else torch.from_numpy(np.packbits(valid_np, bitorder="little")).to(dev)
    # outputs pre-filled with garbage: everything must be written by the call
    perm = torch.full((max(n, 1),), -7, dtype=torch.int64, device=dev)
    offs = torch.full((nparts + 1,), -7, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    st = L.dfgpu_partition_i64_noalloc(C.c_void_p(torch.cuda.current_stream().cuda_stream), C.c_void_p(k.data_ptr()),
                                       C.c_void_p(v.data_ptr()) if v is not None else None, n, nparts,
                                       C.c_void_p(perm.data_ptr()), C.c_void_p(offs.data_ptr()))
    N.check(st)
    torch.cuda.synchronize()
    if check_no_alloc:   # (the first call of a process may page in the module's kernels)
        assert torch.cuda.mem_get_info(0)[0] == free0, "the entry point must not allocate device memory"
    return perm[:n].cpu().numpy(), offs.cpu().numpy()


def test_noalloc_partition_matches_partition_hash_twice_in_a_row():
    import torch
    torch.cuda.set_device(0)
    from gpu_helpers import gpu_ctx
    ctx = gpu_ctx()
    try:
        rng = np.random.default_rng(2024)
        # two calls in a row that share the static scratch: different sizes, partition counts and null patterns, the
        # second SMALLER than the first so stale histogram columns / partition ids of call one lie beyond what it rewrites
        cases = [(700_001, 61, 0.1), (200_003, 7, 0.0), (4097, 256, 0.5), (1 << 20, 64, 0.02), (5, 3, 0.0)]
        for it, (n, nparts, nullf) in enumerate(cases):
            keys = rng.integers(-2**62, 2**62, n, dtype=np.int64)
            keys[rng.random(n) < 0.3] = 42                       # a heavy hitter: one long run inside a partition
            valid = None if nullf == 0.0 else rng.random(n) >= nullf
            arr = pa.array(keys, pa.int64(), mask=None if valid is None else ~valid)
            want_perm, want_offs = _existing(ctx, arr, nparts)
            got_perm, got_offs = _noalloc(keys, valid, nparts, check_no_alloc=it > 0)
            assert (got_offs == want_offs).all(), (it, got_offs[:8], want_offs[:8])   # integer work: bit-exact
            assert (got_perm == want_perm).all(), it
            assert sorted(got_perm.tolist()) == list(range(n)) if n < 10_000 else len(np.unique(got_perm)) == n
        # empty input: offsets all zero, nothing else touched
        _, offs = _noalloc(np.zeros(0, np.int64), None, 5, check_no_alloc=True)
        assert (offs == 0).all()
    finally:
        ctx.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    from datafusion_upstream_b200 import _native as N
    L = N.lib()
    dev = torch.device("cuda:0")
    buf = torch.zeros(16, dtype=torch.int64, device=dev)
    p = C.c_void_p(buf.data_ptr())
    for n, nparts in [((1 << 20) + 1, 4), (16, 257), (16, 0), (-1, 4)]:
        st = L.dfgpu_partition_i64_noalloc(None, p, None, n, nparts, p, p)
        assert st == 5, (n, nparts, st)   # DFGPU_NOT_IMPLEMENTED
        msg = L.dfgpu_last_error().decode()
        assert "exceeds the static scratch" in msg and "1048576" in msg and "256" in msg, msg
