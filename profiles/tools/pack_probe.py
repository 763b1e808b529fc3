"""Where does the time of dfgpu_agg_push_arrow_batches go?  100 M Int64 keys in 8192-row batches, pinned host memory."""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, pyarrow as pa
import datafusion_upstream_b200 as df
from datafusion_upstream_b200 import _native as N

n = 50_000_000
ctx = df.GpuContext(device=0)
k = np.random.default_rng(1).integers(0, 1 << 40, n).astype(np.int64)
b = pa.record_batch([pa.array(k)], names=["k"])
exprs = [df.AggregateExpr(df.AggFunc.Count, None, "c")]
slices = [b.slice(o, 8192) for o in range(0, n, 8192)]
for mode in ("one", "many", "many", "one"):
    op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, b.schema, ["k"], exprs)
    ex = op.export_batches(slices)
    ctx.synchronize()
    t0 = time.perf_counter()
    if mode == "many":
        op.push_exported(ex)
    else:
        L = N.lib()
        for a in ex[0]:
            N.check(L.dfgpu_agg_push_arrow(op.h, ctypes.addressof(a)))
    t1 = time.perf_counter()
    out = op.emit()
    t2 = time.perf_counter()
    m = op.metrics()
    print(mode, "push %.1f ms" % ((t1 - t0) * 1e3), "emit %.1f ms" % ((t2 - t1) * 1e3), "elapsed_compute %.1f ms" % (m["elapsed_compute_ns"] / 1e6), "groups", out.num_rows, flush=True)
    op.release_exported(ex)
    op.close()
