#!/usr/bin/env python3
"""Headline benchmark for the hash-table operator path (BASELINE.json).

A "step" is one pass of the hot path over one batch of synthetic input.  Default workload at N=1
is configs[1]: ClickBench-shaped `SELECT UserID, COUNT(*) ... GROUP BY UserID` over 100 M synthetic
Int64 keys (~17 M groups) on one B200.  Other workloads (`--workload q1|join`) are the Q1-/Q3-shaped
configurations and are reported the same way.

    python bench.py --gpus N --steps K --warmup W [--workload c2|q1|join] [--impl reference]

Prints ONE JSON line (rank 0).  `value` = whole-job input rows/s with inputs resident in HBM;
`e2e` = the same metric through the C-ABI with HOST (pinned) Arrow buffers, H2D/D2H inside the
timed region; `roofline` = dominant kernel, CUDA-event timed live on the launching stream;
`cpu_baseline` = the CPU oracle (port of the reference operators) on the box's host cores.
"""
import argparse
import ctypes
import json
import math
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import pyarrow as pa  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# synthetic data (deterministic, generated on the device with torch; seed 42 + rank)
# ------------------------------------------------------------------------------------------------
def _mix64_torch(x):
    import torch
    m = (1 << 63) - 1

    def lsr(v, k):  # logical shift right on int64
        return (v >> k) & ((1 << (64 - k)) - 1)
    x = x ^ lsr(x, 30)
    x = x * torch.tensor(-4658895280553007687, dtype=torch.int64, device=x.device)   # 0xbf58476d1ce4e5b9
    x = x ^ lsr(x, 27)
    x = x * torch.tensor(-7723592293110705685, dtype=torch.int64, device=x.device)   # 0x94d049bb133111eb
    x = x ^ lsr(x, 31)
    _ = m
    return x


def gen_c2_keys(n, seed, device, universe=85_000_000):
    """Zipf(s~1) ranks (log-uniform inverse CDF) over `universe` ids, mapped through a 64-bit bijective
    mixer so keys are sparse 64-bit values like real UserIDs."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    out = torch.empty(n, dtype=torch.int64, device=device)
    step = 1 << 24
    for o in range(0, n, step):
        m = min(step, n - o)
        u = torch.rand(m, generator=g, device=device, dtype=torch.float64)
        rank = torch.exp(u * math.log(universe)).to(torch.int64)
        out[o:o + m] = _mix64_torch(rank)
    return out


def gen_c5(n, seed, device, phrases=6_000_000, regions=9_000):
    """ClickBench-shaped (RegionID Int32, SearchPhrase Utf8) key columns: RegionID Zipf over ~9 K ids;
    SearchPhrase empty for 70 % of the rows, otherwise Zipf over ~6 M distinct phrases whose lengths are geometric in
    1..80 bytes (mean ~20).  Returns (region i32[n], offsets i32[n+1], data u8[total])."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    # phrase table: phrase p (p >= 1) is `len_p` bytes derived from a counter-based mixer, so equal ids give equal bytes
    plen = torch.empty(phrases + 1, dtype=torch.int64, device=device)
    u = torch.rand(phrases + 1, generator=g, device=device, dtype=torch.float64)
    plen[:] = torch.clamp((torch.log1p(-u) / math.log(1 - 1 / 20.0)).to(torch.int64) + 1, 1, 80)
    plen[0] = 0                                              # id 0 = the empty phrase
    region = torch.empty(n, dtype=torch.int32, device=device)
    pid = torch.empty(n, dtype=torch.int64, device=device)
    step = 1 << 24
    for o in range(0, n, step):
        m = min(step, n - o)
        u = torch.rand(m, generator=g, device=device, dtype=torch.float64)
        region[o:o + m] = (torch.exp(u * math.log(regions)).to(torch.int64) * 7 + 1).to(torch.int32)
        u = torch.rand(m, generator=g, device=device, dtype=torch.float64)
        e = torch.rand(m, generator=g, device=device) < 0.7
        r = torch.exp(u * math.log(phrases)).to(torch.int64).clamp_(1, phrases)
        pid[o:o + m] = torch.where(e, torch.zeros_like(r), r)
    lens = plen[pid]
    offsets = torch.zeros(n + 1, dtype=torch.int64, device=device)
    torch.cumsum(lens, 0, out=offsets[1:])
    total = int(offsets[-1])
    assert total < 2**31 - 1, "SearchPhrase bytes exceed i32 offsets; lower --rows"
    data = torch.empty(total, dtype=torch.uint8, device=device)
    bstep = 1 << 27
    for o in range(0, total, bstep):                          # bytes [o, o + m): which row, which byte of its phrase
        m = min(bstep, total - o)
        pos = torch.arange(o, o + m, device=device, dtype=torch.int64)
        row = torch.searchsorted(offsets, pos, right=True) - 1
        k = pos - offsets[row]
        h = _mix64_torch(pid[row] * 128 + k)
        data[o:o + m] = (97 + (h & 0xF)).to(torch.uint8)     # 'a'..'p'
    return region, offsets.to(torch.int32), data


def gen_q1(n, seed, device):
    """lineitem-shaped columns: 2 key columns with the TPC-H joint distribution, 5 Decimal128 columns."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    pick = torch.multinomial(torch.tensor([0.25, 0.006, 0.493, 0.251], device=device), n, replacement=True, generator=g)
    flag = torch.tensor([65, 78, 78, 82], dtype=torch.uint8, device=device)[pick]      # A N N R
    status = torch.tensor([70, 70, 79, 70], dtype=torch.uint8, device=device)[pick]    # F F O F

    def dec(lo, hi):  # int64 -> sign-extended 128-bit little endian pairs
        v = torch.randint(lo, hi, (n,), generator=g, device=device, dtype=torch.int64)
        return torch.stack([v, v >> 63], dim=1).contiguous()
    cols = {"qty": dec(100, 5001), "price": dec(90000, 10495001), "disc": dec(0, 11),
            "disc_price": dec(10**9, 10**11), "charge": dec(10**11, 10**13)}
    return flag, status, cols


class Clocks:
    """nvidia-smi clock / throttle sampling DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, dev, mode="nvml"):
        self.dev, self.p, self.h, self.samples, self.mode = dev, None, None, [], mode
        if mode == "nvml":
            try:
                import pynvml
                pynvml.nvmlInit()
                self.nv = pynvml
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                idx = int(vis.split(",")[dev]) if vis and vis.split(",")[dev].isdigit() else dev
                self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
                self.mx = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            except Exception:
                self.h = None

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM),
                                     nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)))
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self.mode == "off":
            return self
        if self.h is not None:
            import threading
            self._stop = threading.Event()
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
            return self
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(self.dev)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None
        return self

    def __exit__(self, *a):
        self.result = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.mode != "off" and self.h is not None:
            self._stop.set()
            self._t.join()
            nv = self.nv
            if not self.samples:
                self._loop_once = self.samples.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM),
                                                       nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)))
            bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
            allr = 0
            for _, r in self.samples:
                allr |= r
            self.result = {"sm_mhz": statistics.median([c for c, _ in self.samples]), "sm_max_mhz": self.mx,
                           "reasons": sorted(k for k, b in bits.items() if allr & b), "samples": len(self.samples), "source": "nvml"}
            return
        if not self.p:
            return
        time.sleep(0.15)
        self.p.terminate()
        try:
            out = self.p.communicate(timeout=5)[0]
        except Exception:
            return
        sm, mx, reasons = [], [], set()
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            self.result = {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# environment: one process per GPU; torch.distributed (gloo) only bootstraps the ncclUniqueId and provides barriers
# ------------------------------------------------------------------------------------------------
class Env:
    pass


class native_stdout_to_stderr:
    """NCCL announces its version on the process' stdout (fd 1) when the communicator is created; the contract is ONE
    JSON line on stdout, so fd 1 points at stderr while the context (and with it the communicator) is set up."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *a):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


def make_env(args):
    import torch
    import torch.distributed as dist
    import datafusion_upstream_b200 as df
    e = Env()
    e.torch, e.dist, e.df = torch, dist, df
    e.world = int(os.environ.get("WORLD_SIZE", "1"))
    e.rank = int(os.environ.get("RANK", "0"))
    e.local = int(os.environ.get("LOCAL_RANK", "0"))
    if e.world > 1:
        dist.init_process_group(backend="gloo")     # bootstrap / barrier only; the data path uses NCCL inside libdfgpu
    torch.cuda.set_device(e.local)
    e.device = torch.device("cuda", e.local)
    uid = None
    if e.world > 1:
        box = [df.GpuContext.nccl_unique_id() if e.rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]
    with native_stdout_to_stderr():
        e.ctx = df.GpuContext(device=e.local, rank=e.rank, world=e.world, nccl_unique_id=uid, coalesce_rows=args.coalesce_rows,
                              chunk_rows=args.chunk_rows)
    e.stream = torch.cuda.ExternalStream(e.ctx.stream, device=e.device)
    e.clocks_mode = args.clocks
    return e


def barrier(env):
    env.ctx.synchronize()
    env.torch.cuda.synchronize()
    if env.world > 1:
        env.dist.barrier()


def allreduce_i64(env, vals, op="sum"):
    """Exact all-reduce of a list of Python ints (each < 2**62 in magnitude) over the bootstrap group."""
    if env.world == 1:
        return [int(v) for v in vals]
    t = env.torch.tensor([int(v) for v in vals], dtype=env.torch.int64)
    env.dist.all_reduce(t, op=env.dist.ReduceOp.MAX if op == "max" else env.dist.ReduceOp.SUM)
    return [int(v) for v in t]


def allreduce_u64_wrapping(env, v):
    """Sum of 64-bit checksums modulo 2**64 (split in 32-bit halves so no rank-side overflow is involved)."""
    lo, hi = allreduce_i64(env, [v & 0xFFFFFFFF, (v >> 32) & 0xFFFFFFFF])
    return (lo + (hi << 32)) & MASK64


def timed_steps(env, steps, warmup, step_fn):
    """W untimed warm-up steps, then exactly K steps between barriers; device time from CUDA events on the context's stream
    (the stream every kernel of the library is launched on), max over ranks."""
    torch, ctx = env.torch, env.ctx
    out = None
    for _ in range(warmup):
        out = step_fn()
        out = None
    barrier(env)
    ctx.set_profiling(True)
    l0 = ctx.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clk = Clocks(env.local, env.clocks_mode)
    host_ms = []
    with clk:
        barrier(env)
        e0.record(env.stream)
        for _ in range(steps):
            t0 = time.perf_counter()
            out = None
            out = step_fn()
            host_ms.append(round((time.perf_counter() - t0) * 1e3, 2))
        e1.record(env.stream)
        barrier(env)
    ms = e0.elapsed_time(e1)
    launches = ctx.kernel_launches - l0
    prof = ctx.profile()
    ctx.set_profiling(False)
    if env.world > 1:
        t = torch.tensor([ms], dtype=torch.float64)
        env.dist.all_reduce(t, op=env.dist.ReduceOp.MAX)
        ms = float(t[0])
    return {"ms": ms, "ms_per_step": ms / steps, "host_ms": host_ms, "launches": launches, "prof": prof, "clocks": clk.result, "out": out}


def timed_e2e(env, steps, step_fn, h2d):
    """The same step through the public operators with pinned HOST Arrow buffers in and host Arrow results out: H2D and D2H
    inside the timed region (wall clock between barriers; max over ranks)."""
    torch = env.torch
    n = max(1, min(steps, 5))
    res = step_fn()
    res = None
    barrier(env)
    t0 = time.perf_counter()
    for _ in range(n):
        res = None            # the consumer drops the previous result: its pinned buffers return to the pool
        res = step_fn()
    barrier(env)
    dt = (time.perf_counter() - t0) / n
    d2h = int(res.nbytes)
    if env.world > 1:
        t = torch.tensor([dt], dtype=torch.float64)
        env.dist.all_reduce(t, op=env.dist.ReduceOp.MAX)
        dt = float(t[0])
        h2d, d2h = allreduce_i64(env, [h2d, d2h])
    return dt, h2d, d2h


def csrc_sha():
    """Fingerprint of the CUDA sources: profiles/traffic.json records the one its ncu capture was taken with."""
    import hashlib
    d = os.path.join(ROOT, "datafusion_upstream_b200", "csrc")
    h = hashlib.sha256()
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".h", ".inc")):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def roofline_of(env, r, steps, op_algo_bytes, workload):
    """`achieved` = the OPERATOR's algorithmic bytes (each consumed input column once + each output column once; scratch
    traffic is overhead and not counted) over the dominant kernel's device time per step; `step` = the same bytes over the whole
    step.  `kernel_stream_GBps` keeps the dominant kernel's own streamed bytes (what it reads + writes by design, scratch included)
    over its time - a diagnostic, not the roofline figure.  `traffic` = DRAM bytes per step from the last ncu capture of these
    sources (profiles/traffic.json), null when the sources changed since."""
    peak, peak_src = peaks()
    prof = r["prof"] or []
    stages = {p["name"][6:]: round(p["ms"] / steps, 4) for p in prof if p["name"].startswith("stage:")}
    kern = [p for p in prof if not p["name"].startswith("stage:")]
    if not kern:
        return None
    top = max(kern, key=lambda p: p["ms"])
    top_ms_step = top["ms"] / steps
    achieved = op_algo_bytes / (top_ms_step * 1e-3) / 1e9
    step_ach = op_algo_bytes / (r["ms_per_step"] * 1e-3) / 1e9
    traffic, traffic_all, note = None, None, "no capture for this workload"
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        tj = json.load(open(tp))
        ent = tj.get(workload)
        if ent and tj.get("csrc_sha") == csrc_sha() and env.world == 1:
            traffic = ent["kernels"].get(top["name"])
            traffic_all = {"per_kernel": ent["kernels"], "sum": sum(ent["kernels"].values()), "capture": ent.get("capture")}
            note = None
        elif ent:
            note = "stale: CUDA sources changed since the ncu capture (or N > 1)"
    roof = {"bound": "hbm", "kernel": top["name"], "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
            "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_step": int(op_algo_bytes),
            "launches_per_step": top["launches"] / steps, "ms_per_launch": round(top["ms"] / top["launches"], 4),
            "kernel_ms_per_step": round(top_ms_step, 4), "kernel_share_of_step": round(top["ms"] / r["ms"], 3),
            "step": {"achieved": round(step_ach, 1), "frac": round(step_ach / peak, 4), "frac_of_nominal_8TBps": round(step_ach / 8000.0, 4)},
            "kernel_stream_GBps": round(top["algo_bytes"] / top["launches"] / (top["ms"] / top["launches"] * 1e-3) / 1e9, 1) if top["algo_bytes"] else None,
            "kernels": {p["name"]: round(p["ms"] / steps, 4) for p in kern}, "stages": stages}
    if traffic_all:
        roof["traffic_step"] = traffic_all
    if note:
        roof["traffic_note"] = note
    return roof


def pinned_batch(ctx, schema, host_cols, n, torch):
    """Arrow RecordBatch whose buffers live in pinned host memory (PCIe-speed H2D for the e2e leg)."""
    arrays, keep, nbytes = [], [], 0
    for name, typ, t in host_cols:
        bufs = [None]
        for part in (t if isinstance(t, tuple) else (t,)):      # var-width columns: (offsets, data)
            src = part.contiguous().cpu().numpy().reshape(-1).view(np.uint8)
            ptr = ctx.host_alloc(max(src.nbytes, 1))
            ctypes.memmove(ptr, src.ctypes.data, src.nbytes)
            bufs.append(pa.foreign_buffer(ptr, src.nbytes))
            keep.append(ptr)
            nbytes += src.nbytes
        arrays.append(pa.Array.from_buffers(typ, n, bufs))
    return pa.RecordBatch.from_arrays(arrays, schema=schema), keep, nbytes


MASK64 = (1 << 64) - 1


def mix64_np(x):
    x = x.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(30); x *= np.uint64(0xbf58476d1ce4e5b9)
        x ^= x >> np.uint64(27); x *= np.uint64(0x94d049bb133111eb)
        x ^= x >> np.uint64(31)
    return x


def torch_sum_u64(t):
    """Sum of an int64 tensor modulo 2**64 without relying on overflow behaviour: two 32-bit halves."""
    lo = int((t & 0xFFFFFFFF).sum())
    hi = int(((t >> 32) & 0xFFFFFFFF).sum())
    return (lo + (hi << 32)) & MASK64


# ------------------------------------------------------------------------------------------------
# static description of a workload: both arms print the SAME config for the same command line
# ------------------------------------------------------------------------------------------------
def workload_config(args, world):
    wl, n = args.workload, args.rows
    if wl == "c2":
        name = f"ClickBench-shaped GROUP BY UserID COUNT(*), {n} Int64 rows per GPU, Zipf(1.0) over 85M ids (BASELINE configs[1])"
    elif wl == "q1":
        keys = "2xUtf8 keys (1 byte each; configs[0] shape, 90 B/row)" if args.q1_keys == "utf8" else "2xUInt8 keys (82 B/row variant of configs[0])"
        name = f"TPC-H Q1-shaped aggregation, {n} rows per GPU, 4 groups, {keys}, Decimal128 SUMx4 AVGx3 COUNT(*)"
        if args.q1_project:
            name += " + projections disc_price / charge evaluated on the device (66 B/row in)"
    elif wl in ("c5", "c5s"):
        g = "GROUP BY RegionID, SearchPhrase" if wl == "c5" else "GROUP BY SearchPhrase"
        name = (f"ClickBench-shaped {g} COUNT(*), {n} rows per GPU, Utf8 SearchPhrase (70% empty, Zipf over 6M phrases of 1-80 bytes) "
                "- variable-width GroupValues interning (BASELINE configs[4])")
    elif wl == "join":
        name = (f"TPC-H Q3-shaped join #2 (post-filter): build {args.build_rows} x 16 B, probe {n} x 40 B per GPU, Inner, "
                f"{'CollectLeft' if world == 1 else 'broadcast build (NCCL all-gather)'} (BASELINE configs[2])")
    elif wl == "q3":
        name = (f"TPC-H Q3-shaped chain at SF{n} (post-filter scans): customer x orders -> x lineitem -> GROUP BY (l_orderkey, o_orderdate, "
                f"o_shippriority); lineitem sharded over {world} GPU(s), builds replicated (BASELINE configs[2])")
    else:
        raise SystemExit(f"unknown workload {wl}")
    return {"workload": name, "rows_per_gpu": n, "n_gpus": world, "l2_policy": "inputs larger than L2 (no flush needed)"}


# ------------------------------------------------------------------------------------------------
# aggregation workloads (c2, q1, c5, c5s)
# ------------------------------------------------------------------------------------------------
Q1_GROUPS = [(65, 70), (78, 70), (78, 79), (82, 70)]   # (A,F) (N,F) (N,O) (R,F)


def build_agg_workload(env, args):
    """Returns dict(schema, dev, host_cols, group_by, exprs, in_bytes, out_row_bytes, cpu, verify, ...)."""
    import oracle
    torch, df, ctx, device, rank = env.torch, env.df, env.ctx, env.device, env.rank
    w = {}
    n = args.rows
    if args.workload == "c2":
        keys = gen_c2_keys(n, 42 + rank, device)
        schema = pa.schema([("UserID", pa.int64())])
        w["dev"] = df.DeviceBatch.from_device(ctx, schema, [{"buffers": [None, keys.data_ptr()], "null_count": 0}], n, keep=keys)
        w["schema"] = schema
        w["group_by"] = ["UserID"]
        w["exprs"] = [df.AggregateExpr(df.AggFunc.Count, None, "COUNT(*)")]
        w["in_bytes"] = 8 * n
        w["out_row_bytes"] = 16
        host_cols = [("UserID", pa.int64(), keys)]
        w["cpu"] = lambda threads, sample: oracle.bench_groupby_count_i64(keys[:sample].cpu().numpy(), threads)

        def acero(sample):
            t = pa.table({"UserID": keys[:sample].cpu().numpy()})
            t0 = time.perf_counter()
            r = t.group_by("UserID").aggregate([([], "count_all")])
            dt = time.perf_counter() - t0
            return {"value": sample / dt, "unit": "rows/s", "cores": pa.cpu_count(), "groups": r.num_rows,
                    "sample": f"first {sample} rows, pyarrow {pa.__version__} Table.group_by(UserID).aggregate(count_all)"}
        w["acero"] = acero

        def verify(out, cpu_result):
            """Outside the timed region: the result of the last timed step against (a) what torch computes from the inputs of ALL
            ranks (distinct keys, rows, key-weighted checksum sum(mix64(key) * count) mod 2**64) and (b) at N = 1 the oracle's
            group count and checksum on the same rows."""
            k = out.column(0).to_numpy(zero_copy_only=False)
            c = out.column(1).to_numpy(zero_copy_only=False)
            with np.errstate(over="ignore"):
                cks = int((mix64_np(k.view(np.uint64)) * c.astype(np.uint64)).sum(dtype=np.uint64))
            uniq = int(np.unique(k).size) if env.world > 1 or n <= 20_000_000 else int(torch.unique(torch.from_numpy(k).to(device)).numel())
            G, rows, dup = allreduce_i64(env, [len(k), int(c.sum()), len(k) - uniq])
            cks = allreduce_u64_wrapping(env, cks)
            res = {"checked": True, "groups": G, "sum_count": rows, "checksum": cks, "duplicate_groups": dup}
            if env.rank == 0:
                # rank 0 regenerates every rank's shard (seed 42 + r) and lets torch compute the truth
                exp_cks, parts = 0, []
                for r_ in range(env.world):
                    kk = keys if r_ == rank else gen_c2_keys(n, 42 + r_, device)
                    exp_cks = (exp_cks + torch_sum_u64(_mix64_torch(kk))) & MASK64
                    parts.append(torch.unique(kk))
                    del kk
                exp_G = int(torch.unique(torch.cat(parts)).numel()) if env.world > 1 else int(parts[0].numel())
                del parts
                res.update({"groups_expected": exp_G, "rows_expected": n * env.world, "checksum_expected": exp_cks,
                            "against": "torch.unique / torch.sum over the regenerated inputs of all ranks"})
                ok = G == exp_G and rows == n * env.world and cks == exp_cks and dup == 0
                if cpu_result is not None and env.world == 1:
                    res["oracle"] = {"rows": cpu_result["rows"], "groups": cpu_result["groups"], "checksum": cpu_result["checksum"]}
                    if cpu_result["rows"] == n:
                        ok = ok and cpu_result["groups"] == G and cpu_result["checksum"] == cks
                        res["against"] += "; oracle (CPU port of the reference operators) on the same rows"
                res["ok"] = bool(ok)
            return res
        w["verify"] = verify
    elif args.workload == "q1":
        flag, status, cols = gen_q1(n, 42 + rank, device)
        d152, d384, d386 = pa.decimal128(15, 2), pa.decimal128(38, 4), pa.decimal128(38, 6)
        schema = pa.schema([("l_returnflag", pa.uint8()), ("l_linestatus", pa.uint8()), ("l_quantity", d152), ("l_extendedprice", d152),
                            ("l_discount", d152), ("disc_price", d384), ("charge", d386)])
        if args.q1_keys == "utf8":
            # configs[0] as written: l_returnflag / l_linestatus are Utf8 (one byte each) - the same bytes as the UInt8 variant
            # behind i32 offsets 0..n (90 B/row)
            offs = torch.arange(0, n + 1, device=device, dtype=torch.int32)
            schema = pa.schema([pa.field("l_returnflag", pa.string(), nullable=False), pa.field("l_linestatus", pa.string(), nullable=False)] + list(schema)[2:])
        order = ["qty", "price", "disc", "disc_price", "charge"]
        bufs = [{"buffers": [None, flag.data_ptr()]}, {"buffers": [None, status.data_ptr()]}] + [{"buffers": [None, cols[c].data_ptr()]} for c in order]
        keep_offs = None
        if args.q1_keys == "utf8":
            bufs[0] = {"buffers": [None, offs.data_ptr(), flag.data_ptr()]}
            bufs[1] = {"buffers": [None, offs.data_ptr(), status.data_ptr()]}
            keep_offs = offs
        w["dev"] = df.DeviceBatch.from_device(ctx, schema, bufs, n, keep=(flag, status, cols, keep_offs))
        w["schema"] = schema
        w["group_by"] = ["l_returnflag", "l_linestatus"]
        A = df.AggregateExpr
        w["exprs"] = [A(df.AggFunc.Sum, "l_quantity", "sum_qty"), A(df.AggFunc.Sum, "l_extendedprice", "sum_base_price"),
                      A(df.AggFunc.Sum, "disc_price", "sum_disc_price"), A(df.AggFunc.Sum, "charge", "sum_charge"),
                      A(df.AggFunc.Avg, "l_quantity", "avg_qty"), A(df.AggFunc.Avg, "l_extendedprice", "avg_price"),
                      A(df.AggFunc.Avg, "l_discount", "avg_disc"), A(df.AggFunc.Count, None, "count_order")]
        w["in_bytes"] = (90 if args.q1_keys == "utf8" else 82) * n
        w["out_row_bytes"] = 0
        if args.q1_project:
            # disc_price = l_extendedprice * (1 - l_discount) and charge = disc_price * (1 + l_tax) are computed on
            # the device (dfgpu_batch_eval_decimal) instead of arriving as columns: 4 Decimal128 inputs (66 B/row) instead of 5
            g2 = torch.Generator(device=device)
            g2.manual_seed(4242 + rank)
            tv = torch.randint(0, 9, (n,), generator=g2, device=device, dtype=torch.int64)
            cols["tax"] = torch.stack([tv, tv >> 63], dim=1).contiguous()
            order = ["qty", "price", "disc", "tax"]
            w["host_schema"] = pa.schema(list(schema)[:5] + [pa.field("l_tax", d152)])
            bufs = bufs[:2] + [{"buffers": [None, cols[c].data_ptr()]} for c in order]
            w["dev"] = df.DeviceBatch.from_device(ctx, w["host_schema"], bufs, n, keep=(flag, status, cols, keep_offs))
            schema = pa.schema(list(w["host_schema"]) + [pa.field("disc_price", d384), pa.field("charge", d386)])
            w["schema"] = schema
            disc_price = [("col", 3), ("lit", 1, 20, 0), ("col", 4), ("sub", 1, 2), ("mul", 0, 3)]      # price * (1 - disc)
            charge = [("col", 6), ("lit", 1, 20, 0), ("col", 5), ("add", 1, 2), ("mul", 0, 3)]          # disc_price * (1 + tax)
            w["prepare"] = lambda d: d.eval_decimal(disc_price, "disc_price").eval_decimal(charge, "charge")
            w["in_bytes"] -= 16 * n
        kcols = [("l_returnflag", pa.string(), (offs, flag)), ("l_linestatus", pa.string(), (offs, status))] if args.q1_keys == "utf8" else \
                [("l_returnflag", pa.uint8(), flag), ("l_linestatus", pa.uint8(), status)]
        host_cols = kcols + [(nm, t, cols[c]) for (nm, t), c in zip(list(zip(schema.names, schema.types))[2:], order)]

        def cpu(threads, sample):
            f, s = flag[:sample].cpu().numpy(), status[:sample].cpu().numpy()
            d = [cols[c][:sample].cpu().numpy() for c in ("qty", "price", "disc_price", "charge", "disc")]
            return oracle.bench_q1(f, s, d, threads)
        w["cpu"] = cpu

        def verify(out, cpu_result):
            """Exact expectation from torch: per group, every Decimal128 column summed in two 32-bit halves (the generated values fit
            the low 64-bit word and are non-negative), then SUM / AVG rebuilt with Python integers following the reference's rules
            (AVG(Decimal128(p,s)) = sum * 10**4 / count truncated, scale s + 4: aggregate/utils.rs:108-124)."""
            if args.q1_project:
                return {"checked": False, "why": "projection variant: covered by tests/test_gpu_partition.py (decimal.slt goldens)"}
            gid = torch.full((n,), -1, dtype=torch.int64, device=device)
            for i, (f_, s_) in enumerate(Q1_GROUPS):
                gid[(flag == f_) & (status == s_)] = i
            sums = []
            for c in ("qty", "price", "disc_price", "charge", "disc"):
                v = cols[c][:, 0]
                for g_ in range(4):
                    m = gid == g_
                    sums += [int((v & 0xFFFFFFFF)[m].sum()), int((v >> 32)[m].sum())]
            cnts = [int((gid == g_).sum()) for g_ in range(4)]
            tot = allreduce_i64(env, sums + cnts)
            rows_local = []
            for i in range(out.num_rows):
                rows_local.append([out.column(j)[i].as_py() for j in range(out.num_columns)])
            gathered = [None] * env.world if env.rank == 0 else None
            if env.world > 1:
                env.dist.gather_object(rows_local, gathered, dst=0)
            else:
                gathered = [rows_local]
            res = {"checked": True}
            if env.rank == 0:
                got = {}
                for part in gathered:
                    for r_ in part:
                        k0 = r_[0] if isinstance(r_[0], int) else ord(r_[0])
                        k1 = r_[1] if isinstance(r_[1], int) else ord(r_[1])
                        got[(k0, k1)] = [int(x.scaleb(-x.as_tuple().exponent)) if hasattr(x, "as_tuple") else int(x) for x in r_[2:]]
                want = {}
                S = lambda ci, g_: tot[(ci * 4 + g_) * 2] + (tot[(ci * 4 + g_) * 2 + 1] << 32)
                for g_, key in enumerate(Q1_GROUPS):
                    cnt = tot[40 + g_]
                    if cnt == 0:
                        continue
                    want[key] = [S(0, g_), S(1, g_), S(2, g_), S(3, g_), S(0, g_) * 10**4 // cnt, S(1, g_) * 10**4 // cnt, S(4, g_) * 10**4 // cnt, cnt]
                res.update({"groups": len(got), "groups_expected": len(want), "sum_count": sum(v[-1] for v in got.values()), "rows_expected": sum(tot[40:44]),
                            "against": "torch per-group sums over the inputs of all ranks, SUM / AVG rebuilt in exact integer arithmetic",
                            "ok": bool(got == want)})
                if got != want:
                    res["first_diff"] = str([(k_, got.get(k_), want[k_]) for k_ in want if got.get(k_) != want[k_]][:1])[:400]
            return res
        w["verify"] = verify
    elif args.workload in ("c5", "c5s"):
        region, offsets, data = gen_c5(n, 42 + rank, device)
        schema = pa.schema([pa.field("RegionID", pa.int32(), nullable=False), pa.field("SearchPhrase", pa.string(), nullable=False)])   # ClickBench hits: NOT NULL
        w["dev"] = df.DeviceBatch.from_device(ctx, schema, [{"buffers": [None, region.data_ptr()], "null_count": 0},
                                                            {"buffers": [None, offsets.data_ptr(), data.data_ptr()], "null_count": 0}],
                                              n, keep=(region, offsets, data))
        w["schema"] = schema
        w["group_by"] = ["RegionID", "SearchPhrase"]
        w["exprs"] = [df.AggregateExpr(df.AggFunc.Count, None, "COUNT(*)")]
        w["in_bytes"] = 4 * n + 4 * (n + 1) + int(data.numel())
        if args.workload == "c5s":      # the single-key variant: GROUP BY SearchPhrase (GroupValuesByes, bytes.rs:44-121)
            w["group_by"] = ["SearchPhrase"]
            w["in_bytes"] -= 4 * n
        w["out_row_bytes"] = 4 + 4 + 8          # RegionID + phrase offset + count (the distinct phrases' own bytes are not counted)
        host_cols = [("RegionID", pa.int32(), region), ("SearchPhrase", pa.string(), (offsets, data))]

        def cpu(threads, sample):
            # the oracle's generic operator (GroupValuesRows, row.rs:94-146), single-threaded, 8192-row batches
            hb = pa.record_batch([pa.array(region[:sample].cpu().numpy()),
                                  pa.Array.from_buffers(pa.string(), sample, [None, pa.py_buffer(offsets[:sample + 1].cpu().numpy()),
                                                                             pa.py_buffer(data.cpu().numpy())])], schema=schema)
            o = oracle.OracleAgg("single", w["group_by"], [schema.field(g).type for g in w["group_by"]], [oracle.AggSpec("count", "COUNT(*)", None)])
            t0 = time.perf_counter()
            for off in range(0, sample, 8192):
                s = hb.slice(off, 8192)
                o.push([s.column(schema.get_field_index(g)) for g in w["group_by"]], [None])
            out = o.emit()
            return time.perf_counter() - t0, out.num_rows, 0
        w["cpu"] = cpu
        w["cpu_threads"] = 1
        w["cpu_sample"] = 4_000_000

        def verify(out, cpu_result):
            """Counts sum to the rows; the number of groups equals the number of distinct (RegionID, phrase id) pairs of the generator
            (a phrase id determines its bytes and distinct ids give distinct (length, bytes) with overwhelming probability - the exact
            string parity is tests/test_gpu_two_level.py's)."""
            c = out.column(out.num_columns - 1).to_numpy(zero_copy_only=False)
            G, rows = allreduce_i64(env, [out.num_rows, int(c.sum())])
            res = {"checked": True, "groups": G, "sum_count": rows, "rows_expected": n * env.world}
            if env.rank == 0:
                res["ok"] = bool(rows == n * env.world and G > 0)
                res["against"] = "row conservation (sum of COUNT(*) = input rows of all ranks)"
            return res
        w["verify"] = verify
    else:
        raise SystemExit(f"unknown workload {args.workload}")
    w["host_cols"] = host_cols
    return w


def run_agg(env, args, with_e2e=True, with_cpu=True):
    torch, df, ctx, world, rank = env.torch, env.df, env.ctx, env.world, env.rank
    w = build_agg_workload(env, args)
    n = args.rows
    torch.cuda.synchronize()
    hold = []

    def step(inp, to_host=False):
        """One pass of the hot path over one batch.  inp: DeviceBatch (inputs resident in HBM) or a pinned
        host RecordBatch (H2D inside the step); to_host: read the result back into host Arrow buffers."""
        hold.clear()
        if w.get("prepare"):
            inp = w["prepare"](inp if isinstance(inp, df.DeviceBatch) else df.DeviceBatch.from_arrow(ctx, inp))
        if world == 1 or not args.three_op:
            # N > 1, fused: level 1 on the local shard -> record exchange over NVLink -> final merge of the owned partitions
            op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, w["schema"], w["group_by"], w["exprs"])
            op.push(inp)
            if world > 1:
                op.exchange()
            out = op.emit() if to_host else op.emit_device()
            hold[:] = [(op, out)]
            return out
        # Partial -> hash repartition (NCCL all-to-all over NVLink) -> FinalPartitioned
        part = df.GpuAggregateExec(ctx, df.AggregateMode.Partial, w["schema"], w["group_by"], w["exprs"])
        part.push(inp)
        st = part.emit_device()
        rp = df.GpuRepartitionExec(ctx, part.schema, list(range(len(w["group_by"]))), world)
        recv = rp.exchange(rp.partition_device(st))
        fin = df.GpuAggregateExec(ctx, df.AggregateMode.FinalPartitioned, part.schema, list(range(len(w["group_by"]))),
                                  [df.AggregateExpr(e.func, None, e.name) for e in w["exprs"]])
        if recv is not None:
            fin.push(recv)
        out = fin.emit() if to_host else fin.emit_device()
        hold[:] = [(part, st, recv, fin, out)]
        return out

    r = timed_steps(env, args.steps, args.warmup, lambda: step(w["dev"]))
    out_host = r["out"].to_arrow()          # the result of the LAST TIMED step, read back after the timed region
    groups_local = out_host.num_rows
    r["out"] = None
    hold.clear()
    groups = allreduce_i64(env, [groups_local])[0]
    ms_per_step = r["ms_per_step"]
    value = n * world / (ms_per_step / 1e3)

    # ---- e2e: host (pinned) Arrow buffers through the C-ABI, H2D + D2H inside the timed region ----
    e2e = None
    if with_e2e:
        hb, keep, h2d = pinned_batch(ctx, w.get("host_schema", w["schema"]), w["host_cols"], n, torch)
        exported = []
        e2e_step = lambda: step(hb, to_host=True)
        if args.batch_rows > 0 and world == 1:
            # the reference's batch granularity (8192 rows, common/src/config.rs:215): one dfgpu_agg_push_arrow per slice of
            # the pinned host batch; the C-Data structs are exported before the timed region (a Rust caller has them for free)
            from datafusion_upstream_b200 import _native as N
            for o in range(0, n, args.batch_rows):
                a, sc = N.ArrowArray(), N.ArrowSchema()
                hb.slice(o, args.batch_rows)._export_to_c(ctypes.addressof(a), ctypes.addressof(sc))
                exported.append((a, sc))
            L = N.lib()

            ptrs = (ctypes.c_void_p * len(exported))(*[ctypes.addressof(a) for a, _ in exported])

            def e2e_step():
                hold.clear()
                op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, w["schema"], w["group_by"], w["exprs"])
                if args.batch_api == "many":      # what the stream has ready, in one call (packed by several host threads)
                    N.check(L.dfgpu_agg_push_arrow_batches(op.h, ptrs, len(exported)))
                else:
                    for a, _ in exported:
                        N.check(L.dfgpu_agg_push_arrow(op.h, ctypes.addressof(a)))
                out = op.emit()
                hold[:] = [(op, out)]
                return out
        dt, h2d, d2h = timed_e2e(env, args.steps, e2e_step, h2d)
        hold.clear()
        e2e = {"value": n * world / dt, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": dt * 1e3}
        if exported:
            e2e["batch_rows"] = args.batch_rows
            e2e["pushes_per_step"] = len(exported)
            e2e["batch_api"] = "dfgpu_agg_push_arrow_batches" if args.batch_api == "many" else "dfgpu_agg_push_arrow"
            for a, sc in exported:
                if a.release:
                    ctypes.CFUNCTYPE(None, ctypes.POINTER(N.ArrowArray))(a.release)(ctypes.byref(a))
                N.release_schema(sc)
        del hb
        for p in keep:
            ctx.host_free(p)

    # every consumed input column once + every output column once, per GPU
    algo_bytes = w["in_bytes"] + w["out_row_bytes"] * (groups // world)
    roof = roofline_of(env, r, args.steps, algo_bytes, args.workload)
    # ---- CPU baseline (the oracle port, reference-shaped parallel plan) on rank 0 ----
    cpu, cpu_result = None, None
    if rank == 0 and with_cpu:
        threads = w.get("cpu_threads") or os.cpu_count() or 1
        sample = min(n, w.get("cpu_sample") or args.cpu_sample)
        secs, g, cks = w["cpu"](threads, sample)
        cpu_result = {"rows": sample, "groups": g, "checksum": cks}
        shape = "P=%d Partial -> hash repartition -> FinalPartitioned" % threads if threads > 1 else "one Single-mode operator (scalar port)"
        cpu = {"value": sample / secs, "unit": "rows/s", "cores": threads, "kind": "port",
               "sample": (f"all {sample} rows of rank 0's input" if sample == n else f"first {sample} of the {n} rows") + f", {shape}, 8192-row batches"}
        if w.get("acero") and not args.no_acero:
            # a second, independent CPU data point: pyarrow's Acero hash aggregation on the same Arrow columns
            try:
                cpu["acero"] = w["acero"](min(n, 20_000_000))
            except Exception as e:      # the bench line must not depend on it
                cpu["acero"] = {"unavailable": str(e)[:120]}
    parity = w["verify"](out_host, cpu_result) if not args.no_parity else {"checked": False}
    del out_host
    mode = "Single" if world == 1 else (f"Partial -> NCCL all-to-all -> FinalPartitioned (dp{world})" if args.three_op else
                                        f"fused Partial -> record exchange over NVLink peer memory -> FinalPartitioned in one operator per rank (dp{world})")
    line = {
        "metric": "input rows/s (hash aggregation, device-resident input) and HBM GB/s", "value": value, "unit": "rows/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": {"c2": "int64", "c5": "u8", "c5s": "u8"}.get(args.workload, "i128"), "data": "synthetic",
        "config": workload_config(args, world),
        "result": {"groups": groups, "mode": mode},
        "algorithmic_GBps": algo_bytes * world / (ms_per_step * 1e-3) / 1e9,
        "algorithmic_frac_of_peak": algo_bytes / (ms_per_step * 1e-3) / 1e9 / peaks()[0],
        "host_submit_ms": r["host_ms"], "clocks": r["clocks"], "e2e": e2e, "gpu_launches": r["launches"], "roofline": roof, "cpu_baseline": cpu,
        "parity": parity,
    }
    del w
    return line


# ------------------------------------------------------------------------------------------------
# join (Q3 join #2) and the Q3 chain
# ------------------------------------------------------------------------------------------------
def gen_join(n_build, n_probe, seed, device, rank=0, world=1):
    """Q3-shaped join #2 (post-filter): build = (o_orderkey, o_orderdate, o_shippriority) of the orders that
    survived join #1, probe = (l_orderkey, l_extendedprice, l_discount) of the filtered lineitem rows.  Order keys
    are sparse like TPC-H's (8 of every 32); ~10 % of the probe rows find a build row."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    n_orders = max(15_000_000, 10 * n_build)
    allk = torch.arange(0, n_orders, device=device, dtype=torch.int64)
    allk = (allk // 8) * 32 + (allk % 8) + 1
    perm = torch.randperm(n_orders, generator=g, device=device)
    bkeys = allk[perm[:n_build]].contiguous()
    del perm
    odate = torch.randint(8035, 9204, (n_build,), generator=g, device=device, dtype=torch.int32)
    oprio = torch.zeros(n_build, device=device, dtype=torch.int32)
    gp = torch.Generator(device=device)
    gp.manual_seed(seed + 1000 + rank)
    pk = allk[torch.randint(0, n_orders, (n_probe,), generator=gp, device=device)].contiguous()
    del allk

    def dec(lo, hi):
        v = torch.randint(lo, hi, (n_probe,), generator=gp, device=device, dtype=torch.int64)
        return torch.stack([v, v >> 63], dim=1).contiguous()
    return (bkeys, odate, oprio), (pk, dec(90000, 10495001), dec(0, 11))


def run_join(env, args, with_e2e=True, with_cpu=True):
    """--workload join: HashJoinExec build + probe (Inner, CollectLeft / broadcast build) on the Q3-shaped join."""
    import oracle
    torch, df, ctx, world, rank, device = env.torch, env.df, env.ctx, env.world, env.rank, env.device
    from datafusion_upstream_b200.dist import shard_range
    B, P = args.build_rows, args.rows
    bt, pt = gen_join(B, P, 42, device, rank, world)
    bs = pa.schema([("o_orderkey", pa.int64()), ("o_orderdate", pa.date32()), ("o_shippriority", pa.int32())])
    ps = pa.schema([("l_orderkey", pa.int64()), ("l_extendedprice", pa.decimal128(15, 2)), ("l_discount", pa.decimal128(15, 2))])
    lo, hi = shard_range(B, rank, world)           # every rank holds a row range of the build side; it is all-gathered per step
    bfull = df.DeviceBatch.from_device(ctx, bs, [{"buffers": [None, t.data_ptr()]} for t in bt], B, keep=bt)
    bdev = bfull.slice(lo, hi - lo)
    pdev = df.DeviceBatch.from_device(ctx, ps, [{"buffers": [None, t.data_ptr()]} for t in pt], P, keep=pt)
    hold = []

    def step(b, p, to_host=False):
        hold.clear()
        j = df.GpuHashJoinExec(ctx, bs, ps, [("o_orderkey", "l_orderkey")], df.JoinType.Inner, df.PartitionMode.CollectLeft)
        bb = b if isinstance(b, df.DeviceBatch) else df.DeviceBatch.from_arrow(ctx, b)
        if world > 1:
            bb = df.all_gather(ctx, bs, bb)        # broadcast build: the probe side never moves
        j.build([bb])
        out = j.probe(p) if to_host else j.probe_device(p)
        hold[:] = [(j, bb, out)]
        return out

    r = timed_steps(env, args.steps, args.warmup, lambda: step(bdev, pdev))
    out_dev = r["out"]
    out_rows_local = out_dev.num_rows
    out_rows = allreduce_i64(env, [out_rows_local])[0]
    ms_per_step = r["ms_per_step"]
    rows_in = B + P * world
    value = rows_in / (ms_per_step * 1e-3)
    algo = B * 16 + P * 40 + (out_rows // world) * 56      # per GPU: the build side is read by every rank
    # ---- parity: the joined rows of the last timed step against torch on the same inputs -------------------------------------
    parity = {"checked": False}
    if not args.no_parity:
        oh = out_dev.to_arrow()
        ok_col = oh.column(oh.schema.get_field_index("l_orderkey")).to_numpy(zero_copy_only=False)
        od_col = oh.column(oh.schema.get_field_index("o_orderdate")).cast(pa.int32()).to_numpy(zero_copy_only=False)
        price = oh.column(oh.schema.get_field_index("l_extendedprice"))
        price_lo = np.frombuffer(price.buffers()[1], dtype=np.int64, count=2 * len(price), offset=price.offset * 16)[0::2]
        with np.errstate(over="ignore"):
            got = [int(mix64_np(ok_col.view(np.uint64)).sum(dtype=np.uint64)), int(od_col.astype(np.int64).sum()), int(price_lo.sum())]
        # truth from torch: build keys are unique, so a probe row is output once iff its key is in the build side
        sk, order = torch.sort(bt[0])
        pos = torch.searchsorted(sk, pt[0]).clamp_(max=B - 1)
        hit = sk[pos] == pt[0]
        want = [torch_sum_u64(_mix64_torch(pt[0][hit])), int(bt[1][order][pos[hit]].to(torch.int64).sum()), int(pt[1][:, 0][hit].sum())]
        want_rows = int(hit.sum())
        tot = allreduce_i64(env, [want_rows, got[1], got[2], want[1], want[2]])
        g_cks, w_cks = allreduce_u64_wrapping(env, got[0]), allreduce_u64_wrapping(env, want[0])
        parity = {"checked": True, "output_rows": out_rows, "output_rows_expected": tot[0], "key_checksum": g_cks, "key_checksum_expected": w_cks,
                  "sum_o_orderdate": tot[1], "sum_o_orderdate_expected": tot[3], "sum_l_extendedprice": tot[2], "sum_l_extendedprice_expected": tot[4],
                  "against": "torch.sort / searchsorted membership of every probe key in the build side, payload sums of both sides"}
        parity["ok"] = bool(out_rows == tot[0] and g_cks == w_cks and tot[1] == tot[3] and tot[2] == tot[4])
        del oh, sk, order, pos, hit
    r["out"] = None
    out_dev = None
    hold.clear()
    e2e = None
    if with_e2e:
        host_b = [(f.name, f.type, t[lo:hi]) for f, t in zip(bs, bt)]
        host_p = [(f.name, f.type, t) for f, t in zip(ps, pt)]
        hb, keep_b, h2d_b = pinned_batch(ctx, bs, host_b, hi - lo, torch)
        hp, keep_p, h2d_p = pinned_batch(ctx, ps, host_p, P, torch)
        dt, h2d, d2h = timed_e2e(env, args.steps, lambda: step(hb, hp, True), h2d_b + h2d_p)
        hold.clear()
        e2e = {"value": rows_in / dt, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": dt * 1e3}
        del hb, hp
        for p_ in keep_b + keep_p:
            ctx.host_free(p_)
    roof = roofline_of(env, r, args.steps, algo, "join")
    cpu = None
    if rank == 0 and with_cpu:
        threads = os.cpu_count() or 1
        sample = min(P, args.cpu_sample)
        secs, r_, _ = oracle.bench_join_i64(bt[0].cpu().numpy(), pt[0][:sample].cpu().numpy(), threads)
        cpu = {"value": (B + sample) / secs, "unit": "rows/s", "cores": threads, "kind": "port",
               "sample": f"build {B} + first {sample} probe keys (Int64 keys only, no payload gather), P={threads} Partitioned hash joins, 8192-row batches"}
    return {
        "metric": "input rows/s (hash join build+probe, device-resident input) and HBM GB/s", "value": value, "unit": "rows/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": workload_config(args, world),
        "result": {"build_rows": B, "probe_rows_per_gpu": P, "output_rows": out_rows},
        "algorithmic_GBps": algo * world / (ms_per_step * 1e-3) / 1e9, "algorithmic_frac_of_peak": algo / (ms_per_step * 1e-3) / 1e9 / peaks()[0],
        "host_submit_ms": r["host_ms"], "clocks": r["clocks"], "e2e": e2e, "gpu_launches": r["launches"], "roofline": roof, "cpu_baseline": cpu, "parity": parity}


def gen_q3(sf, seed, device, rank=0, world=1):
    """Post-filter inputs of the three scans of TPC-H Q3 at scale factor `sf`, generated on the device:
    customer (c_custkey; c_mktsegment keeps 1 in 5), orders (o_orderkey sparse 8-of-32, o_custkey, o_orderdate < 1995-03-15,
    o_shippriority = 0), lineitem (l_orderkey, l_extendedprice, l_discount; l_shipdate = o_orderdate + 1..121 > 1995-03-15).
    With world > 1 the lineitem rows are this rank's shard (1/world of the table, its own seed); customer and orders are replicated."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    n_cust, n_orders, n_line = 150_000 * sf, 1_500_000 * sf, 6_000_000 * sf // world
    custkey = torch.arange(1, n_cust + 1, device=device, dtype=torch.int64)
    cust = custkey[torch.rand(n_cust, generator=g, device=device) < 0.2].contiguous()
    ok = torch.arange(0, n_orders, device=device, dtype=torch.int64)
    okey = (ok // 8) * 32 + (ok % 8) + 1
    valid = custkey[custkey % 3 != 0]
    ocust = valid[torch.randint(0, valid.numel(), (n_orders,), generator=g, device=device)]
    odate = torch.randint(8035, 10440, (n_orders,), generator=g, device=device, dtype=torch.int32)
    keep = odate < 9204
    orders = (okey[keep].contiguous(), ocust[keep].contiguous(), odate[keep].contiguous(),
              torch.zeros(int(keep.sum()), device=device, dtype=torch.int32))
    gl = torch.Generator(device=device)
    gl.manual_seed(seed + 1000 + rank)
    oi = torch.randint(0, n_orders, (n_line,), generator=gl, device=device)
    ship = odate[oi] + torch.randint(1, 122, (n_line,), generator=gl, device=device, dtype=torch.int32)
    lkey = okey[oi][ship > 9204].contiguous()
    n_l = lkey.numel()
    del oi, ship

    def dec(lo, hi):
        v = torch.randint(lo, hi, (n_l,), generator=gl, device=device, dtype=torch.int64)
        return torch.stack([v, v >> 63], dim=1).contiguous()
    return (cust,), orders, (lkey, dec(90000, 10495001), dec(0, 11))


def run_q3(env, args, with_e2e=True, with_cpu=True):
    """--workload q3: the whole Q3-shaped operator chain (BASELINE configs[2]): join #1 customer x orders, join #2 x lineitem,
    then GROUP BY (l_orderkey, o_orderdate, o_shippriority) SUM/COUNT - Single on one GPU, Partial -> NCCL all-to-all ->
    FinalPartitioned when sharded (lineitem sharded, customer / orders replicated = CollectLeft builds)."""
    import oracle
    torch, df, ctx, world, rank, device = env.torch, env.df, env.ctx, env.world, env.rank, env.device
    sf = max(1, args.rows)
    ct, ot, lt = gen_q3(sf, 42, device, rank, world)
    cs = pa.schema([("c_custkey", pa.int64())])
    os_ = pa.schema([("o_orderkey", pa.int64()), ("o_custkey", pa.int64()), ("o_orderdate", pa.date32()), ("o_shippriority", pa.int32())])
    ls = pa.schema([("l_orderkey", pa.int64()), ("l_extendedprice", pa.decimal128(15, 2)), ("l_discount", pa.decimal128(15, 2))])

    def devb(schema, ts):
        return df.DeviceBatch.from_device(ctx, schema, [{"buffers": [None, t.data_ptr()]} for t in ts], ts[0].shape[0], keep=ts)
    cdev, odev, ldev = devb(cs, ct), devb(os_, ot), devb(ls, lt)
    nC, nO, nL = ct[0].shape[0], ot[0].shape[0], lt[0].shape[0]
    group = ["l_orderkey", "o_orderdate", "o_shippriority"]
    exprs = [df.AggregateExpr(df.AggFunc.Sum, "l_extendedprice", "SUM(l_extendedprice)"), df.AggregateExpr(df.AggFunc.Sum, "l_discount", "SUM(l_discount)"),
             df.AggregateExpr(df.AggFunc.Count, None, "COUNT(*)")]
    hold = []
    stats = {}

    def step(c, o, l, to_host=False):
        hold.clear()
        j1 = df.GpuHashJoinExec(ctx, cs, os_, [("c_custkey", "o_custkey")], df.JoinType.Inner, df.PartitionMode.CollectLeft)
        j1.build([c])
        o1 = j1.probe_device(o).project(["o_orderkey", "o_orderdate", "o_shippriority"])
        j2 = df.GpuHashJoinExec(ctx, o1.schema, ls, [("o_orderkey", "l_orderkey")], df.JoinType.Inner, df.PartitionMode.CollectLeft)
        j2.build([o1])
        jo = j2.probe_device(l).project(group + ["l_extendedprice", "l_discount"])
        stats["j1"], stats["j2"] = o1.num_rows, jo.num_rows
        if world == 1:
            agg = df.GpuAggregateExec(ctx, df.AggregateMode.Single, jo.schema, group, exprs)
            agg.push(jo)
            out = agg.emit() if to_host else agg.emit_device()
            hold[:] = [(j1, j2, o1, jo, agg, out)]
            return out
        part = df.GpuAggregateExec(ctx, df.AggregateMode.Partial, jo.schema, group, exprs)
        part.push(jo)
        st = part.emit_device()
        rp = df.GpuRepartitionExec(ctx, part.schema, [0, 1, 2], world)
        recv = rp.exchange(rp.partition_device(st))
        fin = df.GpuAggregateExec(ctx, df.AggregateMode.FinalPartitioned, part.schema, [0, 1, 2], [df.AggregateExpr(e.func, None, e.name) for e in exprs])
        if recv is not None:
            fin.push(recv)
        out = fin.emit() if to_host else fin.emit_device()
        hold[:] = [(j1, j2, o1, jo, part, st, recv, fin, out)]
        return out

    r = timed_steps(env, args.steps, args.warmup, lambda: step(cdev, odev, ldev))
    out_dev = r["out"]
    groups_local = out_dev.num_rows
    groups, nL_all, j2_all = allreduce_i64(env, [groups_local, nL, stats["j2"]])
    ms_per_step = r["ms_per_step"]
    rows_in = (nC + nO) * world + nL_all          # every rank reads its replicated build sides
    value = rows_in / (ms_per_step * 1e-3)
    # consumed input columns once + output rows once, per operator, per GPU
    algo = (nC * 8 + nO * 24 + stats["j1"] * 16) + (stats["j1"] * 16 + nL * 40 + stats["j2"] * 56) + (stats["j2"] * 48 + groups_local * 56)
    parity = {"checked": False}
    if not args.no_parity:
        oh = out_dev.to_arrow()
        cnt = oh.column(oh.num_columns - 1).to_numpy(zero_copy_only=False)
        sp = oh.column(oh.schema.get_field_index("SUM(l_extendedprice)"))
        sp_lo = np.frombuffer(sp.buffers()[1], dtype=np.int64, count=2 * len(sp), offset=sp.offset * 16)[0::2]
        okeys = oh.column(0).to_numpy(zero_copy_only=False)
        got = [int(cnt.sum()), int(sp_lo.sum()), len(okeys) - int(np.unique(okeys).size)]
        # truth from torch: orders whose customer survived, lineitems whose order survived; a group = a surviving order with >= 1 lineitem
        cs_sorted = torch.sort(ct[0])[0]
        p1 = torch.searchsorted(cs_sorted, ot[1]).clamp_(max=nC - 1)
        o_hit = cs_sorted[p1] == ot[1]
        ok1 = torch.sort(ot[0][o_hit])[0]
        p2 = torch.searchsorted(ok1, lt[0]).clamp_(max=max(int(ok1.numel()) - 1, 0))
        l_hit = ok1[p2] == lt[0]
        want = [int(l_hit.sum()), int(lt[1][:, 0][l_hit].sum())]
        # distinct surviving orders with a lineitem, over ALL ranks: a mark per surviving order, max-reduced
        mark = torch.zeros(int(ok1.numel()), dtype=torch.uint8, device=device)
        mark[p2[l_hit]] = 1
        if world > 1:
            mc = mark.cpu()
            env.dist.all_reduce(mc, op=env.dist.ReduceOp.MAX)
            want_groups = int(mc.sum())
        else:
            want_groups = int(mark.sum())
        tot = allreduce_i64(env, got + want)
        parity = {"checked": True, "join1_rows": stats["j1"], "join1_rows_expected": int(o_hit.sum()), "groups": groups, "groups_expected": want_groups,
                  "sum_count": tot[0], "join2_rows_expected": tot[3], "sum_l_extendedprice": tot[1], "sum_l_extendedprice_expected": tot[4],
                  "duplicate_groups": tot[2], "against": "torch.sort / searchsorted semi-joins of the same inputs on all ranks"}
        parity["ok"] = bool(stats["j1"] == int(o_hit.sum()) and groups == want_groups and tot[0] == tot[3] and tot[1] == tot[4] and tot[2] == 0 and j2_all == tot[3])
        del oh, cs_sorted, p1, o_hit, ok1, p2, l_hit, mark
    r["out"] = None
    out_dev = None
    hold.clear()
    e2e = None
    if with_e2e:
        hc, kc, b1 = pinned_batch(ctx, cs, [(f.name, f.type, t) for f, t in zip(cs, ct)], nC, torch)
        ho, ko, b2 = pinned_batch(ctx, os_, [(f.name, f.type, t) for f, t in zip(os_, ot)], nO, torch)
        hl, kl, b3 = pinned_batch(ctx, ls, [(f.name, f.type, t) for f, t in zip(ls, lt)], nL, torch)
        dt, h2d, d2h = timed_e2e(env, args.steps, lambda: step(hc, ho, hl, True), b1 + b2 + b3)
        hold.clear()
        e2e = {"value": rows_in / dt, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": dt * 1e3}
        del hc, ho, hl
        for p_ in kc + ko + kl:
            ctx.host_free(p_)
    roof = roofline_of(env, r, args.steps, algo, "q3")
    cpu = None
    if rank == 0 and with_cpu:
        # the oracle's partition-parallel drivers, operator by operator (keys only for the joins; the aggregation is its Int64
        # GROUP BY ... COUNT driver over the joined l_orderkey column - the 3-column key is a function of l_orderkey here)
        threads = os.cpu_count() or 1
        s1, r1, _ = oracle.bench_join_i64(ct[0].cpu().numpy(), ot[1].cpu().numpy(), threads)
        cs_sorted = torch.sort(ct[0])[0]
        o_hit = cs_sorted[torch.searchsorted(cs_sorted, ot[1]).clamp_(max=nC - 1)] == ot[1]
        ok1 = ot[0][o_hit]
        s2, r2, _ = oracle.bench_join_i64(ok1.cpu().numpy(), lt[0].cpu().numpy(), threads)
        oks = torch.sort(ok1)[0]
        l_hit = oks[torch.searchsorted(oks, lt[0]).clamp_(max=max(int(oks.numel()) - 1, 0))] == lt[0]
        s3, g3, _ = oracle.bench_groupby_count_i64(lt[0][l_hit].cpu().numpy(), threads)
        cpu = {"value": (nC + nO + nL) / (s1 + s2 + s3), "unit": "rows/s", "cores": threads, "kind": "port",
               "sample": f"rank 0's inputs ({nC} + {nO} + {nL} rows): join #1 {s1 * 1e3:.0f} ms + join #2 {s2 * 1e3:.0f} ms (Int64 keys only, no payload gather) + "
                         f"GROUP BY l_orderkey COUNT over the {int(l_hit.sum())} joined rows {s3 * 1e3:.0f} ms, P={threads} partitions each"}
    return {
        "metric": "input rows/s (Q3-shaped join + join + aggregation chain, device-resident input) and HBM GB/s", "value": value, "unit": "rows/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "i128", "data": "synthetic",
        "config": workload_config(args, world),
        "result": {"customer_rows": nC, "orders_rows": nO, "lineitem_rows_per_gpu": nL, "join1_rows": stats["j1"], "join2_rows": j2_all, "groups": groups,
                   "mode": "Single" if world == 1 else f"CollectLeft builds replicated, lineitem sharded, Partial -> NCCL all-to-all -> FinalPartitioned (dp{world})"},
        "algorithmic_GBps": algo * world / (ms_per_step * 1e-3) / 1e9, "algorithmic_frac_of_peak": algo / (ms_per_step * 1e-3) / 1e9 / peaks()[0],
        "host_submit_ms": r["host_ms"], "clocks": r["clocks"], "e2e": e2e, "gpu_launches": r["launches"], "roofline": roof, "cpu_baseline": cpu, "parity": parity}


# ------------------------------------------------------------------------------------------------
# the north_star targets, inside the default line (the driver only runs the default command)
# ------------------------------------------------------------------------------------------------
def run_targets(env, args):
    """Device-timed Q1 (both key variants) and Q3 (join #2 alone, and the whole chain) at TPC-H SF100 TOTAL size, sharded over the
    N GPUs of the run (N = 1: the whole of SF100 on one B200; N = 2/4/8: BASELINE configs[3]).  Each entry: ms per step, rows/s,
    SURVEY 8(d) algorithmic GB/s and fraction of the measured HBM peak, dominant kernel, and the parity verdict."""
    import copy
    torch = env.torch
    out = {}
    sf = args.targets_sf
    specs = [
        ("q1_sf%d_uint8_keys" % sf, dict(workload="q1", rows=6_000_379 * sf // env.world, q1_keys="u8")),
        ("q1_sf%d_utf8_keys" % sf, dict(workload="q1", rows=6_000_379 * sf // env.world, q1_keys="utf8")),
        ("q3_join2_sf%d" % sf, dict(workload="join", rows=3_240_000 * sf // env.world, build_rows=146_000 * sf)),
        ("q3_chain_sf%d" % sf, dict(workload="q3", rows=sf)),
    ]
    for name, kw in specs:
        a = copy.copy(args)
        a.steps, a.warmup, a.q1_project, a.batch_rows, a.three_op = args.targets_steps, 3, False, 0, False
        for k, v in kw.items():
            setattr(a, k, v)
        try:
            fn = {"q1": run_agg, "join": run_join, "q3": run_q3}[a.workload]
            ln = fn(env, a, with_e2e=False, with_cpu=False)
            roof = ln["roofline"] or {}
            out[name] = {"workload": ln["config"]["workload"], "rows_in": int(round(ln["value"] * ln["ms_per_step"] / 1e3)), "n_gpus": env.world,
                         "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ln["ms_per_step"], 4), "rows_per_s": ln["value"],
                         "algorithmic_bytes_per_gpu": roof.get("algorithmic_bytes_per_step"), "algorithmic_GBps_per_gpu": (roof.get("step") or {}).get("achieved"),
                         "frac_of_measured_hbm_peak": (roof.get("step") or {}).get("frac"), "peak": roof.get("peak"),
                         "dominant_kernel": roof.get("kernel"), "dominant_kernel_ms": roof.get("kernel_ms_per_step"), "kernels": roof.get("kernels"),
                         "stages": roof.get("stages"), "result": ln.get("result"), "parity": ln.get("parity"), "clocks": ln.get("clocks")}
        except Exception as e:      # a target that does not fit / fails must not take the headline line down with it
            out[name] = {"error": f"{type(e).__name__}: {str(e)[:300]}"}
        barrier(env)
        import gc
        gc.collect()
        torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------
# reference arm: the oracle port of the reference's CPU operators on the box's host cores
# ------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's own CPU implementation of the path, on the box's host cores.  The reference is
    Rust and cannot be built here (no cargo/rustc, crates not vendored: DESIGN.md), so this arm times
    the oracle port in the reference's parallel shape (P x Partial -> hash repartition -> P x Final) on the SAME
    workload size the GPU arm runs per GPU (all `--rows` rows; same `config`)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    import oracle
    threads = os.cpu_count() or 1
    n = args.rows
    rng = np.random.default_rng(42)
    dtype = "int64"
    if args.workload == "c2":
        u = rng.random(n)
        keys = mix64_np(np.exp(u * math.log(85_000_000)).astype(np.int64).view(np.uint64)).view(np.int64)
        del u
        run = lambda: oracle.bench_groupby_count_i64(keys, threads)
        rows_in = n
        sample = f"all {n} rows of one GPU's input per step (numpy generator, same distribution), P={threads} threads: Partial -> hash repartition -> FinalPartitioned, 8192-row batches"
    elif args.workload == "q1":
        pick = rng.choice(4, n, p=[0.25, 0.006, 0.493, 0.251])
        flag = np.array([65, 78, 78, 82], np.uint8)[pick]
        status = np.array([70, 70, 79, 70], np.uint8)[pick]

        def dec(lo, hi):
            v = rng.integers(lo, hi, n)
            return np.stack([v, v >> 63], axis=1).copy()
        d = [dec(100, 5001), dec(90000, 10495001), dec(10**9, 10**11), dec(10**11, 10**13), dec(0, 11)]
        run = lambda: oracle.bench_q1(flag, status, d, threads)
        rows_in = n
        dtype = "i128"
        sample = f"all {n} rows per step, 2 x UInt8 keys, P={threads} threads: Partial -> hash repartition -> FinalPartitioned, 8192-row batches"
    elif args.workload == "join":
        B = args.build_rows
        n_orders = max(15_000_000, 10 * B)
        allk = np.arange(n_orders, dtype=np.int64)
        allk = (allk // 8) * 32 + (allk % 8) + 1
        bk = allk[rng.permutation(n_orders)[:B]].copy()
        pk = allk[rng.integers(0, n_orders, n)].copy()
        del allk
        run = lambda: oracle.bench_join_i64(bk, pk, threads)
        rows_in = B + n
        sample = f"build {B} + all {n} probe keys per step (Int64 keys only, no payload gather), P={threads} Partitioned hash joins, 8192-row batches"
    elif args.workload == "q3":
        sf = max(1, n)
        n_cust, n_orders, n_line = 150_000 * sf, 1_500_000 * sf, 6_000_000 * sf      # the WHOLE job (the GPU arm shards lineitem over its ranks)
        custkey = np.arange(1, n_cust + 1, dtype=np.int64)
        cust = custkey[rng.random(n_cust) < 0.2]
        ok = np.arange(n_orders, dtype=np.int64)
        okey = (ok // 8) * 32 + (ok % 8) + 1
        valid = custkey[custkey % 3 != 0]
        ocust = valid[rng.integers(0, valid.size, n_orders)]
        odate = rng.integers(8035, 10440, n_orders).astype(np.int32)
        keep = odate < 9204
        o_key, o_cust = okey[keep], ocust[keep]
        oi = rng.integers(0, n_orders, n_line)
        ship = odate[oi] + rng.integers(1, 122, n_line).astype(np.int32)
        lkey = okey[oi][ship > 9204]
        del oi, ship
        ok1 = o_key[np.isin(o_cust, cust)]
        joined = lkey[np.isin(lkey, ok1)]
        rows_in = cust.size + o_key.size + lkey.size

        def run():
            s1, _, _ = oracle.bench_join_i64(cust, o_cust, threads)
            s2, _, _ = oracle.bench_join_i64(ok1, lkey, threads)
            s3, g, c = oracle.bench_groupby_count_i64(joined, threads)
            return s1 + s2 + s3, g, c
        dtype = "i128"
        sample = (f"the whole job's inputs ({cust.size} + {o_key.size} + {lkey.size} rows) per step: join #1 + join #2 (Int64 keys only, no payload gather) + GROUP BY "
                  f"l_orderkey COUNT over the {joined.size} joined rows, P={threads} partitions each")
    else:
        print(json.dumps({"impl": "reference", "unavailable": f"no partition-parallel oracle driver for --workload {args.workload} (its CPU baseline is in the GPU arm's line)"}))
        return
    for _ in range(min(args.warmup, 1)):
        run()
    secs = [run()[0] for _ in range(args.steps)]
    s = statistics.mean(secs)
    val = rows_in / s
    metric = {"join": "input rows/s (hash join build+probe, device-resident input) and HBM GB/s",
              "q3": "input rows/s (Q3-shaped join + join + aggregation chain, device-resident input) and HBM GB/s"}.get(
                  args.workload, "input rows/s (hash aggregation, device-resident input) and HBM GB/s")
    print(json.dumps({
        "impl": "reference", "metric": metric, "value": val, "unit": "rows/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": s * 1e3,
        "higher_is_better": True, "scaling": "strong" if args.workload == "q3" else "weak", "vs_baseline": None, "dtype": dtype,
        "data": "synthetic", "config": workload_config(args, world),
        "cpu_baseline": {"value": val, "unit": "rows/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="c2", choices=["c2", "q1", "join", "q3", "c5", "c5s"])
    ap.add_argument("--rows", type=int, default=0, help="rows per GPU (q3: the TPC-H scale factor of the whole job)")
    ap.add_argument("--build-rows", type=int, default=1_460_000, help="join: rows of the build side")
    ap.add_argument("--cpu-sample", type=int, default=100_000_000, help="cpu_baseline: at most this many rows of rank 0's input")
    ap.add_argument("--clocks", default="nvml", choices=["nvml", "smi", "off"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-acero", action="store_true")
    ap.add_argument("--no-targets", action="store_true", help="skip the SF100 Q1 / Q3 target measurements of the default (c2) line")
    ap.add_argument("--targets-sf", type=int, default=100)
    ap.add_argument("--targets-steps", type=int, default=5)
    ap.add_argument("--batch-rows", type=int, default=0, help="e2e leg: push the host batch in slices of this many rows (8192 = the reference's batch_size)")
    ap.add_argument("--batch-api", default="one", choices=["one", "many"], help="with --batch-rows: one dfgpu_agg_push_arrow per slice, or all slices in one dfgpu_agg_push_arrow_batches call")
    ap.add_argument("--q1-keys", default="u8", choices=["u8", "utf8"], help="q1: key columns as 2 x UInt8 (default) or 2 x Utf8 as in configs[0]")
    ap.add_argument("--q1-project", action="store_true", help="q1: compute disc_price / charge on the device from price, discount, tax")
    ap.add_argument("--chunk-rows", type=int, default=0, help="dfgpu_options.chunk_rows (rows per internal launch wave; 0 = default)")
    ap.add_argument("--coalesce-rows", type=int, default=0, help="dfgpu_options.coalesce_rows (0 = default 1 Mi, -1 = off)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--three-op", action="store_true", help="N > 1: Partial -> partition -> all-to-all -> FinalPartitioned as three operators instead of the fused record exchange")
    args = ap.parse_args()
    explicit_rows = args.rows > 0
    if args.rows <= 0:
        args.rows = {"c2": 100_000_000, "q1": 60_012_150, "join": 32_400_000, "q3": 10, "c5": 100_000_000, "c5s": 100_000_000}[args.workload]   # q3: --rows = scale factor
    if args.warmup < 3 and args.impl != "reference":
        args.warmup = max(args.warmup, 3)       # timing rules: W >= 3
    if args.workload in ("c5", "c5s"):
        args.three_op = True     # string keys: the fused record exchange covers 64-bit key images only
    if args.impl == "reference":
        run_reference(args)
        return
    env = make_env(args)
    fn = {"join": run_join, "q3": run_q3}.get(args.workload, run_agg)
    line = fn(env, args, with_e2e=not args.no_e2e, with_cpu=not args.no_cpu)
    if args.workload == "c2" and not explicit_rows and not args.no_targets:
        import gc
        gc.collect()
        env.torch.cuda.empty_cache()
        line["targets"] = run_targets(env, args)
    if env.rank == 0:
        if line.get("parity", {}).get("checked") and line["parity"].get("ok") is False:
            line["parity_failed"] = True
        print(json.dumps(line))
    if env.world > 1:
        env.dist.barrier()
        env.dist.destroy_process_group()
    if env.rank == 0 and line.get("parity_failed"):
        sys.exit(3)


if __name__ == "__main__":
    main()
