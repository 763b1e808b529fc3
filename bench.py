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
# workloads
# ------------------------------------------------------------------------------------------------
def build_workload(args, df, ctx, torch, device, rank):
    """Returns dict(schema, dev_batch, host_batch_factory, make_op, rows, algo_bytes_fn, name, cpu)."""
    import oracle
    w = {}
    n = args.rows
    if args.workload == "c2":
        keys = gen_c2_keys(n, 42 + rank, device)
        schema = pa.schema([("UserID", pa.int64())])
        w["name"] = f"ClickBench-shaped GROUP BY UserID COUNT(*), {n} Int64 rows, Zipf(1.0) over 85M ids (BASELINE configs[1])"
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
    elif args.workload == "q1":
        flag, status, cols = gen_q1(n, 42 + rank, device)
        d152, d384, d386 = pa.decimal128(15, 2), pa.decimal128(38, 4), pa.decimal128(38, 6)
        schema = pa.schema([("l_returnflag", pa.uint8()), ("l_linestatus", pa.uint8()), ("l_quantity", d152), ("l_extendedprice", d152),
                            ("l_discount", d152), ("disc_price", d384), ("charge", d386)])
        w["name"] = f"TPC-H Q1-shaped aggregation, {n} rows, 4 groups, 2xUInt8 keys, Decimal128 SUMx4 AVGx3 COUNT(*) (82 B/row variant of configs[0])"
        if args.q1_keys == "utf8":
            # configs[0] as written: l_returnflag / l_linestatus are Utf8 (one byte each) - the same bytes as the UInt8 variant
            # behind i32 offsets 0..n (90 B/row)
            offs = torch.arange(0, n + 1, device=device, dtype=torch.int32)
            schema = pa.schema([pa.field("l_returnflag", pa.string(), nullable=False), pa.field("l_linestatus", pa.string(), nullable=False)] + list(schema)[2:])
            w["name"] = w["name"].replace("2xUInt8 keys", "2xUtf8 keys (1 byte each)").replace("82 B/row variant of configs[0]", "configs[0] shape, 90 B/row")
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
            w["name"] += " + projections disc_price / charge evaluated on the device (66 B/row in)"
        kcols = [("l_returnflag", pa.string(), (offs, flag)), ("l_linestatus", pa.string(), (offs, status))] if args.q1_keys == "utf8" else \
                [("l_returnflag", pa.uint8(), flag), ("l_linestatus", pa.uint8(), status)]
        host_cols = kcols + \
                    [(nm, t, cols[c]) for (nm, t), c in zip(list(zip(schema.names, schema.types))[2:], order)]

        def cpu(threads, sample):
            f, s = flag[:sample].cpu().numpy(), status[:sample].cpu().numpy()
            d = [cols[c][:sample].cpu().numpy() for c in ("qty", "price", "disc_price", "charge", "disc")]
            return oracle.bench_q1(f, s, d, threads)
        w["cpu"] = cpu
    elif args.workload in ("c5", "c5s"):
        region, offsets, data = gen_c5(n, 42 + rank, device)
        schema = pa.schema([pa.field("RegionID", pa.int32(), nullable=False), pa.field("SearchPhrase", pa.string(), nullable=False)])   # ClickBench hits: NOT NULL
        w["name"] = (f"ClickBench-shaped GROUP BY RegionID, SearchPhrase COUNT(*), {n} rows, Int32 x Utf8 keys (70% empty phrase, Zipf over "
                     f"6M phrases of 1-80 bytes, {int(data.numel())} string bytes) - variable-width GroupValues interning (BASELINE configs[4])")
        w["dev"] = df.DeviceBatch.from_device(ctx, schema, [{"buffers": [None, region.data_ptr()], "null_count": 0},
                                                            {"buffers": [None, offsets.data_ptr(), data.data_ptr()], "null_count": 0}],
                                              n, keep=(region, offsets, data))
        w["schema"] = schema
        w["group_by"] = ["RegionID", "SearchPhrase"]
        w["exprs"] = [df.AggregateExpr(df.AggFunc.Count, None, "COUNT(*)")]
        w["in_bytes"] = 4 * n + 4 * (n + 1) + int(data.numel())
        if args.workload == "c5s":      # the single-key variant: GROUP BY SearchPhrase (GroupValuesByes, bytes.rs:44-121)
            w["name"] = w["name"].replace("GROUP BY RegionID, SearchPhrase", "GROUP BY SearchPhrase").replace("Int32 x Utf8 keys", "one Utf8 key")
            w["group_by"] = ["SearchPhrase"]
            w["in_bytes"] -= 4 * n
        w["out_row_bytes"] = 4 + 4 + 8          # RegionID + phrase offset + count (the distinct phrases' own bytes are not counted)
        host_cols = [("RegionID", pa.int32(), region), ("SearchPhrase", pa.string(), (offsets, data))]

        def cpu(threads, sample):
            # the oracle's generic operator (GroupValuesRows, row.rs:94-146), single-threaded, 8192-row batches
            import time as _t
            hb = pa.record_batch([pa.array(region[:sample].cpu().numpy()),
                                  pa.Array.from_buffers(pa.string(), sample, [None, pa.py_buffer(offsets[:sample + 1].cpu().numpy()),
                                                                             pa.py_buffer(data.cpu().numpy())])], schema=schema)
            o = oracle.OracleAgg("single", ["RegionID", "SearchPhrase"], [pa.int32(), pa.string()], [oracle.AggSpec("count", "COUNT(*)", None)])
            t0 = _t.perf_counter()
            for off in range(0, sample, 8192):
                s = hb.slice(off, 8192)
                o.push([s.column(0), s.column(1)], [None])
            out = o.emit()
            return _t.perf_counter() - t0, out.num_rows, 0
        w["cpu"] = cpu
        w["cpu_threads"] = 1
        w["cpu_sample"] = 4_000_000
    else:
        raise SystemExit(f"unknown workload {args.workload}")
    w["host_cols"] = host_cols
    return w


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


def run_ours(args):
    import torch
    import torch.distributed as dist
    import datafusion_upstream_b200 as df

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group(backend="gloo")     # bootstrap / barrier only; the data path uses NCCL inside libdfgpu
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    uid = None
    if world > 1:
        box = [df.GpuContext.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]
    with native_stdout_to_stderr():
        ctx = df.GpuContext(device=local, rank=rank, world=world, nccl_unique_id=uid, coalesce_rows=args.coalesce_rows, chunk_rows=args.chunk_rows)
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)
    w = build_workload(args, df, ctx, torch, device, rank)
    n = args.rows
    torch.cuda.synchronize()

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    hold = []

    def step(inp, to_host=False):
        """One pass of the hot path over one batch.  inp: DeviceBatch (inputs resident in HBM) or a pinned
        host RecordBatch (H2D inside the step); to_host: read the result back into host Arrow buffers."""
        if w.get("prepare"):
            inp = w["prepare"](inp if isinstance(inp, df.DeviceBatch) else df.DeviceBatch.from_arrow(ctx, inp))
        if world == 1:
            op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, w["schema"], w["group_by"], w["exprs"])
            op.push(inp)
            out = op.emit() if to_host else op.emit_device()
            hold[:] = [(op, out)]
            return out
        if not args.three_op:
            # fused: level 1 on the local shard -> record exchange (NCCL, one contiguous run per peer) -> level 2 on the owned partitions
            op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, w["schema"], w["group_by"], w["exprs"])
            op.push(inp)
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

    def step_device():
        return step(w["dev"])

    for _ in range(args.warmup):
        out = step_device()
        groups = out.num_rows
        hold.clear()
    barrier()
    ctx.set_profiling(True)
    launches0 = ctx.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clk = Clocks(local, args.clocks)
    host_ms = []
    with clk:
        barrier()
        e0.record(stream)
        for _ in range(args.steps):
            t0 = time.perf_counter()
            out = step_device()
            host_ms.append(round((time.perf_counter() - t0) * 1e3, 2))
        e1.record(stream)
        barrier()
    ms = e0.elapsed_time(e1)
    groups = out.num_rows
    launches = ctx.kernel_launches - launches0
    prof = ctx.profile()
    ctx.set_profiling(False)
    hold.clear()
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        gt = torch.tensor([groups], dtype=torch.int64)
        dist.all_reduce(gt)
        groups = int(gt[0])
    ms_per_step = ms / args.steps
    value = n * world / (ms_per_step / 1e3)

    # ---- e2e: host (pinned) Arrow buffers through the C-ABI, H2D + D2H inside the timed region ----
    e2e = None
    if not args.no_e2e:
        hb, keep, h2d = pinned_batch(ctx, w.get("host_schema", w["schema"]), w["host_cols"], n, torch)
        e2e_steps = max(1, min(args.steps, 5))
        exported = []
        if args.batch_rows > 0 and world == 1:
            # the reference's batch granularity (8192 rows, common/src/config.rs:215): one dfgpu_agg_push_arrow per slice of
            # the pinned host batch; the C-Data structs are exported before the timed region (a Rust caller has them for free)
            from datafusion_upstream_b200 import _native as N
            for o in range(0, n, args.batch_rows):
                a, sc = N.ArrowArray(), N.ArrowSchema()
                hb.slice(o, args.batch_rows)._export_to_c(ctypes.addressof(a), ctypes.addressof(sc))
                exported.append((a, sc))
            L = N.lib()

            def step(inp, to_host=True):  # noqa: F811 - e2e leg only
                op = df.GpuAggregateExec(ctx, df.AggregateMode.Single, w["schema"], w["group_by"], w["exprs"])
                for a, _ in exported:
                    N.check(L.dfgpu_agg_push_arrow(op.h, ctypes.addressof(a)))
                out = op.emit()
                hold[:] = [(op, out)]
                return out
        step(hb, to_host=True)
        hold.clear()
        barrier()
        t0 = time.perf_counter()
        res = None
        for _ in range(e2e_steps):
            res = None
            hold.clear()            # the consumer drops the previous result: its pinned buffers return to the pool
            res = step(hb, to_host=True)
        barrier()
        dt = (time.perf_counter() - t0) / e2e_steps
        d2h = int(res.nbytes)
        hold.clear()
        if world > 1:
            t = torch.tensor([dt, float(h2d), float(d2h)], dtype=torch.float64)
            tmax = t.clone()
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(t)
            dt, h2d, d2h = float(tmax[0]), int(t[1]), int(t[2])
        e2e = {"value": n * world / dt, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": dt * 1e3}
        if exported:
            e2e["batch_rows"] = args.batch_rows
            e2e["pushes_per_step"] = len(exported)
            for a, sc in exported:
                if a.release:
                    ctypes.CFUNCTYPE(None, ctypes.POINTER(N.ArrowArray))(a.release)(ctypes.byref(a))
                N.release_schema(sc)
        del hb
        for p in keep:
            ctx.host_free(p)

    # ---- roofline of the dominant kernel (CUDA events recorded live by the library) ----
    peak, peak_src = peaks()
    roof = None
    stages = {r["name"][6:]: round(r["ms"] / args.steps, 4) for r in prof if r["name"].startswith("stage:")} if prof else {}
    prof = [r for r in prof if not r["name"].startswith("stage:")] if prof else prof
    if prof:
        top = max(prof, key=lambda r: r["ms"])
        per_launch_ms = top["ms"] / top["launches"]
        achieved = top["algo_bytes"] / top["launches"] / (per_launch_ms * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get(args.workload, {}).get(top["name"])
        roof = {"bound": "hbm", "kernel": top["name"], "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
                "launches_per_step": top["launches"] / args.steps, "ms_per_launch": round(per_launch_ms, 4),
                "kernel_share_of_step": round(top["ms"] / ms, 3),
                "kernels": {r["name"]: round(r["ms"] / args.steps, 4) for r in prof}, "stages": stages}
    algo_bytes = w["in_bytes"] + w["out_row_bytes"] * (groups // world)
    # ---- CPU baseline (the oracle port, reference-shaped parallel plan) on rank 0 ----
    cpu = None
    if rank == 0 and not args.no_cpu:
        threads = w.get("cpu_threads") or os.cpu_count() or 1
        sample = min(n, w.get("cpu_sample") or args.cpu_sample)
        secs, g, _ = w["cpu"](threads, sample)
        shape = "P=%d Partial -> hash repartition -> FinalPartitioned" % threads if threads > 1 else "one Single-mode operator (scalar port)"
        cpu = {"value": sample / secs, "unit": "rows/s", "cores": threads, "kind": "port",
               "sample": f"first {sample} rows of the same workload, {shape}, 8192-row batches"}
        if w.get("acero"):
            # a second, independent CPU data point: pyarrow's Acero hash aggregation on the same Arrow columns
            try:
                cpu["acero"] = w["acero"](min(n, 20_000_000))
            except Exception as e:      # the bench line must not depend on it
                cpu["acero"] = {"unavailable": str(e)[:120]}
    if rank == 0:
        line = {
            "metric": "input rows/s (hash aggregation, device-resident input) and HBM GB/s", "value": value, "unit": "rows/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": {"c2": "int64", "c5": "u8", "c5s": "u8"}.get(args.workload, "i128"), "data": "synthetic",
            "config": {"workload": w["name"], "rows_per_gpu": n, "groups": groups, "l2_policy": "inputs larger than L2 (no flush needed)"
                       if w["in_bytes"] > 256e6 else "inputs smaller than L2: table/state reset each step", "mode": "Single" if world == 1 else
                       (f"Partial -> NCCL all-to-all -> FinalPartitioned (dp{world})" if args.three_op else
                       f"fused Partial -> NCCL record exchange -> FinalPartitioned in one operator per rank (dp{world})")},
            "algorithmic_GBps": algo_bytes * world / (ms_per_step * 1e-3) / 1e9,
            "algorithmic_frac_of_peak": algo_bytes / (ms_per_step * 1e-3) / 1e9 / peak,
            "host_submit_ms": host_ms, "clocks": clk.result, "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu,
        }
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def gen_join(n_build, n_probe, seed, device, rank=0, world=1):
    """Q3-shaped join #2 at SF10 (post-filter): build = (o_orderkey, o_orderdate, o_shippriority) of the orders that
    survived join #1, probe = (l_orderkey, l_extendedprice, l_discount) of the filtered lineitem rows.  Order keys
    are sparse like TPC-H's (8 of every 32); ~10 % of the probe rows find a build row."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    n_orders = 15_000_000
    allk = torch.arange(0, n_orders, device=device, dtype=torch.int64)
    allk = (allk // 8) * 32 + (allk % 8) + 1
    perm = torch.randperm(n_orders, generator=g, device=device)
    bkeys = allk[perm[:n_build]].contiguous()
    odate = torch.randint(8035, 9204, (n_build,), generator=g, device=device, dtype=torch.int32)
    oprio = torch.zeros(n_build, device=device, dtype=torch.int32)
    gp = torch.Generator(device=device)
    gp.manual_seed(seed + 1000 + rank)
    pk = allk[torch.randint(0, n_orders, (n_probe,), generator=gp, device=device)].contiguous()

    def dec(lo, hi):
        v = torch.randint(lo, hi, (n_probe,), generator=gp, device=device, dtype=torch.int64)
        return torch.stack([v, v >> 63], dim=1).contiguous()
    return (bkeys, odate, oprio), (pk, dec(90000, 10495001), dec(0, 11))


def run_join(args):
    """--workload join: HashJoinExec build + probe (Inner, CollectLeft / broadcast build) on the Q3-shaped join."""
    import torch
    import torch.distributed as dist
    import datafusion_upstream_b200 as df
    import oracle
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group(backend="gloo")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    from datafusion_upstream_b200.dist import bootstrap_nccl_id, shard_range
    uid = bootstrap_nccl_id(dist, df.GpuContext.nccl_unique_id) if world > 1 else None
    with native_stdout_to_stderr():
        ctx = df.GpuContext(device=local, rank=rank, world=world, nccl_unique_id=uid)
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)
    B, P = 1_460_000, args.rows
    bt, pt = gen_join(B, P, 42, device, rank, world)
    bs = pa.schema([("o_orderkey", pa.int64()), ("o_orderdate", pa.date32()), ("o_shippriority", pa.int32())])
    ps = pa.schema([("l_orderkey", pa.int64()), ("l_extendedprice", pa.decimal128(15, 2)), ("l_discount", pa.decimal128(15, 2))])
    lo, hi = shard_range(B, rank, world)           # every rank holds a row range of the build side; it is all-gathered per step
    bfull = df.DeviceBatch.from_device(ctx, bs, [{"buffers": [None, t.data_ptr()]} for t in bt], B, keep=bt)
    bdev = bfull.slice(lo, hi - lo)
    pdev = df.DeviceBatch.from_device(ctx, ps, [{"buffers": [None, t.data_ptr()]} for t in pt], P, keep=pt)
    host_b = [(f.name, f.type, t[lo:hi]) for f, t in zip(bs, bt)]
    host_p = [(f.name, f.type, t) for f, t in zip(ps, pt)]
    hold = []

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def step(b, p, to_host=False):
        j = df.GpuHashJoinExec(ctx, bs, ps, [("o_orderkey", "l_orderkey")], df.JoinType.Inner, df.PartitionMode.CollectLeft)
        bb = b if isinstance(b, df.DeviceBatch) else df.DeviceBatch.from_arrow(ctx, b)
        if world > 1:
            bb = df.all_gather(ctx, bs, bb)        # broadcast build: the probe side never moves
        j.build([bb])
        out = j.probe(p) if to_host else j.probe_device(p)
        hold[:] = [(j, bb, out)]
        return out

    for _ in range(args.warmup):
        out = step(bdev, pdev)
    barrier()
    ctx.set_profiling(True)
    l0 = ctx.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clk = Clocks(local, args.clocks)
    with clk:
        barrier()
        e0.record(stream)
        for _ in range(args.steps):
            out = step(bdev, pdev)
        e1.record(stream)
        barrier()
    ms = e0.elapsed_time(e1)
    out_rows = out.num_rows
    launches = ctx.kernel_launches - l0
    prof = ctx.profile()
    ctx.set_profiling(False)
    hold.clear()
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        ot = torch.tensor([out_rows], dtype=torch.int64)
        dist.all_reduce(ot)
        out_rows = int(ot[0])
    ms_per_step = ms / args.steps
    rows_in = B + P * world
    value = rows_in / (ms_per_step * 1e-3)
    algo = B * 16 + P * world * 40 + out_rows * 56
    e2e = None
    if not args.no_e2e:
        hb, keep_b, h2d_b = pinned_batch(ctx, bs, host_b, hi - lo, torch)
        hp, keep_p, h2d_p = pinned_batch(ctx, ps, host_p, P, torch)
        n_e2e = max(1, min(args.steps, 5))
        step(hb, hp, True)
        hold.clear()
        barrier()
        t0 = time.perf_counter()
        res = None
        for _ in range(n_e2e):
            res = None
            hold.clear()
            res = step(hb, hp, True)
        barrier()
        dt = (time.perf_counter() - t0) / n_e2e
        h2d, d2h = h2d_b + h2d_p, int(res.nbytes)
        hold.clear()
        if world > 1:
            t = torch.tensor([dt, float(h2d), float(d2h)], dtype=torch.float64)
            tm = t.clone()
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            dist.all_reduce(t)
            dt, h2d, d2h = float(tm[0]), int(t[1]), int(t[2])
        e2e = {"value": rows_in / dt, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": dt * 1e3}
        del hb, hp, res
        for p_ in keep_b + keep_p:
            ctx.host_free(p_)
    peak, peak_src = peaks()
    roof = None
    stages = {r["name"][6:]: round(r["ms"] / args.steps, 4) for r in prof if r["name"].startswith("stage:")} if prof else {}
    prof = [r for r in prof if not r["name"].startswith("stage:")] if prof else prof
    if prof:
        top = max(prof, key=lambda r: r["ms"])
        per = top["ms"] / top["launches"]
        ach = top["algo_bytes"] / top["launches"] / (per * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": top["name"], "achieved": round(ach, 1), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                "traffic": None, "peak_source": peak_src, "launches_per_step": top["launches"] / args.steps, "ms_per_launch": round(per, 4),
                "kernel_share_of_step": round(top["ms"] / ms, 3), "kernels": {r["name"]: round(r["ms"] / args.steps, 4) for r in prof}, "stages": stages}
    cpu = None
    if rank == 0 and not args.no_cpu:
        threads = os.cpu_count() or 1
        sample = min(P, args.cpu_sample)
        secs, r_, _ = oracle.bench_join_i64(bt[0].cpu().numpy(), pt[0][:sample].cpu().numpy(), threads)
        cpu = {"value": (B + sample) / secs, "unit": "rows/s", "cores": threads, "kind": "port",
               "sample": f"build {B} + first {sample} probe keys (Int64 keys only), P={threads} Partitioned hash joins, 8192-row batches"}
    if rank == 0:
        print(json.dumps({
            "metric": "input rows/s (hash join build+probe, device-resident input) and HBM GB/s", "value": value, "unit": "rows/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": f"TPC-H Q3-shaped join #2 at SF10 (post-filter): build {B} x 16 B, probe {P} x 40 B per GPU, Inner, "
                                   f"{'CollectLeft' if world == 1 else 'broadcast build (NCCL all-gather)'} (BASELINE configs[2])",
                       "build_rows": B, "probe_rows_per_gpu": P, "output_rows": out_rows, "l2_policy": "inputs larger than L2 (no flush needed)"},
            "algorithmic_GBps": algo / (ms_per_step * 1e-3) / 1e9, "algorithmic_frac_of_peak": algo / world / (ms_per_step * 1e-3) / 1e9 / peak,
            "clocks": clk.result, "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def gen_q3(sf, seed, device, rank=0, world=1):
    """Post-filter inputs of the three scans of TPC-H Q3 at scale factor `sf`, generated on the device:
    customer (c_custkey; c_mktsegment keeps 1 in 5), orders (o_orderkey sparse 8-of-32, o_custkey, o_orderdate < 1995-03-15,
    o_shippriority = 0), lineitem (l_orderkey, l_extendedprice, l_discount; l_shipdate = o_orderdate + 1..121 > 1995-03-15).
    With world > 1 the lineitem rows are this rank's shard (its own seed); customer and orders are replicated."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    n_cust, n_orders, n_line = 150_000 * sf, 1_500_000 * sf, 6_000_000 * sf
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


def run_q3(args):
    """--workload q3: the whole Q3-shaped operator chain (BASELINE configs[2]): join #1 customer x orders, join #2 x lineitem,
    then GROUP BY (l_orderkey, o_orderdate, o_shippriority) SUM/COUNT - Single on one GPU, Partial -> NCCL all-to-all ->
    FinalPartitioned when sharded (lineitem sharded, customer / orders replicated = CollectLeft builds)."""
    import torch
    import torch.distributed as dist
    import datafusion_upstream_b200 as df
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group(backend="gloo")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    from datafusion_upstream_b200.dist import bootstrap_nccl_id
    uid = bootstrap_nccl_id(dist, df.GpuContext.nccl_unique_id) if world > 1 else None
    with native_stdout_to_stderr():
        ctx = df.GpuContext(device=local, rank=rank, world=world, nccl_unique_id=uid)
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)
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

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def step(c, o, l, to_host=False):
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

    for _ in range(args.warmup):
        out = step(cdev, odev, ldev)
    barrier()
    ctx.set_profiling(True)
    l0 = ctx.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clk = Clocks(local, args.clocks)
    with clk:
        barrier()
        e0.record(stream)
        for _ in range(args.steps):
            out = step(cdev, odev, ldev)
        e1.record(stream)
        barrier()
    ms = e0.elapsed_time(e1)
    groups = out.num_rows
    launches = ctx.kernel_launches - l0
    prof = ctx.profile()
    ctx.set_profiling(False)
    hold.clear()
    rows_local = nC + nO + nL
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        gt = torch.tensor([groups, nL, stats["j2"]], dtype=torch.int64)
        dist.all_reduce(gt)
        groups, nL_all, j2_all = int(gt[0]), int(gt[1]), int(gt[2])
    else:
        nL_all, j2_all = nL, stats["j2"]
    ms_per_step = ms / args.steps
    rows_in = (nC + nO) * world + nL_all          # every rank reads its replicated build sides
    value = rows_in / (ms_per_step * 1e-3)
    # consumed input columns once + output rows once, per operator
    algo = world * (nC * 8 + nO * 24 + stats["j1"] * 16) + (world * stats["j1"] * 16 + nL_all * 40 + j2_all * 56) + (j2_all * 48 + groups * 56)
    e2e = None
    if not args.no_e2e:
        hc, kc, b1 = pinned_batch(ctx, cs, [(f.name, f.type, t) for f, t in zip(cs, ct)], nC, torch)
        ho, ko, b2 = pinned_batch(ctx, os_, [(f.name, f.type, t) for f, t in zip(os_, ot)], nO, torch)
        hl, kl, b3 = pinned_batch(ctx, ls, [(f.name, f.type, t) for f, t in zip(ls, lt)], nL, torch)
        n_e2e = max(1, min(args.steps, 5))
        step(hc, ho, hl, True)
        hold.clear()
        barrier()
        t0 = time.perf_counter()
        res = None
        for _ in range(n_e2e):
            res = None
            hold.clear()
            res = step(hc, ho, hl, True)
        barrier()
        dt = (time.perf_counter() - t0) / n_e2e
        h2d, d2h = b1 + b2 + b3, int(res.nbytes)
        hold.clear()
        if world > 1:
            t = torch.tensor([dt, float(h2d), float(d2h)], dtype=torch.float64)
            tm = t.clone()
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            dist.all_reduce(t)
            dt, h2d, d2h = float(tm[0]), int(t[1]), int(t[2])
        e2e = {"value": rows_in / dt, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": dt * 1e3}
        del hc, ho, hl, res
        for p_ in kc + ko + kl:
            ctx.host_free(p_)
    peak, peak_src = peaks()
    roof = None
    stages = {r["name"][6:]: round(r["ms"] / args.steps, 4) for r in prof if r["name"].startswith("stage:")} if prof else {}
    prof = [r for r in prof if not r["name"].startswith("stage:")] if prof else prof
    if prof:
        top = max(prof, key=lambda r: r["ms"])
        per = top["ms"] / top["launches"]
        ach = top["algo_bytes"] / top["launches"] / (per * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": top["name"], "achieved": round(ach, 1), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                "traffic": None, "peak_source": peak_src, "launches_per_step": top["launches"] / args.steps, "ms_per_launch": round(per, 4),
                "kernel_share_of_step": round(top["ms"] / ms, 3), "kernels": {r["name"]: round(r["ms"] / args.steps, 4) for r in prof}, "stages": stages}
    if rank == 0:
        print(json.dumps({
            "metric": "input rows/s (Q3-shaped join + join + aggregation chain, device-resident input) and HBM GB/s", "value": value, "unit": "rows/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "i128", "data": "synthetic",
            "config": {"workload": f"TPC-H Q3-shaped chain at SF{sf} (post-filter scans): customer {nC} x orders {nO} -> {stats['j1']} rows, x lineitem "
                                   f"{nL} per GPU -> {j2_all} rows, GROUP BY (l_orderkey, o_orderdate, o_shippriority) -> {groups} groups (BASELINE configs[2])",
                       "customer_rows": nC, "orders_rows": nO, "lineitem_rows_per_gpu": nL, "join1_rows": stats["j1"], "join2_rows": j2_all, "groups": groups,
                       "l2_policy": "inputs larger than L2 (no flush needed)",
                       "mode": "Single" if world == 1 else f"CollectLeft builds replicated, lineitem sharded, Partial -> NCCL all-to-all -> FinalPartitioned (dp{world})"},
            "algorithmic_GBps": algo / (ms_per_step * 1e-3) / 1e9, "algorithmic_frac_of_peak": algo / world / (ms_per_step * 1e-3) / 1e9 / peak,
            "clocks": clk.result, "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": None}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()



def run_reference(args):
    """The reference's own CPU implementation of the path, on the box's host cores.  The reference is
    Rust and cannot be built here (no cargo/rustc, crates not vendored: DESIGN.md), so this arm times
    the oracle port in the reference's parallel shape (P x Partial -> hash repartition -> P x Final)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.workload not in ("c2", "q1"):
        # the headline workload (c2) and q1 have a partition-parallel oracle driver; the other workloads report their CPU
        # baseline inside their own bench line (cpu_baseline)
        print(json.dumps({"impl": "reference", "unavailable": f"reference arm covers --workload c2 and q1, not {args.workload}"}))
        return
    import oracle
    threads = os.cpu_count() or 1
    n = min(args.rows, args.cpu_sample)
    rng = np.random.default_rng(42)
    if args.workload == "c2":
        u = rng.random(n)
        rank_ = np.exp(u * math.log(85_000_000)).astype(np.int64)
        with np.errstate(over="ignore"):
            x = rank_.astype(np.uint64)
            x ^= x >> np.uint64(30); x *= np.uint64(0xbf58476d1ce4e5b9)
            x ^= x >> np.uint64(27); x *= np.uint64(0x94d049bb133111eb)
            x ^= x >> np.uint64(31)
        keys = x.view(np.int64)
        run = lambda: oracle.bench_groupby_count_i64(keys, threads)
        name = f"ClickBench-shaped GROUP BY UserID COUNT(*), {args.rows} Int64 rows, Zipf(1.0) over 85M ids (BASELINE configs[1])"
    else:
        pick = rng.choice(4, n, p=[0.25, 0.006, 0.493, 0.251])
        flag = np.array([65, 78, 78, 82], np.uint8)[pick]
        status = np.array([70, 70, 79, 70], np.uint8)[pick]

        def dec(lo, hi):
            v = rng.integers(lo, hi, n)
            return np.stack([v, v >> 63], axis=1).copy()
        d = [dec(100, 5001), dec(90000, 10495001), dec(10**9, 10**11), dec(10**11, 10**13), dec(0, 11)]
        run = lambda: oracle.bench_q1(flag, status, d, threads)
        name = f"TPC-H Q1-shaped aggregation, {args.rows} rows"
    for _ in range(min(args.warmup, 1)):
        run()
    secs = [run()[0] for _ in range(args.steps)]
    s = statistics.mean(secs)
    val = n / s
    sample = f"{n} rows of the workload per step, P={threads} threads, 8192-row batches"
    print(json.dumps({
        "impl": "reference", "metric": "input rows/s (hash aggregation, device-resident input) and HBM GB/s", "value": val, "unit": "rows/s",
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup, "ms_per_step": s * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64" if args.workload == "c2" else "i128",
        "data": "synthetic", "config": {"workload": name},
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
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--cpu-sample", type=int, default=50_000_000)
    ap.add_argument("--clocks", default="nvml", choices=["nvml", "smi", "off"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--batch-rows", type=int, default=0, help="e2e leg: push the host batch in slices of this many rows (8192 = the reference's batch_size)")
    ap.add_argument("--q1-keys", default="u8", choices=["u8", "utf8"], help="q1: key columns as 2 x UInt8 (default) or 2 x Utf8 as in configs[0]")
    ap.add_argument("--q1-project", action="store_true", help="q1: compute disc_price / charge on the device from price, discount, tax")
    ap.add_argument("--chunk-rows", type=int, default=0, help="dfgpu_options.chunk_rows (rows per internal launch wave; 0 = default)")
    ap.add_argument("--coalesce-rows", type=int, default=0, help="dfgpu_options.coalesce_rows (0 = default 1 Mi, -1 = off)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--three-op", action="store_true", help="N > 1: Partial -> partition -> all-to-all -> FinalPartitioned as three operators instead of the fused record exchange")
    args = ap.parse_args()
    if args.rows <= 0:
        args.rows = {"c2": 100_000_000, "q1": 60_012_150, "join": 32_400_000, "q3": 10, "c5": 100_000_000, "c5s": 100_000_000}[args.workload]   # q3: --rows = scale factor
    if args.warmup < 3 and args.impl != "reference":
        args.warmup = max(args.warmup, 3)       # timing rules: W >= 3
    if args.workload in ("c5", "c5s"):
        args.three_op = True     # string keys: the fused record exchange covers 64-bit key images only
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "join":
        run_join(args)
    elif args.workload == "q3":
        run_q3(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
