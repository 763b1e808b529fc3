#!/usr/bin/env python3
"""Turns an ncu CSV (one row per launch and metric) into the per-kernel DRAM bytes per step that bench.py prints as
`roofline.traffic`.  Usage, on the GPU box (after the same command exited 0 without ncu):

    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/traffic_c2.csv python bench.py --workload c2 --steps 1 --warmup 3 --no-e2e --no-cpu --no-parity --no-targets
    python profiles/capture_traffic.py c2 gpurun_out/traffic_c2.csv 4 [more: q1 file steps ...]

`steps` = warm-up + timed steps of the profiled command (every step launches the same kernels).  The file records the
fingerprint of the CUDA sources (bench.csrc_sha); bench.py refuses to print traffic captured from other sources."""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def parse(path, steps):
    rows = list(csv.reader(l for l in open(path, errors="replace") if l.startswith('"')))
    hdr = rows[0]
    ik, im, iu, iv = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Unit"), hdr.index("Metric Value")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    tscale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}
    by, ms, launches = {}, {}, {}
    for r in rows[1:]:
        name = re.match(r"(?:void )?(?:dfgpu::)?([A-Za-z_0-9]+)", r[ik]).group(1)
        v = float(r[iv].replace(",", ""))
        if r[im].startswith("dram__bytes"):
            by[name] = by.get(name, 0.0) + v * scale.get(r[iu], 1)
            if r[im] == "dram__bytes_read.sum":
                launches[name] = launches.get(name, 0) + 1
        elif r[im] == "gpu__time_duration.sum":
            ms[name] = ms.get(name, 0.0) + v * tscale.get(r[iu], 1e-6)
    return ({k: int(v / steps) for k, v in by.items()}, {k: round(v / steps, 4) for k, v in ms.items()}, {k: v / steps for k, v in launches.items()})


def main():
    import bench
    out_path = os.path.join(ROOT, "profiles", "traffic.json")
    tj = json.load(open(out_path)) if os.path.exists(out_path) else {}
    sha = bench.csrc_sha()
    if tj.get("csrc_sha") != sha:
        tj = {"csrc_sha": sha}
    a = sys.argv[1:]
    for i in range(0, len(a), 3):
        wl, path, steps = a[i], a[i + 1], int(a[i + 2])
        kern, ms, launches = parse(path, steps)
        tj[wl] = {"kernels": kern, "ncu_ms_per_step": ms, "launches_per_step": launches,
                  "capture": {"file": os.path.basename(path), "steps": steps, "metrics": "dram__bytes_read.sum + dram__bytes_write.sum (cold-cache, serialised replays)"}}
    json.dump(tj, open(out_path, "w"), indent=1)
    print(json.dumps(tj, indent=1))


if __name__ == "__main__":
    main()
