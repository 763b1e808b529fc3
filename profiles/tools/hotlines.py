#!/usr/bin/env python3
"""Attribute executed SASS instructions of one kernel in an .ncu-rep to CUDA source lines.

    python profiles/tools/hotlines.py <report.ncu-rep> <kernel regex> <object.o> [top N] [mangled-name regex]

ncu's source page gives instruction counts per SASS address; nvdisasm -g on the cubin of the object file gives the source
line of every SASS offset (the library is built with -lineinfo).  Prints the N source lines that executed most warp
instructions, with the average number of active threads per instruction (divergence)."""
import csv
import os
import re
import subprocess
import sys
import tempfile


def main():
    rep, kre, obj = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
    mre = sys.argv[5] if len(sys.argv) > 5 else kre     # regex on the MANGLED name (template instantiations: k_fooILb1E)
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, check=True, capture_output=True)
    cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout.splitlines()
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = start = None
    for i, r in enumerate(rows):
        if "Source" in r and "Instructions Executed" in r:
            hdr, start = r, i + 1
            break
    idx = {h: i for i, h in enumerate(hdr)}
    sass = {}
    for r in rows[start:]:
        try:
            sass[int(r[idx["Address"]], 16)] = (float(r[idx["Instructions Executed"]]), float(r[idx["Thread Instructions Executed"]]), r[idx["Source"]])
        except Exception:
            pass
    base = min(sass)
    first = sass[base][2].split()[0]
    # the function in the disassembly: the section whose kernel name matches and whose first opcode agrees
    cur, seq, infn = None, [], False
    for l in dis:
        if l.startswith("//--------------------- .text."):
            infn = re.search(mre, l) is not None
            if infn and seq:
                break
            continue
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            seq.append((int(m.group(1), 16), cur))
    by, tot = {}, 0.0
    for off, cur in seq:
        if base + off in sass:
            ie, te, _ = sass[base + off]
            tot += ie
            v = by.setdefault(cur, [0.0, 0.0])
            v[0] += ie
            v[1] += te
    srcdir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "datafusion_upstream_b200", "csrc")
    cache = {}

    def line_of(k):
        if not k:
            return ""
        if k[0] not in cache:
            try:
                cache[k[0]] = open(os.path.join(srcdir, k[0])).read().splitlines()
            except Exception:
                cache[k[0]] = []
        ls = cache[k[0]]
        return ls[k[1] - 1].strip() if 0 < k[1] <= len(ls) else ""
    print(f"kernel {kre}: {tot:.0f} warp instructions attributed ({first} ...)")
    for k, v in sorted(by.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{v[0] / tot * 100:5.1f}%  thr/inst={v[1] / max(v[0], 1):4.1f}  {k[0] if k else '?'}:{k[1] if k else 0:<4d} {line_of(k)[:110]}")


if __name__ == "__main__":
    main()
