# A/B of the rows per level-1 launch for Q1 at SF100, one box
for c in 33554432 67108864 134217728 268435456; do
  timeout 300 python bench.py --workload q1 --rows 600037900 --steps 5 --warmup 3 --no-e2e --no-cpu --chunk-rows $c > gpurun_out/r2s_q1_chunk$c.json 2> gpurun_out/r2s_q1_chunk$c.err
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2s_q1_chunk$c.json") if l.startswith("{")][0])
k=d["roofline"]["kernels"]
print($c, round(d["ms_per_step"],3), "kernels", round(sum(k.values()),3), k, d["parity"]["ok"])
PY
done
