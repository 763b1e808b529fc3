# A/B of the one-byte-key form of the register-mode key phase, Q1 at SF100 on one box (UInt8 and Utf8 keys)
for keys in u8 utf8; do for nok8 in 1 0; do
  DFGPU_FA_NOK8=$nok8 timeout 300 python bench.py --workload q1 --q1-keys $keys --rows 600037900 --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2u_q1_${keys}_nok8_$nok8.json 2> gpurun_out/r2u_q1_${keys}_nok8_$nok8.err
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2u_q1_${keys}_nok8_$nok8.json") if l.startswith("{")][0])
k=d["roofline"]["kernels"]
print("$keys", "general_keys=$nok8", "ms_per_step", round(d["ms_per_step"],3), "kernel_ms", round(sum(k.values()),3), k, d["parity"]["ok"])
PY
done; done
