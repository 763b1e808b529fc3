set -x
timeout 600 python -m pytest tests/test_gpu_join.py tests/test_gpu_q3.py -m gpu -q -x 2>&1 | tail -2
timeout 300 python bench.py --workload join --rows 324000000 --build-rows 14600000 --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_final_join_sf100.json 2>/dev/null
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2_final_join_sf100.json") if l.startswith("{")][0])
print(d["ms_per_step"], d["roofline"]["kernels"], d["parity"]["ok"])
PY
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_final_traffic_c2.csv python bench.py --workload c2 --steps 1 --warmup 3 --no-e2e --no-cpu --no-parity --no-targets > /dev/null 2>&1; echo ncu-traffic rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_fa_l1_reg" --launch-skip 3 --launch-count 1 -o gpurun_out/r2_final_q1 -f python bench.py --workload q1 --rows 134217728 --steps 1 --warmup 3 --no-e2e --no-cpu --no-parity > /dev/null 2>&1; echo ncu-q1 rc=$?
