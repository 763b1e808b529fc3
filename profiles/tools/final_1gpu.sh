# end-of-round evidence on one GPU: tests, smoke, bench lines, launch list, DRAM traffic, ncu summaries
set -x
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2_final_pytest.log 2>&1; echo pytest rc=$?; tail -2 gpurun_out/r2_final_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r2_final_bench_default.json 2> gpurun_out/r2_final_bench_default.err; echo bench rc=$?
timeout 900 python bench.py --impl reference > gpurun_out/r2_final_bench_reference.json 2> gpurun_out/r2_final_bench_reference.err; echo ref rc=$?
timeout 300 python bench.py --workload join --steps 10 --warmup 3 > gpurun_out/r2_final_bench_join.json 2>/dev/null
timeout 300 python bench.py --workload q3 --steps 10 --warmup 3 > gpurun_out/r2_final_bench_q3.json 2>/dev/null
timeout 300 python bench.py --workload q1 --steps 10 --warmup 3 > gpurun_out/r2_final_bench_q1.json 2>/dev/null
# launch list + DRAM bytes per launch of the C2 step (cold-cache, serialised)
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_final_traffic_c2.csv python bench.py --workload c2 --steps 1 --warmup 3 --no-e2e --no-cpu --no-parity --no-targets > /dev/null 2>&1; echo ncu-traffic rc=$?
python profiles/capture_traffic.py c2 gpurun_out/r2_final_traffic_c2.csv 4 > /dev/null; echo capture rc=$?
# full sets for the kernels the docs quote
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_hp_|k_fa_scan" --launch-skip 24 --launch-count 8 -o gpurun_out/r2_final_c2 -f python bench.py --workload c2 --steps 1 --warmup 3 --no-e2e --no-cpu --no-parity --no-targets > /dev/null 2>&1; echo ncu-c2 rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_fa_l1_reg|k_strlen" --launch-skip 3 --launch-count 2 -o gpurun_out/r2_final_q1 -f python bench.py --workload q1 --q1-keys utf8 --rows 134217728 --steps 1 --warmup 3 --no-e2e --no-cpu --no-parity > /dev/null 2>&1; echo ncu-q1 rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_jbp|k_jp_|k_take_multi|k_pack_rows" --launch-skip 21 --launch-count 7 -o gpurun_out/r2_final_join -f python bench.py --workload join --rows 324000000 --build-rows 14600000 --steps 1 --warmup 3 --no-e2e --no-cpu --no-parity > /dev/null 2>&1; echo ncu-join rc=$?
ls -la gpurun_out/r2_final_*
