#!/bin/bash
# round 2, call F: GPU suite (reset_env kernel, d=3 writer), reset timings + ncu, bench, c5 at N=1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
for args in "7 0.1" "7 0.0005" "7 0.005" "5 0.1" "3 0.1" "8 0.1" "9 0.1"; do
  timeout 200 python scripts/micro/reset_only.py $args 2>/dev/null | tail -1 >> gpurun_out/reset_times.txt
done
cat gpurun_out/reset_times.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:reset_env -s 3 -c 1 -o gpurun_out/prof_reset_env -f python scripts/micro/reset_only.py 7 0.1 > gpurun_out/ncu_reset.log 2>&1
tail -2 gpurun_out/ncu_reset.log
timeout 300 python scripts/kernel_table.py --ds 3,5,7,9 --scale 4 > gpurun_out/kernel_table_4M.jsonl 2> gpurun_out/kernel_table.err
grep -i "reset" gpurun_out/kernel_table_4M.jsonl | cut -c1-200
timeout 200 python bench.py --d 3 --no-cpu-baseline --no-e2e --steps 100 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('d', j['config']['d'], 'ms/step %.4f writer %.4f ms %.0f GB/s'%(j['ms_per_step'], r['kernel_ms'], r['achieved']))"
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench.json
bash scripts/gpu_r2_multi.sh 1 > gpurun_out/multi_n1.log 2>&1; grep -v "^$" gpurun_out/multi_n1.log | cut -c1-300 | tail -12
