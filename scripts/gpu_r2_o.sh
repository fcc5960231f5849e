#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
rm -f gpurun_out/reset_times.txt
for args in "7 0.1" "9 0.1" "9 0.0005" "12 0.1" "15 0.1" "16 0.1" "17 0.1" "25 0.1"; do
  timeout 200 python scripts/micro/reset_only.py $args 2>/dev/null | tail -1 >> gpurun_out/reset_times.txt
done
cat gpurun_out/reset_times.txt
timeout 300 python scripts/kernel_table.py --ds 9,15 --scale 4 > gpurun_out/kernel_table_d9_15.jsonl 2> gpurun_out/kernel_table.err
grep -i "reset" gpurun_out/kernel_table_d9_15.jsonl | cut -c1-200
