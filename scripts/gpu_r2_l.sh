#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python scripts/kernel_table.py --ds 5,7 --scale 4 > gpurun_out/kernel_table_4M.jsonl 2> gpurun_out/kernel_table.err
cut -c1-200 gpurun_out/kernel_table_4M.jsonl
timeout 300 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read()); print('value %.4g ms %.4f e2e %.4g'%(j['value'], j['ms_per_step'], j['e2e']['value']))"
