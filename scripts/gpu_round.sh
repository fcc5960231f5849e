#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench, then ncu launch list + full capture of the top kernel.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
cat gpurun_out/bench_ref.json
NCU_CMD="python bench.py --steps 5 --warmup 80 --no-cpu-baseline --no-e2e"   # steady state: P/env has saturated after ~75 steps
timeout 300 $NCU_CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches.csv $NCU_CMD > gpurun_out/ncu_launch.log 2>&1
timeout 300 $NCU_CMD > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:persp_write -s 81 -c 2 -o gpurun_out/prof_write -f $NCU_CMD > gpurun_out/ncu_full.log 2>&1
timeout 300 $NCU_CMD > gpurun_out/plain3.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:actor_step -s 81 -c 2 -o gpurun_out/prof_actor -f $NCU_CMD > gpurun_out/ncu_full2.log 2>&1
timeout 300 python scripts/kernel_table.py > gpurun_out/kernel_table.jsonl 2> gpurun_out/kernel_table.err
timeout 900 python scripts/run_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err
for dt in bfloat16 float16 uint8; do
  timeout 300 python bench.py --out-dtype $dt --no-cpu-baseline > gpurun_out/bench_$dt.json 2> gpurun_out/bench_$dt.err
done
ls -la gpurun_out
