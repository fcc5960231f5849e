#!/bin/bash
# ncu --set full capture of the perspective writer at a given element type (after a plain run exits 0).
DT=${1:-bfloat16}
mkdir -p gpurun_out
CMD="python bench.py --out-dtype $DT --steps 5 --warmup 80 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain_$DT.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:persp_write -s 81 -c 1 -o gpurun_out/prof_write_$DT -f $CMD > gpurun_out/ncu_$DT.log 2>&1
tail -2 gpurun_out/ncu_$DT.log
