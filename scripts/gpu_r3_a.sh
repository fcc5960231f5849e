#!/bin/bash
# round 3, call A: reference-loop tests after the uint8 shortcut, bench with e2e_reference_api, then the ncu launch
# list of a short bench command and one full capture of the writer (numbers printed under ncu are never bench values)
python -m pytest tests/test_gpu_reference_loop.py tests/test_gpu_parity.py tests/test_gpu_policy.py -m gpu -q -x > gpurun_out/r3a_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r3a_tests.log; tail -n 3 gpurun_out/r3a_tests.log
python bench.py --steps 20 --warmup 3 > gpurun_out/r3a_bench.json 2> gpurun_out/r3a_bench.err; echo "bench rc=$?"
NCU_CMD="python bench.py --steps 5 --warmup 5 --burnin 20 --repeats 1 --kernel-samples 0 --no-cpu-baseline --no-e2e"
$NCU_CMD > gpurun_out/r3a_ncu_cmd_plain.json 2>&1; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r3a_launches.csv $NCU_CMD > gpurun_out/r3a_ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:persp_write -s 27 -c 1 -o gpurun_out/r3a_prof_write -f $NCU_CMD > gpurun_out/r3a_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | tail -n 8
