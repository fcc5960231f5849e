#!/bin/bash
# gpurun call for the graphed small-batch decoder: its tests, then configs[0] timing, then the whole GPU suite.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_rollout.py -m gpu -x -q > gpurun_out/pytest_rollout.log 2>&1; echo "rollout rc=$?" >> gpurun_out/pytest_rollout.log
tail -30 gpurun_out/pytest_rollout.log
timeout 300 python scripts/run_configs.py --only c1 > gpurun_out/configs_c1.jsonl 2> gpurun_out/configs_c1.err; echo "c1 rc=$?"
cat gpurun_out/configs_c1.jsonl; tail -5 gpurun_out/configs_c1.err
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
