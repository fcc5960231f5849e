#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python scripts/run_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
cut -c1-330 gpurun_out/configs.jsonl; tail -3 gpurun_out/configs.err
