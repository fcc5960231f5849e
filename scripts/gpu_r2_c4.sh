#!/bin/bash
# BASELINE configs[3] on N GPUs: bash scripts/gpu_r2_c4.sh N COUNT COUNT_HIGH_P
N=${1:-8}; mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 scripts/c4_big.py --count $2 --count-high-p $3 > gpurun_out/c4_n$N.jsonl 2> gpurun_out/c4_n$N.err; echo "c4 rc=$?"
cut -c1-600 gpurun_out/c4_n$N.jsonl; tail -3 gpurun_out/c4_n$N.err
