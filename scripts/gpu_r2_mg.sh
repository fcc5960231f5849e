#!/bin/bash
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 scripts/gpu_multi_check.py > gpurun_out/multi_check.log 2>&1; echo "rc=$?"
grep -v "^$" gpurun_out/multi_check.log | tail -40 | cut -c1-400
