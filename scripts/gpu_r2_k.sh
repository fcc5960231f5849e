#!/bin/bash
# A/B: CTA size of the step / count kernels (TC_COUNT_THREADS)
mkdir -p gpurun_out; rm -f gpurun_out/ab_count_threads.txt
for rep in 1 2; do for v in - ct128 ct64; do
  if [ "$v" = "-" ]; then unset TORIC_B200_LIB; else export TORIC_B200_LIB=$PWD/toric_rl_decoder_b200/variants/$v.so; fi
  echo -n "$v " >> gpurun_out/ab_count_threads.txt
  timeout 200 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('ms/step %.4f value %.4g writer %.4f ms rest %.4f ms'%(j['ms_per_step'], j['value'], r['kernel_ms'], j['ms_per_step']-r['kernel_ms']))" >> gpurun_out/ab_count_threads.txt
done; done
unset TORIC_B200_LIB
cat gpurun_out/ab_count_threads.txt
timeout 300 python -m pytest tests/test_gpu_actor.py -m gpu -x -q -k "trajectory or full_size" 2>&1 | tail -2
