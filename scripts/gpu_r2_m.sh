#!/bin/bash
# A/B: L2 eviction-priority hint on the writer's TMA bulk stores (TC_TMA_L2_HINT)
mkdir -p gpurun_out; rm -f gpurun_out/ab_tma_l2_hint.txt
for rep in 1 2; do for v in - hint1 hint2 hint3; do
  if [ "$v" = "-" ]; then unset TORIC_B200_LIB; else export TORIC_B200_LIB=$PWD/toric_rl_decoder_b200/variants/$v.so; fi
  echo -n "$v " >> gpurun_out/ab_tma_l2_hint.txt
  timeout 200 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('ms/step %.4f value %.4g writer %.4f ms %.0f GB/s'%(j['ms_per_step'], j['value'], r['kernel_ms'], r['achieved']))" >> gpurun_out/ab_tma_l2_hint.txt
done; done
unset TORIC_B200_LIB
cat gpurun_out/ab_tma_l2_hint.txt
