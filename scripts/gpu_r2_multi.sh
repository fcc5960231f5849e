#!/bin/bash
# round 2, multi-GPU call: bash scripts/gpu_r2_multi.sh N [c4_n] [c4_n_high] [c3_syndromes]
# bench at N ranks (value + the three e2e forms), BASELINE configs[4] (c5), and at N=8 configs[3] (c4) and configs[2] (c3)
N=${1:-8}; C4N=${2:-0}; C4H=${3:-0}; C3S=${4:-0}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
if [ "$N" = "1" ]; then RUN="python"; fi
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi_n$N.csv
timeout 600 $RUN bench.py --gpus $N --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"
python - <<PY
import json
try:
    j=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1])
    print('N=$N value %.4g ms %.4f'%(j['value'], j['ms_per_step']), ' '.join('%s %.4g (%.4f ms)'%(k, j[k]['value'], j[k]['ms_per_step']) for k in ('e2e','e2e_open_loop','e2e_u8_full')))
except Exception as e: print('bench parse failed', e)
PY
tail -2 gpurun_out/bench_n$N.err
timeout 900 $RUN scripts/c5_sharded.py > gpurun_out/c5_n$N.jsonl 2> gpurun_out/c5_n$N.err; echo "c5 rc=$?"
cut -c1-260 gpurun_out/c5_n$N.jsonl; tail -2 gpurun_out/c5_n$N.err
if [ "$C4N" != "0" ]; then
  timeout 1500 $RUN scripts/c4_big.py --count $C4N --count-high-p $C4H > gpurun_out/c4_n$N.jsonl 2> gpurun_out/c4_n$N.err; echo "c4 rc=$?"
  cut -c1-420 gpurun_out/c4_n$N.jsonl; tail -2 gpurun_out/c4_n$N.err
fi
if [ "$C3S" != "0" ]; then
  timeout 900 $RUN scripts/c3_sharded.py --syndromes $C3S > gpurun_out/c3_n$N.jsonl 2> gpurun_out/c3_n$N.err; echo "c3 rc=$?"
  cut -c1-900 gpurun_out/c3_n$N.jsonl; tail -3 gpurun_out/c3_n$N.err
fi
