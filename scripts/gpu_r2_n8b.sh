#!/bin/bash
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523"
timeout 300 $RUN bench.py --impl reference --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_ref_n8.json 2> gpurun_out/bench_ref_n8.err; echo "ref rc=$?"; grep '^{' gpurun_out/bench_ref_n8.json | cut -c1-200
timeout 600 $RUN bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_n8_20.json 2> gpurun_out/bench_n8_20.err; echo "bench20 rc=$?"
timeout 600 $RUN bench.py --gpus 8 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "bench rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_n8_20.json','gpurun_out/bench_n8.json'):
    try:
        j=json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f, 'steps', j['steps'], 'value %.4g ms %.4f'%(j['value'], j['ms_per_step']), ' '.join('%s %.4g'%(k, j[k]['value']) for k in ('e2e','e2e_open_loop','e2e_u8_full')), j['clocks']['sm_mhz'], j['clocks']['reasons'])
    except Exception as e: print(f, 'parse failed', e)
PY
tail -2 gpurun_out/bench_n8.err
