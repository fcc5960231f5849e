#!/bin/bash
# round 2, call C: GPU suite (new sampler, single-pass, group writer, native host step), bench, diagnostics
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    j=json.loads(open('gpurun_out/bench.json').read())
    print('value %.4g ms %.4f P %.1f'%(j['value'], j['ms_per_step'], j['config']['mean_perspectives_per_env']))
    for k in ('e2e','e2e_open_loop','e2e_u8_full'):
        print(k, '%.4g'%j[k]['value'], 'ms %.4f'%j[k]['ms_per_step'])
    print('roofline', j['roofline']['achieved'], j['roofline']['write_peak'], 'cpu', j['cpu_baseline']['value'])
except Exception as e: print('bench parse failed', e)
PY
tail -3 gpurun_out/bench.err
# single-pass diagnostics: per-env pre-phase and look-back wait
TORIC_B200_LIB=$PWD/toric_rl_decoder_b200/variants/timing.so timeout 200 python scripts/micro/fused_timing.py > gpurun_out/fused_timing.json 2> gpurun_out/fused_timing.err
TORIC_B200_LIB=$PWD/toric_rl_decoder_b200/variants/timing_nolb.so timeout 200 python scripts/micro/fused_timing.py >> gpurun_out/fused_timing.json 2>> gpurun_out/fused_timing.err
cat gpurun_out/fused_timing.json; tail -3 gpurun_out/fused_timing.err
for sp in 0 1; do
  timeout 300 python bench.py --single-pass $sp --no-cpu-baseline --no-e2e --steps 100 >> gpurun_out/ab_single_pass.jsonl 2>> gpurun_out/ab.err
done
# small lattices: group writer vs one env per warp
for d in 3 4 5; do
  timeout 200 python bench.py --d $d --no-cpu-baseline --no-e2e --steps 100 >> gpurun_out/ab_group_writer.jsonl 2>> gpurun_out/ab.err
  TORIC_B200_LIB=$PWD/toric_rl_decoder_b200/variants/nogroup.so timeout 200 python bench.py --d $d --no-cpu-baseline --no-e2e --steps 100 >> gpurun_out/ab_group_writer.jsonl 2>> gpurun_out/ab.err
done
python - <<'PY'
import json
for f in ('gpurun_out/ab_single_pass.jsonl','gpurun_out/ab_group_writer.jsonl'):
    for l in open(f):
        j=json.loads(l); r=j['roofline']
        print(f[-22:], 'd', j['config']['d'], j.get('step_form','')[:12], 'ms/step %.4f'%j['ms_per_step'], 'value %.4g'%j['value'], 'kernel %.4f ms %.0f GB/s'%(r['kernel_ms'], r['achieved']))
PY
# reset kernel: plain run, then ncu
timeout 200 python scripts/micro/reset_only.py 7 0.1 > gpurun_out/reset_plain.log 2>&1 && cat gpurun_out/reset_plain.log | tail -1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:reset_kernel -s 3 -c 1 -o gpurun_out/prof_reset -f python scripts/micro/reset_only.py 7 0.1 > gpurun_out/ncu_reset.log 2>&1
tail -2 gpurun_out/ncu_reset.log
timeout 300 python scripts/kernel_table.py --ds 5,7,9 --scale 4 > gpurun_out/kernel_table_4M.jsonl 2> gpurun_out/kernel_table.err
grep -i "reset\|actor_pre" gpurun_out/kernel_table_4M.jsonl | cut -c1-200
