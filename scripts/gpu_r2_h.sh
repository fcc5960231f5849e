#!/bin/bash
# round 2, call H: what the driver runs at round end (GPU suite, smoke, bench, reference arm) + reset timings / ncu
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    j=json.loads(open('gpurun_out/bench.json').read())
    print('value %.4g ms %.4f P %.1f'%(j['value'], j['ms_per_step'], j['config']['mean_perspectives_per_env']))
    for k in ('e2e','e2e_open_loop','e2e_u8_full'):
        print(k, '%.4g'%j[k]['value'], 'ms %.4f'%j[k]['ms_per_step'])
    r=j['roofline']; print('roofline', r['achieved'], r['frac'], r['write_peak'], r['traffic'], 'cpu', j['cpu_baseline']['value'])
except Exception as e: print('bench parse failed', e)
PY
tail -3 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cut -c1-200 gpurun_out/bench_ref.json
timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_20.json 2>gpurun_out/bench_20.err; cut -c1-200 gpurun_out/bench_20.json
rm -f gpurun_out/reset_times.txt
for args in "7 0.1" "7 0.0005" "5 0.1" "3 0.1" "8 0.1" "9 0.1" "15 0.1" "25 0.1"; do
  timeout 200 python scripts/micro/reset_only.py $args 2>/dev/null | tail -1 >> gpurun_out/reset_times.txt
done
cat gpurun_out/reset_times.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:reset_env -s 3 -c 1 -o gpurun_out/prof_reset_env -f python scripts/micro/reset_only.py 7 0.1 > gpurun_out/ncu_reset.log 2>&1
tail -1 gpurun_out/ncu_reset.log
timeout 300 python scripts/kernel_table.py --ds 3,5,7,9,15,25 --scale 4 > gpurun_out/kernel_table_4M.jsonl 2> gpurun_out/kernel_table.err
grep -i "reset p=0.1\|actor_pre" gpurun_out/kernel_table_4M.jsonl | cut -c1-200
