#!/bin/bash
# round 2, call E: GPU suite, bench (priority streams + rows from the step kernel), small-lattice writer variants
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
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
for spec in "3 -" "3 g3_4" "3 s1" "4 -" "4 s1" "5 -" "5 s1"; do
  set -- $spec
  if [ "$2" = "-" ]; then unset TORIC_B200_LIB; else export TORIC_B200_LIB=$PWD/toric_rl_decoder_b200/variants/$2.so; fi
  echo -n "$2 " >> gpurun_out/group_variants2.txt
  timeout 200 python bench.py --d $1 --no-cpu-baseline --no-e2e --steps 100 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('d', j['config']['d'], 'ms/step %.4f writer %.4f ms %.0f GB/s'%(j['ms_per_step'], r['kernel_ms'], r['achieved']))" >> gpurun_out/group_variants2.txt
done
unset TORIC_B200_LIB
cat gpurun_out/group_variants2.txt
timeout 300 python scripts/kernel_table.py --ds 5,7,9 --scale 4 > gpurun_out/kernel_table_4M.jsonl 2> gpurun_out/kernel_table.err
grep -i "reset\|actor_pre" gpurun_out/kernel_table_4M.jsonl | cut -c1-200
