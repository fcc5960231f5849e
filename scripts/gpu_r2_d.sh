#!/bin/bash
# round 2, call D: GPU suite, bench (native host step), reset-kernel variants + ncu
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
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
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cut -c1-250 gpurun_out/bench_ref.json
for v in "" reset3 reset4 resetD; do
  if [ -n "$v" ]; then export TORIC_B200_LIB=$PWD/toric_rl_decoder_b200/variants/$v.so; else unset TORIC_B200_LIB; fi
  for args in "7 0.1" "7 0.0005" "5 0.1" "9 0.1" "25 0.1"; do
    echo -n "variant=${v:-default} " >> gpurun_out/reset_variants.txt
    timeout 200 python scripts/micro/reset_only.py $args 2>/dev/null | tail -1 >> gpurun_out/reset_variants.txt
  done
done
unset TORIC_B200_LIB
cat gpurun_out/reset_variants.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:reset_kernel -s 3 -c 1 -o gpurun_out/prof_reset -f python scripts/micro/reset_only.py 7 0.1 > gpurun_out/ncu_reset.log 2>&1
tail -2 gpurun_out/ncu_reset.log
timeout 300 python scripts/kernel_table.py --ds 5,7,9 --scale 4 > gpurun_out/kernel_table_4M.jsonl 2> gpurun_out/kernel_table.err
grep -i "reset\|actor_pre" gpurun_out/kernel_table_4M.jsonl | cut -c1-200
timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_20.json 2>gpurun_out/bench_20.err; cut -c1-200 gpurun_out/bench_20.json
# group-writer sizes (variants built with -DTC_GROUP_D{3,4,5}=K); default build = 16 / 8 / 4
for spec in "3 g3_8" "3 g3_32" "4 g4_4" "4 g4_16" "5 g5_2" "5 g5_8"; do
  set -- $spec
  if [ -f toric_rl_decoder_b200/variants/$2.so ]; then
    echo -n "$2 " >> gpurun_out/group_variants.txt
    TORIC_B200_LIB=$PWD/toric_rl_decoder_b200/variants/$2.so timeout 200 python bench.py --d $1 --no-cpu-baseline --no-e2e --steps 100 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('d', j['config']['d'], 'ms/step %.4f writer %.4f ms %.0f GB/s'%(j['ms_per_step'], r['kernel_ms'], r['achieved']))" >> gpurun_out/group_variants.txt
  fi
done
cat gpurun_out/group_variants.txt
