#!/bin/bash
# A/B timing of the writer-kernel variants built by scripts/micro/build_variants.py (run on the GPU box).
mkdir -p gpurun_out
for so in scripts/micro/variants/lib_*.so; do
  name=$(basename $so .so)
  TORIC_B200_LIB=$PWD/$so timeout 300 python bench.py --steps 60 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/var_$name.json 2> gpurun_out/var_$name.err
  python - "$name" gpurun_out/var_$name.json <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[2]))
    r = j["roofline"]
    print("%-14s step %.4f ms  writer %.4f ms  %.0f GB/s  value %.4g" % (sys.argv[1], j["ms_per_step"], r["kernel_ms"], r["achieved"], j["value"]))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
if [ -n "$NCU_VARIANT" ]; then
  TORIC_B200_LIB=$PWD/scripts/micro/variants/lib_$NCU_VARIANT.so timeout 600 ncu --set full --clock-control none --import-source on -k regex:persp_write -s 12 -c 1 -o gpurun_out/prof_var -f python bench.py --steps 5 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/ncu_var.log 2>&1
  tail -3 gpurun_out/ncu_var.log
fi
