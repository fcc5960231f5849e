#!/bin/bash
# Alternating A/B timing of every scripts/micro/variants/lib_*.so on the same box (3 rounds).
for round in 1 2 3; do
for so in scripts/micro/variants/lib_*.so; do
  name=$(basename $so .so)
  TORIC_B200_LIB=$PWD/$so timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu-baseline --no-e2e > gpurun_out/ab_$name.json 2>/dev/null
  python - "$name" gpurun_out/ab_$name.json <<'PY'
import json, sys
j = json.load(open(sys.argv[2])); r = j["roofline"]
print("%-10s step %.4f ms  writer %.4f ms (%.0f GB/s)  rest %.4f ms" % (sys.argv[1], j["ms_per_step"], r["kernel_ms"], r["achieved"], j["ms_per_step"] - r["kernel_ms"]))
PY
done; done
