#!/bin/bash
# A/B of scripts/micro/variants/lib_*.so at several lattice sizes / element types: ab_d.sh "5 7" "float32 bfloat16"
for d in ${1:-7}; do for dt in ${2:-float32}; do for round in 1 2; do
for so in scripts/micro/variants/lib_*.so; do
  name=$(basename $so .so)
  TORIC_B200_LIB=$PWD/$so timeout 300 python bench.py --d $d --out-dtype $dt --no-cpu-baseline --no-e2e > gpurun_out/ab_$name.json 2>/dev/null
  python - "$name" gpurun_out/ab_$name.json $d $dt <<'PY'
import json, sys
j = json.load(open(sys.argv[2])); r = j["roofline"]
print("d=%s %-8s %-10s step %.4f ms  writer %.4f ms (%.0f GB/s)" % (sys.argv[3], sys.argv[4], sys.argv[1], j["ms_per_step"], r["kernel_ms"], r["achieved"]))
PY
done; done; done; done
