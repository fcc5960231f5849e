#!/bin/bash
# GPU parity tests, then the headline bench at every perspective element type (float32 = headline).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
for dt in float32 bfloat16 float16 uint8 float32; do
  timeout 300 python bench.py --out-dtype $dt --no-cpu-baseline > gpurun_out/bench_$dt.json 2> gpurun_out/bench_$dt.err; echo "bench $dt rc=$?"
  python - gpurun_out/bench_$dt.json <<'PY'
import json, sys
j = json.load(open(sys.argv[1])); r = j["roofline"]
print("%s: %.3e env-steps/s  step %.4f ms  writer %.4f ms (%.0f GB/s, frac %.3f)  e2e %.3e" % (
    j["config"]["out_dtype"], j["value"], j["ms_per_step"], r["kernel_ms"], r["achieved"], r["frac"], j["e2e"]["value"]))
PY
done
