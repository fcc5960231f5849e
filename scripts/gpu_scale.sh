#!/bin/bash
# Scaling check on one box: bench.py at N = 1, 2, 4, 8 (torchrun for N > 1), as the driver launches it.
mkdir -p gpurun_out
NG=${1:-8}
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > gpurun_out/smi_scale.csv
for n in 1 2 4 8; do
  [ $n -gt $NG ] && continue
  if [ $n -eq 1 ]; then
    timeout 300 python bench.py --gpus 1 --no-cpu-baseline > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $n --no-cpu-baseline > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
  echo "N=$n rc=$?"
  python - gpurun_out/scale_n$n.json <<'PY'
import json, sys
try:
    j = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print("N=%d value %.4e env-steps/s  step %.4f ms  e2e %.4e  e2e(state rows) %.4e  writer frac %.3f" % (
        j["n_gpus"], j["value"], j["ms_per_step"], j["e2e"]["value"], j["e2e_state_rows"]["value"],
        j["roofline"]["frac"]))
except Exception as e:
    print("parse failed", e)
PY
done
