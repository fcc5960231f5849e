#!/bin/bash
# One gpurun call: DRAM traffic of ONE persp_write_kernel launch whose sum P is close to the bench average, with the
# algorithmic bytes of exactly that launch beside it.  (1) plain run of the command that lists sum P per timed
# launch; (2) pick the launch; (3) ncu --set full on that launch of the same (deterministic) command.
mkdir -p gpurun_out
CMD="python bench.py --steps 76 --warmup 80 --sample-every 1 --dump-launches --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/traffic_plain.json 2> gpurun_out/traffic_plain.err || { tail -5 gpurun_out/traffic_plain.err; exit 1; }
K=$(python - <<'PY'
import json
j = json.load(open("gpurun_out/traffic_plain.json"))
r = j["roofline"]
target = 74.0 * j["config"]["envs_per_gpu"]
sp = r["timed_launch_sum_p"]
k = min(range(len(sp)), key=lambda i: abs(sp[i] - target))
json.dump({"k": k, "sum_p": sp[k], "algorithmic_bytes": r["timed_launch_algorithmic_bytes"][k],
           "envs": j["config"]["envs_per_gpu"]}, open("gpurun_out/traffic_pick.json", "w"))
print(k)
PY
)
echo "picked timed launch $K: $(cat gpurun_out/traffic_pick.json)"
# writer launches before the timed region: the 80 warm-up steps, so timed launch k is launch 80 + k of the process;
# launches K-1, K, K+1 are captured so that the step <-> launch mapping can be checked (sum P grows ~0.5 % per step)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:persp_write -s $((79 + K)) -c 3 \
  -o gpurun_out/prof_write_avg -f $CMD > gpurun_out/ncu_traffic.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_traffic.log
ncu -i gpurun_out/prof_write_avg.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum > gpurun_out/traffic_raw.csv 2>&1
tail -4 gpurun_out/traffic_raw.csv
