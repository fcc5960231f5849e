#!/bin/bash
# round 2, profiling call: launch list of a short bench, ncu --set full of the writer (DRAM traffic of one launch with
# the algorithmic bytes of exactly that launch) and of the step kernel
mkdir -p gpurun_out
SHORT="python bench.py --steps 5 --warmup 5 --burnin 20 --no-cpu-baseline --no-e2e"
timeout 300 $SHORT > gpurun_out/short_plain.json 2> gpurun_out/short_plain.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu_launch.log 2>&1
tail -2 gpurun_out/ncu_launch.log
CMD="python bench.py --steps 40 --warmup 20 --sample-every 1 --dump-launches --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/traffic_plain.json 2> gpurun_out/traffic_plain.err || { tail -5 gpurun_out/traffic_plain.err; exit 1; }
K=$(python - <<'PY'
import json
j = json.load(open("gpurun_out/traffic_plain.json"))
r = j["roofline"]
sp = r["timed_launch_sum_p"]
target = sum(sp) / len(sp)
k = min(range(len(sp)), key=lambda i: abs(sp[i] - target))
json.dump({"k": k, "sum_p": sp[k], "algorithmic_bytes": r["timed_launch_algorithmic_bytes"][k],
           "envs": j["config"]["envs_per_gpu"], "mean_sum_p": target}, open("gpurun_out/traffic_pick.json", "w"))
print(k)
PY
)
echo "picked timed launch $K: $(cat gpurun_out/traffic_pick.json)"
# writer launches before the timed region: 152 burn-in + 20 warm-up steps -> timed launch k is writer launch 172 + k
timeout 600 ncu --set full --clock-control none --import-source on -k regex:persp_write -s $((171 + K)) -c 3 \
  -o gpurun_out/prof_write_avg -f $CMD > gpurun_out/ncu_traffic.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/ncu_traffic.log
ncu -i gpurun_out/prof_write_avg.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum > gpurun_out/traffic_raw.csv 2>&1
tail -4 gpurun_out/traffic_raw.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:actor_step_kernel -s 180 -c 1 \
  -o gpurun_out/prof_actor -f $CMD > gpurun_out/ncu_actor.log 2>&1; echo "ncu actor rc=$?"
