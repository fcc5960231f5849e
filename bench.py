#!/usr/bin/env python
"""Headline benchmark: env-steps/s including perspectives at d=7 (BASELINE.json configs[1]).

  python bench.py --gpus N --steps K --warmup W            # B200 arm (this repo's CUDA engine)
  python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the oracle port on the host cores

One "step" = one fused actor iteration over every env of the rank: on-device random action -> step
(incremental syndrome, reward, terminal) -> auto-reset of finished envs -> all perspectives of the new
states written as the float32 (sumP,2,d,d) network-input tensor + positions + counts/offsets.
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "env_steps_per_sec_incl_perspectives_d7"
UNIT = "env-steps/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=("b200", "reference"))
    ap.add_argument("--d", type=int, default=7)
    ap.add_argument("--p", type=float, default=0.1)
    ap.add_argument("--envs", type=int, default=65536, help="envs per GPU (weak scaling)")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--cpu-envs", type=int, default=8192, help="envs in the bounded CPU sample")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--out-dtype", default="float32", choices=("float32", "bfloat16", "float16", "uint8"),
                    help="element type of the perspective tensor; float32 = the reference's (headline)")
    ap.add_argument("--sample-every", type=int, default=8,
                    help="time the writer kernel with CUDA events on every k-th step of the timed region")
    ap.add_argument("--dump-launches", action="store_true",
                    help="add the sum P and algorithmic bytes of every timed writer launch to the roofline object "
                         "(to match an ncu capture of one launch of the same command, scripts/gpu_traffic.sh)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def workload_name(a):
    short = {"float32": "fp32", "bfloat16": "bf16", "float16": "fp16", "uint8": "u8"}[a.out_dtype]
    return ("d=%d p=%g %d envs/GPU random actions: step + auto-reset + perspectives (%s)"
            % (a.d, a.p, a.envs, short))


def b_alg(d, n_envs, sum_p, s=4):
    """Algorithmic bytes of one whole step (SURVEY.md section 8d / BASELINE.md section 4)."""
    per_env = 2 * ((4 * d * d + 7) // 8) + 2 * ((2 * d * d + 7) // 8) + 30
    return n_envs * per_env + sum_p * (2 * d * d * s + 4)


def write_kernel_bytes(d, n_envs, sum_p, s=4):
    """Algorithmic bytes of one persp_write_kernel launch: perspectives + positions out, syndrome + offsets."""
    return sum_p * (2 * d * d * s + 4) + n_envs * (((2 * d * d + 7) // 8) + 8)


# ----------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons sampled through NVML from a background thread DURING the timed
    region (the timed region is tens of milliseconds, too short for `nvidia-smi -lms`)."""

    def __init__(self, gpu_index=0, period_s=0.002):
        self.gpu_index, self.period = gpu_index, period_s
        self.sm, self.reasons, self.sm_max = [], set(), None
        self._stop = False
        self._thread = None
        self._h = None

    def _loop(self):
        import pynvml as nv
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown")
                else nv.nvmlClocksThrottleReasonHwSlowdown,
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown",
                                               getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40)),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown",
                                               getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20)),
                "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap",
                                        getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4))}
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons",
                              getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons", None))
        while not self._stop:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                if get_reasons is not None:
                    r = int(get_reasons(self._h))
                    for name, bit in bits.items():
                        if r & int(bit):
                            self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def __enter__(self):
        try:
            import threading
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.gpu_index
            if vis:
                try:
                    idx = int(vis.split(",")[self.gpu_index])
                except Exception:
                    pass
            self._h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.sm_max = float(nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM))
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        except Exception:
            self._thread = None
        return self

    def __exit__(self, *exc):
        self._stop = True
        if self._thread is not None:
            self._thread.join(timeout=2)

    def summary(self):
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "how": "NVML, 2 ms period, timed region only"}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------------------------
def cpu_arm(a, seconds, n_envs, threads=0):
    """Time the oracle's C port of the same loop on the host cores (bounded sample)."""
    from oracle import c_oracle as C
    loop = C.ActorLoop(a.d, n_envs, a.p, seed=a.seed, nthreads=threads)
    loop.run(2)                                  # warm caches / page in the buffers
    t0 = time.perf_counter()
    loop.run(2)
    per_step = max((time.perf_counter() - t0) / 2, 1e-6)
    steps = max(3, int(seconds / per_step))
    c0 = loop.counters.copy()
    t0 = time.perf_counter()
    loop.run(steps)
    dt = time.perf_counter() - t0
    c = loop.counters - c0
    return {"value": float(c[0]) / dt, "unit": UNIT, "cores": loop.nthreads, "kind": "port",
            "sample": "%d envs x %d steps, d=%d p=%g, C port of EnvSet.step + generatePerspectiveBatch + float32 "
                      "concatenate (oracle/toric_oracle.c), one slice per thread, %.1f s; mean perspectives/env %.1f"
                      % (n_envs, steps, a.d, a.p, dt, float(c[3]) / max(float(c[0]), 1.0)),
            "seconds": dt, "steps": steps, "envs": n_envs}


def run_reference(a):
    """--impl reference: the CPU arm.  The reference is pure Python/numba and cannot travel to the GPU box,
    so this times the oracle's C port of the same loop (kind "port") with every host core; each step
    advances a bounded sample of --cpu-envs environments once."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import c_oracle as C
    loop = C.ActorLoop(a.d, a.cpu_envs, a.p, seed=a.seed, nthreads=0)
    loop.run(max(a.warmup, 1))
    c0 = loop.counters.copy()
    t0 = time.perf_counter()
    loop.run(a.steps)
    dt = time.perf_counter() - t0
    c = loop.counters - c0
    value = float(c[0]) / dt
    sample = ("%d envs x %d steps, d=%d p=%g, C port of EnvSet.step + generatePerspectiveBatch + float32 concatenate "
              "(oracle/toric_oracle.c), one slice per thread; mean perspectives/env %.1f"
              % (a.cpu_envs, a.steps, a.d, a.p, float(c[3]) / max(float(c[0]), 1.0)))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic", "config": {"workload": workload_name(a),
                                            "note": "CPU arm: bounded sample of the same loop on the host cores"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": loop.nthreads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
def run_b200(a):
    import numpy as np
    import torch
    import toric_rl_decoder_b200 as pkg

    from toric_rl_decoder_b200 import sharding
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    rank, world, local = sharding.init_from_env("nccl")     # one process per GPU (torchrun env)
    dist = None
    if world > 1:
        import torch.distributed as dist
    dev = torch.device("cuda", local)
    d, n = a.d, a.envs

    out_dtype = getattr(torch, a.out_dtype)
    esz = torch.empty((), dtype=out_dtype).element_size()
    eng = pkg.ToricEngine(d, n, device=dev, seed=a.seed, env_base=rank * n, p_error=a.p,
                          out_dtype=out_dtype)                                              # envs shard by rank
    eng.ensure_capacity(eng.worst_case_capacity())
    eng.reset()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        eng.actor_step()
    barrier()
    eng.counters.zero_()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # The step is the fused call a user makes (tc_actor_step: three back-to-back launches).  Every
    # `a.sample_every`-th step is issued as its two halves with CUDA events around the writer launch, which
    # gives the dominant kernel's duration inside the timed region without putting two event records (and a
    # break of the programmatic dependent launch chain) into every step.
    sampled = [k for k in range(a.steps) if k % a.sample_every == 0]
    evb = {k: torch.cuda.Event(enable_timing=True) for k in sampled}
    evc = {k: torch.cuda.Event(enable_timing=True) for k in sampled}
    totals = torch.zeros(len(sampled), dtype=torch.int64, device=dev)      # sum P of the sampled launches
    barrier()
    with ClockSampler(local) as clk:
        ev0.record()
        for k in range(a.steps):
            if k in evb:
                eng.actor_pre()            # actor_step_kernel + scan_blocks_kernel
                evb[k].record()
                eng.persp_write()          # persp_write_kernel
                evc[k].record()
                totals[k // a.sample_every].copy_(eng.offsets[n], non_blocking=True)
            else:
                eng.actor_step()           # the same three kernels, one call
        ev1.record()
        barrier()
    ms = ev0.elapsed_time(ev1)
    write_ms = sum(evb[k].elapsed_time(evc[k]) for k in sampled) / len(sampled)
    sampled_sum_p = float(totals.sum().item()) / len(sampled)
    per_launch_sum_p = [int(x) for x in totals.tolist()] if a.dump_launches else None
    ctr = eng.counters.clone()
    assert int(eng.overflow.item()) == 0
    ms = sharding.max_over_ranks(ms, dev)                  # time = max over ranks
    sharding.allreduce_counters(ctr)                       # the path's only collective: counter all-reduce
    c = {name: int(ctr[i].item()) for i, name in enumerate(pkg.engine.COUNTER_NAMES)}
    total_steps = c["env_steps"]
    assert total_steps == n * a.steps * world
    value = total_steps / (ms * 1e-3)
    sum_p_local = int(eng.counters[5].item())
    p_bar = sum_p_local / float(n * a.steps)

    peak, peak_src = load_peaks()
    wbytes = write_kernel_bytes(d, n, sampled_sum_p, esz)      # the launches that were timed
    achieved = wbytes / (write_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "persp_write_kernel<%d,%s>" % (d, a.out_dtype), "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                "kernel_ms": write_ms, "kernel_share_of_step": write_ms * a.steps / ms,
                "kernel_launches_timed": len(sampled),
                "step_achieved": b_alg(d, n, sum_p_local / a.steps, esz) / (ms / a.steps * 1e-3) / 1e9}
    if per_launch_sum_p is not None:
        roofline["timed_launch_steps"] = sampled
        roofline["timed_launch_sum_p"] = per_launch_sum_p
        roofline["timed_launch_algorithmic_bytes"] = [write_kernel_bytes(d, n, sp, esz) for sp in per_launch_sum_p]
    # dram read+write bytes of ONE writer launch from the committed `ncu --set full` capture, next to the
    # algorithmic bytes of that same launch (the launch is chosen with sum P close to the bench average, so the
    # figure is comparable with achieved x kernel_ms)
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr) and a.out_dtype == "float32" and d == 7:
        try:
            tj = json.load(open(tr))
            roofline["traffic"] = tj.get("persp_write_bytes_per_launch")
            if tj.get("launch_algorithmic_bytes"):
                roofline["traffic_launch_algorithmic_bytes"] = tj["launch_algorithmic_bytes"]
                roofline["traffic_over_algorithmic"] = tj["persp_write_bytes_per_launch"] / tj["launch_algorithmic_bytes"]
        except Exception:
            pass

    # ---- e2e: reference-facing call with HOST buffers (actions in, step results out) ----------
    e2e = e2e_rows = None
    if not a.no_e2e:
        del eng
        torch.cuda.empty_cache()

        def make_engine():
            e = pkg.ToricEngine(d, n, device=dev, seed=a.seed, env_base=rank * n, p_error=a.p, out_dtype=out_dtype)
            e.ensure_capacity(e.worst_case_capacity())
            return e
        recorded = record_device_policy(a, make_engine, max(3, a.warmup) + a.steps)
        e2e = run_e2e(a, make_engine, recorded, dev, dist, barrier, world, "u8")
        e2e_rows = run_e2e(a, make_engine, recorded, dev, dist, barrier, world, "rows")

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cpu = cpu_arm(a, a.cpu_seconds, a.cpu_envs)
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u32 bit planes -> %s" % {"float32": "f32", "bfloat16": "bf16",
                                                                        "float16": "f16", "uint8": "u8"}[a.out_dtype],
                "data": "synthetic",
                "config": {"workload": workload_name(a), "d": d, "p_error": a.p, "envs_per_gpu": n,
                           "global_envs": n * world, "parallelism": "env-sharded x%d, no data-path collective" % world,
                           "mean_perspectives_per_env": p_bar, "out_dtype": a.out_dtype,
                           "l2": "no flush needed: each step streams %.2f GB of new output (>> 126 MB L2)"
                                 % (wbytes / 1e9),
                           "b_alg_bytes_per_env_step": b_alg(d, n, sum_p_local / a.steps, esz) / n},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "e2e_state_rows": e2e_rows,
                "gpu_launches": 3 * a.steps,
                "clocks": clk.summary(),
                "counters": {k: c[k] for k in ("episodes", "solved", "ground_kept", "timeouts", "perspectives")}}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def record_device_policy(a, make_engine, total_steps):
    """The device's random-policy trajectory (the one `value` times), recorded once: actions (steps,n,4) int32 in
    pinned host memory plus the running perspective total after every step."""
    import torch
    rec = make_engine()
    rec.reset()
    host = torch.empty((total_steps, a.envs, 4), dtype=torch.int32).pin_memory()
    sum_p = []
    for k in range(total_steps):
        rec.actor_step()
        host[k].copy_(rec.actions, non_blocking=True)
        sum_p.append(rec.counters[5].clone())
    torch.cuda.synchronize()
    sum_p = [int(x.item()) for x in sum_p]
    del rec
    torch.cuda.empty_cache()
    return host, sum_p


def run_e2e(a, make_engine, recorded, dev, dist, barrier, world, states):
    """Same metric through the host-facing call (toric_rl_decoder_b200.HostActorPipeline): every step copies
    that step's actions (n,4) int32 from pinned host memory, runs the fused actor step, and reads rewards,
    terminals, perspective counts and the new states (uint8 n,2,d,d) back to pinned host memory.  Uploads
    and downloads run on their own streams with up to three steps in flight; every step's results are read on
    the host inside the timed region.
    The host policy replays the actions the device's random policy chose on the same seed (recorded beforehand,
    record_device_policy), so the envs walk exactly the trajectory `value` was timed on -- same resets, same
    perspectives per env -- and the perspective totals are checked against the recording.
    The perspective tensor stays on the device, where the reference also puts it (`.to(device)`,
    src/numba/util_actor.py:39).  states="u8" (the `e2e` key) downloads the reference's one-byte-per-stabiliser
    array; states="rows" (`e2e_state_rows`) downloads np.packbits(states, axis=-1, bitorder="little") instead --
    at 8 ranks the box's aggregate device->host rate (~90 GB/s, profiles/r01_e2e_scale_probe.txt), not the
    GPUs, bounds the u8 form."""
    import torch
    from toric_rl_decoder_b200 import HostActorPipeline, sharding
    n = a.envs
    host_actions, rec_sum_p = recorded
    warm = host_actions.shape[0] - a.steps
    eng = make_engine()
    eng.reset()
    depth = 3
    pipe = HostActorPipeline(eng, depth=depth, states=states)
    skey = "states" if states == "u8" else "state_rows"
    checksum = 0.0

    def run(k0, k1):
        nonlocal checksum
        for k in range(k0, k1):
            if pipe.submitted - pipe.collected >= depth:
                res = pipe.collect()                    # the oldest step in flight is in host memory now
                checksum += float(res["reward"][0]) + float(res[skey][0, 0, 0, 0])
            pipe.submit(host_actions[k], auto_reset=True)
        while pipe.collected < pipe.submitted:
            res = pipe.collect()
            checksum += float(res["reward"][0])

    run(0, warm)
    barrier()
    c0 = eng.counters.clone()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    run(warm, warm + a.steps)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    ms = sharding.max_over_ranks(ms, dev)
    dc = eng.counters - c0
    assert int(eng.counters[5].item()) == rec_sum_p[-1], "host-buffer run left the recorded trajectory"
    out = {"value": n * world * a.steps / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": pipe.h2d_bytes,
           "d2h_bytes_per_step": pipe.d2h_bytes, "ms_per_step": ms / a.steps,
           "mean_perspectives_per_env": float(dc[5].item()) / float(n * a.steps),
           "call": "HostActorPipeline.submit(pinned host actions) / collect() -> host reward, terminal, counts, "
                   "%s; uploads and downloads on their own streams, up to 3 steps in flight"
                   % ("states uint8 (n,2,d,d)" if states == "u8" else
                      "state_rows uint8 (n,2,d,ceil(d/8)) = packbits(states, little)"),
           "policy": "host-supplied actions = the device random policy's choices on the same seed (recorded), "
                     "i.e. the trajectory `value` is timed on"}
    del pipe, eng
    torch.cuda.empty_cache()
    return out


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
