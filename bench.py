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
    ap.add_argument("--burnin", type=int, default=152,
                    help="untimed, uncounted steps before --warmup (episode clocks are pre-staggered over "
                         "[0, max_steps], so the timed region sees the stationary reset rate whatever --steps is)")
    ap.add_argument("--no-stagger", action="store_true", help="keep every env's episode clock at 0 after reset")
    ap.add_argument("--cpu-envs", type=int, default=0,
                    help="envs the CPU arm walks per pass; 0 = --envs, the batch the GPU arm steps (one slice per thread)")
    ap.add_argument("--repeats", type=int, default=5, help="how often the --steps region is timed (value = the first)")
    ap.add_argument("--kernel-samples", type=int, default=16,
                    help="least number of event-timed writer launches behind roofline.kernel_ms (a post-region pass "
                         "on the same stationary state supplies what the timed region did not)")
    ap.add_argument("--e2e-min-steps", type=int, default=1024,
                    help="the e2e regions run max(--steps, this) steps (bounded by 1 GiB of pinned actions)")
    ap.add_argument("--ref-step-seconds", type=float, default=0.25,
                    help="--impl reference: CPU time one reference 'step' (a whole number of passes over the "
                         "--cpu-envs sample) is calibrated to")
    ap.add_argument("--ref-warm-seconds", type=float, default=2.0)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--out-dtype", default="float32", choices=("float32", "bfloat16", "float16", "uint8"),
                    help="element type of the perspective tensor; float32 = the reference's (headline)")
    ap.add_argument("--sample-every", type=int, default=8,
                    help="time the writer kernel with CUDA events on every k-th step of the timed region")
    ap.add_argument("--dump-launches", action="store_true",
                    help="add the sum P and algorithmic bytes of every timed writer launch to the roofline object "
                         "(to match an ncu capture of one launch of the same command, scripts/gpu_traffic.sh)")
    ap.add_argument("--single-pass", type=int, default=-1, choices=(-1, 0, 1),
                    help="1: the one-kernel step (TC_ACTOR_SINGLE_PASS), 0: the three-kernel step, -1: the engine's default")
    ap.add_argument("--ref-api-steps", type=int, default=8,
                    help="timed iterations of the reference actor's loop body through the reference-named mirrors "
                         "(e2e_reference_api, N = 1 only); 0 = skip")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    a = ap.parse_args()
    if a.cpu_envs <= 0:
        a.cpu_envs = a.envs
    return a


NOMINAL_HBM_GBS = 8000.0      # B200 HBM3e nominal, the figure north_star's ">= 60 %" is quoted on


def bench_config(a, world):
    """The `config` object of BOTH arms (the driver compares them): only what the command line fixes.  What a run
    measures about its workload (perspectives per env, bytes per env-step) goes into `workload_stats`."""
    return {"workload": workload_name(a), "d": a.d, "p_error": a.p, "envs_per_gpu": a.envs,
            "global_envs": a.envs * world, "parallelism": "env-sharded x%d, no data-path collective" % world,
            "out_dtype": a.out_dtype,
            "regime": ("stationary: episode clocks staggered over [0, 75], untimed burn-in of at least one episode "
                       "length before the warm-up") if not a.no_stagger else
                      "all envs start their first episode at the first warm-up step",
            "l2": "no flush: every step streams a fresh perspective tensor far larger than the 126 MB L2 "
                  "(bytes per step: workload_stats)"}


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
                "reasons": sorted(self.reasons), "samples": len(self.sm), "how": "NVML, 2 ms period, over the timed region and its repeats"}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------------------------
def _cpu_loop(a, n_envs, threads=0):
    """The oracle's C port of the actor loop on the host cores, brought to the same stationary regime as the GPU
    arm: episode clocks staggered over [0, max_steps], then burned in for one full episode length."""
    import numpy as np
    from oracle import c_oracle as C
    loop = C.ActorLoop(a.d, n_envs, a.p, seed=a.seed, nthreads=threads)
    if not a.no_stagger:
        base = 0
        for sl in loop.slices:
            sl.steps[:] = stagger_clocks(np.arange(base, base + sl.n), 75)
            base += sl.n
    return loop


def stagger_clocks(env_ids, max_steps):
    """Episode clock of env e at the start: a fixed pseudo-random value in [0, max_steps] (Knuth hash of the
    global env id) -- in a long-running actor the envs are at independent points of their episodes."""
    import numpy as np
    h = (np.asarray(env_ids, dtype=np.uint64) * np.uint64(2654435761)) & np.uint64(0xffffffff)
    return ((h * np.uint64(max_steps + 1)) >> np.uint64(32)).astype(np.int32)


def _timed_cpu(loop, warm_seconds, seconds, min_passes=1):
    """Warm for >= warm_seconds (thread pool, caches, pages, stationary regime), then time >= `seconds`."""
    t0 = time.perf_counter()
    loop.run(2)
    per_pass = max((time.perf_counter() - t0) / 2, 1e-6)
    warm_passes = max(76, int(warm_seconds / per_pass))       # at least one full episode length
    t0 = time.perf_counter()
    loop.run(warm_passes)
    per_pass = max((time.perf_counter() - t0) / warm_passes, 1e-6)
    passes = max(min_passes, int(round(seconds / per_pass)))
    return per_pass, passes


def cpu_arm(a, seconds, n_envs, threads=0):
    """Time the oracle's C port of the same loop on the host cores (bounded sample)."""
    loop = _cpu_loop(a, n_envs, threads)
    _, steps = _timed_cpu(loop, a.ref_warm_seconds, seconds, 3)
    c0 = loop.counters.copy()
    t0 = time.perf_counter()
    loop.run(steps)
    dt = time.perf_counter() - t0
    c = loop.counters - c0
    return {"value": float(c[0]) / dt, "unit": UNIT, "cores": loop.nthreads, "kind": "port",
            "sample": "%d envs x %d steps, d=%d p=%g, C port of EnvSet.step + generatePerspectiveBatch + float32 "
                      "concatenate (oracle/toric_oracle.c), one slice per thread, %.1f s after a >= %.0f s warm-up in "
                      "the stationary regime; mean perspectives/env %.1f, %d episodes ended"
                      % (n_envs, steps, a.d, a.p, dt, a.ref_warm_seconds, float(c[3]) / max(float(c[0]), 1.0),
                         int(c[1])),
            "seconds": dt, "steps": steps, "envs": n_envs}


def run_reference(a):
    """--impl reference: the CPU arm.  The reference is pure Python/numba and cannot travel to the GPU box,
    so this times the oracle's C port of the same loop (kind "port") with every host core.  One reference
    'step' = a bounded sample: `passes` passes over --cpu-envs environments (each pass advances every env of
    the sample once), with `passes` calibrated so a step takes ~--ref-step-seconds; the worker threads are
    created once and warmed for >= --ref-warm-seconds before the --warmup steps."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    loop = _cpu_loop(a, a.cpu_envs, 0)
    per_pass, passes = _timed_cpu(loop, a.ref_warm_seconds, a.ref_step_seconds, 1)
    for _ in range(max(a.warmup, 1)):
        loop.run(passes)
    c0 = loop.counters.copy()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        loop.run(passes)
    dt = time.perf_counter() - t0
    c = loop.counters - c0
    value = float(c[0]) / dt
    sample = ("%d reference steps x %d passes over %d envs (%.2f s per step, %.1f s timed), d=%d p=%g, C port of "
              "EnvSet.step + generatePerspectiveBatch + float32 concatenate (oracle/toric_oracle.c), one slice per "
              "thread, threads created once, >= %.0f s warm-up in the stationary regime; mean perspectives/env %.1f, "
              "%d episodes ended"
              % (a.steps, passes, a.cpu_envs, dt / a.steps, dt, a.d, a.p, a.ref_warm_seconds,
                 float(c[3]) / max(float(c[0]), 1.0), int(c[1])))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic", "config": bench_config(a, max(a.gpus, 1)),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": loop.nthreads, "kind": "port", "sample": sample,
                             "passes_per_step": passes, "cpu_envs": a.cpu_envs},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "nonce": _nonce()}
    print(json.dumps(line), flush=True)


def _nonce():
    return "%s-%d" % (os.urandom(6).hex(), time.time_ns())


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
    # N > 1: pin the process to the cores next to its GPU BEFORE any pinned host buffer exists, so that eight ranks
    # downloading at once do it into NUMA-local memory.  Not at N = 1: the cpu_baseline leg of that run needs every core.
    numa_cpus = sharding.bind_to_gpu_numa(local) if world > 1 else 0

    out_dtype = getattr(torch, a.out_dtype)
    esz = torch.empty((), dtype=out_dtype).element_size()
    ekw = {} if a.single_pass < 0 else {"single_pass": bool(a.single_pass)}
    eng = pkg.ToricEngine(d, n, device=dev, seed=a.seed, env_base=rank * n, p_error=a.p,
                          out_dtype=out_dtype, **ekw)                                       # envs shard by rank
    one_kernel = eng.single_pass and out_dtype is torch.float32 and d <= pkg.engine.SINGLE_PASS_MAX_D
    eng.ensure_capacity(eng.worst_case_capacity())
    max_steps = eng.max_steps
    prepare(a, eng)                             # reset, staggered episode clocks, untimed burn-in

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    write_peak = measure_write_peak(dev)
    # warm-up: the fused call, and once the split form with its event records and the device-side copy of sum P, so that
    # nothing is used for the first time (lazy module loads, allocator growth) inside a short timed region
    warm_ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    warm_tot = torch.zeros(1, dtype=torch.int64, device=dev)
    for w in range(a.warmup):
        if w == 0 and not one_kernel:
            eng.actor_pre()
            warm_ev[0].record()
            eng.persp_write()
            warm_ev[1].record()
            warm_tot[0].copy_(eng.offsets[n], non_blocking=True)
        else:
            eng.actor_step()
    barrier()
    if not one_kernel and a.warmup > 0:
        warm_ev[0].elapsed_time(warm_ev[1])
    eng.counters.zero_()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # The step is the fused call a user makes (tc_actor_step: three back-to-back launches).  Every
    # `a.sample_every`-th step is issued as its two halves with CUDA events around the writer launch, which
    # gives the dominant kernel's duration inside the timed region without putting two event records (and a
    # break of the programmatic dependent launch chain) into every step.
    sampled = [] if one_kernel else [k for k in range(a.steps) if k % a.sample_every == 0]
    evb = {k: torch.cuda.Event(enable_timing=True) for k in sampled}
    evc = {k: torch.cuda.Event(enable_timing=True) for k in sampled}
    totals = torch.zeros(max(len(sampled), 1), dtype=torch.int64, device=dev)      # sum P of the sampled launches
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
        ctr = eng.counters.clone()                             # the counters of the region `value` is quoted on
        sum_p_local = int(ctr[5].item())
        per_launch_sum_p = [int(x) for x in totals.tolist()] if a.dump_launches else None
        ms = sharding.max_over_ranks(ms, dev)                  # time = max over ranks

        # the same region again (same stationary state, it just keeps running): run-to-run spread of `value`; the
        # clock sampler stays on, so that a short --steps region still gets more than a couple of NVML samples
        repeat_ms = [ms]
        for _ in range(max(a.repeats, 1) - 1):
            barrier()
            ev0.record()
            for k in range(a.steps):
                eng.actor_step()
            ev1.record()
            barrier()
            repeat_ms.append(sharding.max_over_ranks(ev0.elapsed_time(ev1), dev))

    # writer-launch samples: those of the timed region plus a post-region pass in the split form until there are
    # at least --kernel-samples of them, whatever --steps was
    if one_kernel:          # the step IS the kernel: its launches run back to back, so step time = kernel time
        launch_ms = [ms / a.steps]
        sampled_sum_p = float(sum_p_local) / a.steps
        n_launches = a.steps
    else:
        launch_ms = [evb[k].elapsed_time(evc[k]) for k in sampled]
        launch_sum_p = [int(x) for x in totals.tolist()][:len(sampled)]
        extra = max(a.kernel_samples - len(sampled), 0)
        xb = [torch.cuda.Event(enable_timing=True) for _ in range(extra)]
        xc = [torch.cuda.Event(enable_timing=True) for _ in range(extra)]
        xtot = torch.zeros(max(extra, 1), dtype=torch.int64, device=dev)
        for k in range(extra):
            eng.actor_pre()
            xb[k].record()
            eng.persp_write()
            xc[k].record()
            xtot[k].copy_(eng.offsets[n], non_blocking=True)
        torch.cuda.synchronize()
        launch_ms += [xb[k].elapsed_time(xc[k]) for k in range(extra)]
        launch_sum_p += [int(x) for x in xtot.tolist()][:extra]
        sampled_sum_p = float(sum(launch_sum_p)) / len(launch_sum_p)
        n_launches = len(launch_ms)
    write_ms = sum(launch_ms) / len(launch_ms)
    srt = sorted(launch_ms)
    assert int(eng.overflow.item()) == 0
    sharding.allreduce_counters(ctr)                       # the path's only collective: counter all-reduce
    c = {name: int(ctr[i].item()) for i, name in enumerate(pkg.engine.COUNTER_NAMES)}
    total_steps = c["env_steps"]
    assert total_steps == n * a.steps * world
    if not a.no_stagger and a.steps >= 8:
        assert c["episodes"] > 0, "the timed region saw no episode end: not the stationary regime"
    value = total_steps / (ms * 1e-3)
    p_bar = sum_p_local / float(n * a.steps)

    peak, peak_src = load_peaks()
    wbytes = (b_alg(d, n, sampled_sum_p, esz) if one_kernel else
              write_kernel_bytes(d, n, sampled_sum_p, esz))    # the launches that were timed
    achieved = wbytes / (write_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": ("fused_step_kernel<%d,%s>" if one_kernel else "persp_write_kernel<%d,%s>")
                % (d, a.out_dtype), "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                # the kernel is a write-only stream and `peak` is a COPY (read+write) figure, so frac can exceed 1;
                # write_peak is the write-only ceiling measured live (cudaMemsetAsync / float32 fill, best of 6)
                "write_peak": write_peak, "frac_of_write_peak": achieved / write_peak if write_peak else None,
                "nominal": NOMINAL_HBM_GBS, "frac_of_nominal": achieved / NOMINAL_HBM_GBS,
                "kernel_ms": write_ms, "kernel_ms_min": srt[0], "kernel_ms_median": srt[len(srt) // 2],
                "kernel_ms_max": srt[-1], "kernel_share_of_step": write_ms * a.steps / ms,
                "kernel_launches_timed": n_launches,
                "kernel_launches_in_region": a.steps if one_kernel else len(sampled)}
    step_achieved = b_alg(d, n, sum_p_local / a.steps, esz) / (ms / a.steps * 1e-3) / 1e9
    roofline["step_achieved"] = step_achieved
    roofline["step_frac"] = {"of_measured_copy_peak": step_achieved / peak,
                             "of_write_peak": step_achieved / write_peak if write_peak else None,
                             "of_nominal": step_achieved / NOMINAL_HBM_GBS}
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
    e2e = e2e_open = e2e_rows = None
    if not a.no_e2e:
        del eng
        torch.cuda.empty_cache()

        def make_engine():
            e = pkg.ToricEngine(d, n, device=dev, seed=a.seed, env_base=rank * n, p_error=a.p, out_dtype=out_dtype,
                                **ekw)
            e.ensure_capacity(e.worst_case_capacity())
            return e
        # long enough to be a measurement of its own (>= a few hundred ms at the BENCH config), bounded by 1 GiB of
        # pinned recorded actions
        e2e_steps = max(a.steps, min(a.e2e_min_steps, (1 << 30) // (n * 16)))
        recorded = record_device_policy(a, make_engine, max(3, a.warmup) + e2e_steps)
        # `e2e` is the reference's own host contract: EnvSet.step hands back the uint8 (n,2,d,d) states, one byte per
        # stabiliser, every step (src/EnvSet.py:38-47) -- that array is downloaded whole.  The cheaper host formats
        # (row-packed states behind a LazyStates view; three steps in flight) are reported under their own keys.
        e2e = run_e2e(a, make_engine, recorded, dev, dist, barrier, world, "u8", 1, e2e_steps)
        e2e_rows = run_e2e(a, make_engine, recorded, dev, dist, barrier, world, "lazy", 1, e2e_steps)
        e2e_open = run_e2e(a, make_engine, recorded, dev, dist, barrier, world, "lazy", 3, e2e_steps)
        if world > 1:
            for o in (e2e, e2e_rows, e2e_open):
                o["numa_bound_cpus"] = numa_cpus
            # the uint8 states are 7 MB per rank and step; when that form falls behind the row-packed one, the
            # limiter is the host's aggregate device->host rate, not the GPUs (no collective is involved)
            if e2e["value"] < 0.9 * e2e_rows["value"]:
                e2e["limit"] = ("host D->H rate: %d ranks x %.1f MB per step into one host's memory (%.0f GB/s asked "
                                "at the row-packed form's step time)"
                                % (world, e2e["d2h_bytes_per_step"] / 1e6,
                                   world * e2e["d2h_bytes_per_step"] / (e2e_rows["ms_per_step"] * 1e-3) / 1e9))

    ref_api = None
    if world == 1 and a.ref_api_steps > 0 and not a.no_e2e:
        ref_api = {"loop": "src/Actor_mp.py:104-183 literally, network replaced by a uniformly random perspective and "
                           "operator: generatePerspectiveBatch(state) -> tc_select_action(epsilon=1) -> EnvSet.step -> "
                           "generateTransitionParallel -> resetTerminalEnvs(idx, p_errors) -> next_state[idx] = "
                           "reset_states; wall clock around the loop, device idle at both ends",
                   "numpy_compat_true": run_reference_api(a, dev, True, a.ref_api_steps),
                   "numpy_compat_false": run_reference_api(a, dev, False, a.ref_api_steps)}

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
                "config": bench_config(a, world),
                "workload_stats": {"mean_perspectives_per_env": p_bar, "burnin_steps": a.burnin,
                                   "max_steps_per_episode": max_steps,
                                   "writer_bytes_per_step": wbytes,
                                   "b_alg_bytes_per_env_step": b_alg(d, n, sum_p_local / a.steps, esz) / n},
                "value_repeats": [n * a.steps * world / (m * 1e-3) for m in repeat_ms],
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "e2e_rows": e2e_rows,
                "e2e_open_loop": e2e_open, "e2e_reference_api": ref_api,
                "gpu_launches": (1 if one_kernel else 3) * a.steps,
                "step_form": "single-pass (1 kernel)" if one_kernel else "three kernels (step+count, scan, writer)",
                "clocks": clk.summary(), "nonce": _nonce(),
                "counters": {k: c[k] for k in ("episodes", "solved", "ground_kept", "timeouts", "perspectives")}}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def prepare(a, eng):
    """Fresh syndromes, then the stationary regime of a long-running actor: every env's episode clock starts at a
    fixed pseudo-random point of [0, max_steps] (so time-outs, hence auto-resets, are spread evenly over the steps
    instead of hitting all envs in the same step) and --burnin untimed steps of the device's random policy are
    run and their counters dropped."""
    import numpy as np
    import torch
    eng.reset()
    if not a.no_stagger:
        clocks = stagger_clocks(np.arange(eng.env_base, eng.env_base + eng.n), eng.max_steps)
        eng.steps.copy_(torch.from_numpy(clocks))
        for _ in range(a.burnin):
            eng.actor_step()
    eng.counters.zero_()


def measure_write_peak(dev):
    """Write-only HBM ceiling, measured live: the better of cudaMemsetAsync (torch zero_) and a vectorised float32 fill
    over 2 GiB, best of 6 each.  (A uint8 fill_ kernel reaches only ~3.9 TB/s on B200 and is not used.)"""
    import torch
    try:
        buf = torch.empty(1 << 29, dtype=torch.float32, device=dev)
        best = None
        for fn in (buf.zero_, lambda: buf.fill_(1.0)):
            for _ in range(6):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1)
                best = ms if best is None or ms < best else best
        del buf
        torch.cuda.empty_cache()
        return (2 << 30) / (best * 1e-3) / 1e9
    except Exception:
        return None


def record_device_policy(a, make_engine, total_steps):
    """The device's random-policy trajectory (the one `value` times), recorded once: actions (steps,n,4) int32 in
    pinned host memory plus the running perspective total after every step."""
    import torch
    rec = make_engine()
    prepare(a, rec)
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


def run_e2e(a, make_engine, recorded, dev, dist, barrier, world, states, depth, steps):
    """Same metric through the host-facing call (toric_rl_decoder_b200.HostActorPipeline): every step copies
    that step's actions (n,4) int32 from pinned host memory, runs the fused actor step, and reads rewards,
    terminals, perspective counts and the new states back to pinned host memory; every step's results are read
    on the host inside the timed region (reward[0], the whole counts vector -- summed and checked against the
    recorded trajectory -- and env 0's state through the `states` view).

    depth=1 is the CLOSED loop of the reference's EnvSet.step contract (src/Actor_mp.py:116-125): the action of
    step k+1 is submitted only after step k's results have been collected and read on the host.  The step's
    results are final before the perspective writer runs, so their download (and the host's turn-around) overlaps
    the writer of the same step.  depth=3 (`e2e_open_loop`) keeps three steps in flight, i.e. no action depends
    on a downloaded result.

    states="u8" (`e2e`, the headline): the reference's own contract -- the one-byte-per-stabiliser (n,2,d,d) array is
    expanded on the device and downloaded whole every step; at 8 ranks the box's aggregate device->host rate
    (~90 GB/s, profiles/r01_e2e_scale_probe.txt), not the GPUs, bounds that form.
    states="lazy" (`e2e_rows`, `e2e_open_loop`): the states cross PCIe row-packed (ceil(d/8) bytes per lattice row)
    and the host sees the reference's (n,2,d,d) array as a LazyStates view that expands only the envs it indexes.

    The host policy replays the actions the device's random policy chose on the same seed (recorded beforehand,
    record_device_policy), so the envs walk exactly the trajectory `value` was timed on -- same resets, same
    perspectives per env.  The perspective tensor stays on the device, where the reference also puts it
    (`.to(device)`, src/numba/util_actor.py:39)."""
    import torch
    from toric_rl_decoder_b200 import HostActorPipeline, sharding
    n = a.envs
    host_actions, rec_sum_p = recorded
    warm = host_actions.shape[0] - steps
    eng = make_engine()
    prepare(a, eng)
    pipe = HostActorPipeline(eng, depth=depth, states=states)
    checksum = 0.0
    collected = 0
    prev_total = 0

    def read(res):
        nonlocal checksum, collected, prev_total
        total = prev_total + int(res["counts"].sum(dtype="int64"))
        assert total == rec_sum_p[collected], "host-buffer run left the recorded trajectory"
        prev_total = total
        collected += 1
        checksum += float(res["reward"][0]) + float(res["states"][0][0, 0, 0])

    def run(k0, k1):
        for k in range(k0, k1):
            if pipe.submitted - pipe.collected >= depth:
                read(pipe.collect())                    # the oldest step in flight is in host memory now
            pipe.submit(host_actions[k], auto_reset=True)
        while pipe.collected < pipe.submitted:
            read(pipe.collect())

    run(0, warm)
    barrier()
    c0 = eng.counters.clone()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record()
    run(warm, warm + steps)
    ev1.record()
    torch.cuda.synchronize()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    barrier()
    ms_raw = ev0.elapsed_time(ev1)
    ms = sharding.max_over_ranks(ms_raw, dev)
    dc = eng.counters - c0
    assert int(eng.counters[5].item()) == rec_sum_p[-1], "host-buffer run left the recorded trajectory"
    out = {"value": n * world * steps / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": pipe.h2d_bytes,
           "d2h_bytes_per_step": pipe.d2h_bytes, "ms_per_step": ms / steps, "steps": steps,
           "elapsed_ms_raw": ms_raw, "wall_ms": wall_ms, "nonce": _nonce(),
           "steps_in_flight": depth, "loop": "closed" if depth == 1 else "open",
           "mean_perspectives_per_env": float(dc[5].item()) / float(n * steps),
           "episodes": int(dc[1].item()),
           "call": "HostActorPipeline(depth=%d, states=%r).submit(pinned host actions) / collect() -> host reward, "
                   "terminal, counts, %s; uploads and downloads on their own streams; %s"
                   % (depth, states,
                      "states uint8 (n,2,d,d) downloaded whole" if states == "u8" else
                      "states as a LazyStates view over the row-packed download (ceil(d/8) bytes per lattice row), "
                      "expanded on the host for the envs that are indexed",
                      "the next action is submitted only after this step's results were collected and read"
                      if depth == 1 else "up to %d steps in flight" % depth),
           "host_touches_per_step": "reward[0]; counts.sum() over all n envs (checked against the recorded "
                                    "trajectory); states[0] expanded through the view",
           "policy": "host-supplied actions = the device random policy's choices on the same seed (recorded), "
                     "i.e. the trajectory `value` is timed on"}
    del pipe, eng
    torch.cuda.empty_cache()
    return out


def run_reference_api(a, dev, numpy_compat, steps, warm=2):
    """The loop body of the reference's actor (src/Actor_mp.py:104-183) through the mirrors that carry the
    reference's names, as a reference user would write it after changing only the imports: numpy `state` in,
    `EnvSetB200.step`, `generateTransitionParallel`, `resetTerminalEnvs`, `next_state[idx] = reset_states`.  The
    Q-network is left out (a uniformly random perspective and operator are drawn instead, by tc_select_action with
    epsilon = 1 over a zero Q table), everything else is the reference's sequence.

    numpy_compat=True hands the caller the reference's dtypes: every step converts 65 536 x 98 stabilisers to int64
    on the host twice over (EnvSet.step's states, 51 MB, and the transition records, 2 x 51 MB) -- that conversion,
    single-threaded numpy, is the bound of this form, not the GPU.  numpy_compat=False keeps states, rewards and
    terminals on the device as tensors; the states still make the round trip device -> uint8 tensor -> pack ->
    perspectives that the array-in / array-out signatures imply."""
    import numpy as np
    import torch
    import toric_rl_decoder_b200 as pkg
    d, n, g = a.d, a.envs, a.d // 2
    es = pkg.EnvSetB200({"size": d, "p_error": a.p}, n, device=dev, seed=a.seed, numpy_compat=numpy_compat)
    trans_type = pkg.transition_type(d) if numpy_compat else None
    state = es.resetAll()
    max_actions = 75
    if numpy_compat:
        steps_per_episode = stagger_clocks(np.arange(n), max_actions).astype(np.int64)
    else:
        steps_per_episode = torch.from_numpy(stagger_clocks(np.arange(n), max_actions).astype(np.int64)).to(dev)
    sum_p = resets = 0
    t0 = 0.0
    for k in range(warm + steps):
        if k == warm:
            torch.cuda.synchronize()
            sum_p = resets = 0
            t0 = time.perf_counter()
        steps_per_episode += 1
        persp, pos, counts, offsets = pkg.generatePerspectiveBatch(g, d, state, device=dev, raw_positions=True)
        q = torch.zeros((persp.shape[0], 3), dtype=torch.float32, device=dev)
        step_no = torch.full((n,), k, dtype=torch.int32, device=dev)
        act, q_values = pkg.select_action_device(q, pos, offsets, n, epsilon=1.0, steps=step_no, seed=a.seed)
        action = act.cpu().numpy().astype(np.int64) if numpy_compat else act
        next_state, reward, terminal_state, _ = es.step(action)
        transition = pkg.generateTransitionParallel(action, reward, state, next_state, terminal_state, g, trans_type,
                                                    device=dev)
        del transition
        too_many_steps = steps_per_episode > max_actions
        if numpy_compat:
            idx = np.argwhere(np.logical_or(terminal_state, too_many_steps)).flatten()
        else:
            idx = torch.nonzero(terminal_state | too_many_steps).flatten()
        if len(idx):
            p_errors = np.full(len(idx), a.p) if numpy_compat else None
            reset_states = es.resetTerminalEnvs(idx, p_errors=p_errors)
            next_state[idx] = reset_states
            steps_per_episode[idx] = 0
        sum_p += int(persp.shape[0])
        resets += int(len(idx))
        state = next_state
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"value": n * steps / dt, "unit": UNIT, "ms_per_step": 1e3 * dt / steps, "steps": steps, "warmup": warm,
            "mean_perspectives_per_env": sum_p / float(n * steps), "resets_per_step": resets / float(steps),
            "host_conversion_bytes_per_step": (3 * n * 2 * d * d * 8) if numpy_compat else 0}


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
