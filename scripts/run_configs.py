#!/usr/bin/env python
"""BASELINE.json configs other than the headline (configs[1] is bench.py), run at bounded sizes on one GPU.
Prints one JSON line per config; the driver-facing numbers stay in bench.py.

  python scripts/run_configs.py [--only c1,c2,c3,c4,c5] [--scale 1.0]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import toric_rl_decoder_b200 as pkg  # noqa: E402
from toric_rl_decoder_b200 import evaluation, networks  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def emit(**kw):
    print(json.dumps(kw), flush=True)


def gpu_time(fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize()
    return out, e0.elapsed_time(e1) * 1e-3


def c1(scale):
    """configs[0]: d=5, p=0.1 greedy rollouts with Size_5_NN_11 (state_dict of network/Size_5_NN_11.pt)."""
    networks.strict_fp32()
    d = 5
    model = networks.load_state(networks.NN_11(d, 3), os.path.join(GOLD, "Size_5_NN_11_state.pt"), device="cuda")
    cfg = {"size": d, "min_qubit_errors": 0, "p_error": 0.1}
    # single env, the reference's own use (one episode at a time): episodes one after the other, after a warm-up;
    # eager loop, then one CUDA graph per decoding step (evaluation.GraphedGreedyDecoder, the default)
    for use_graph, episodes in ((False, 20), (True, 200)):
        for seed in range(3):
            evaluation.evaluate(model, "toric-code-v0", cfg, d // 2, "cuda", [0.1], num_of_episodes=1, num_of_steps=75,
                                seed=100 + seed, keep_failed=False, use_graph=use_graph)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        tot_steps, solved = 0.0, 0.0
        for seed in range(episodes):
            ec, gs, steps, mq, _ = evaluation.evaluate(model, "toric-code-v0", cfg, d // 2, "cuda", [0.1],
                                                       num_of_episodes=1, num_of_steps=75, seed=seed,
                                                       keep_failed=False, use_graph=use_graph)
            tot_steps += float(steps[0])
            solved += float(ec[0])
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        emit(config="c1 d=5 p=0.1 greedy NN_11 (fp32, TF32 off), single env, episodes one at a time, %s"
             % ("one CUDA graph per decoding step" if use_graph else "eager loop"), episodes=episodes,
             success=solved / episodes, mean_steps=tot_steps / episodes, env_steps_per_s=tot_steps / wall,
             ms_per_episode=1e3 * wall / episodes)
    for n in (int(4096 * scale),):
        evaluation.evaluate(model, "toric-code-v0", cfg, d // 2, "cuda", [0.1], num_of_episodes=min(n, 64),
                            num_of_steps=75, keep_failed=False)
        t0 = time.perf_counter()
        (ec, gs, steps, mq, _), dt = gpu_time(lambda: evaluation.evaluate(
            model, "toric-code-v0", cfg, d // 2, "cuda", [0.1], num_of_episodes=n, num_of_steps=75, seed=1,
            keep_failed=False))
        wall = time.perf_counter() - t0
        emit(config="c1 d=5 p=0.1 greedy NN_11 (fp32, TF32 off)", episodes=n, success=float(ec[0]),
             ground_state=float(gs[0]), mean_steps=float(steps[0]), mean_q=float(mq[0]),
             env_steps_per_s=float(steps[0]) * n / wall, seconds=wall)


def c2(scale):
    """configs[1]: d=7, p=0.1, 65536 envs, seeds {0,1,2}.  Two variants: the actor loop with
    random actions (what bench.py times) and "fresh syndromes only" (reset + perspectives of the new syndromes)."""
    d, n = 7, int(65536 * scale)
    per_persp = 8 * d * d + 4
    for seed in (0, 1, 2):
        eng = pkg.ToricEngine(d, n, seed=seed, p_error=0.1)
        eng.ensure_capacity(eng.worst_case_capacity())
        eng.reset()
        for _ in range(20):
            eng.actor_step()
        c0 = eng.read_counters()
        steps = 200

        def loop():
            for _ in range(steps):
                eng.actor_step()
        _, dt = gpu_time(loop)
        c = eng.read_counters()
        sum_p = c["perspectives"] - c0["perspectives"]
        emit(config="c2 actor loop, random actions", d=d, envs=n, seed=seed, steps=steps,
             mean_perspectives_per_env=sum_p / (n * steps), env_steps_per_s=n * steps / dt,
             step_GBps=(sum_p * per_persp + n * steps * 154) / dt / 1e9, ms_per_step=1e3 * dt / steps)
        eng.reset()
        eng.perspectives()
        iters = 100
        tot = torch.zeros((), dtype=torch.int64, device="cuda")

        def fresh():
            for _ in range(iters):
                eng.reset()
                eng.perspectives()
                tot.add_(eng.offsets[n])
        fresh()                                   # warm-up
        tot.zero_()
        _, dt = gpu_time(fresh)
        sum_p = int(tot.item())
        emit(config="c2 fresh syndromes only (reset + perspectives)", d=d, envs=n, seed=seed, iters=iters,
             mean_perspectives_per_env=sum_p / (n * iters), syndromes_per_s=n * iters / dt,
             GBps=(sum_p * per_persp + n * iters * (168 + 56 + 26 + 8)) / dt / 1e9, ms_per_iter=1e3 * dt / iters)
        del eng


def c3(scale):
    """configs[2]: d=9 batched greedy evaluation with NN_17.  Size_9_NN_17.pt is not in the reference checkout
    (.MISSING_LARGE_BLOBS), so the weights are seeded random-init: rates are NOT comparable with RL_9.txt."""
    d = 9
    torch.manual_seed(0)
    # channel counts padded to a multiple of 8 with zero weights: same function, tensor-core aligned convolutions
    # (profiles/r01_net_forward.jsonl: 2.9e5 -> 4.6e5 perspectives/s under bf16 autocast)
    model = networks.pad_channels(networks.NN_17(d, 3).cuda().eval(), 8).to(memory_format=torch.channels_last)
    cfg = {"size": d, "min_qubit_errors": 0, "p_error": 0.1}
    n = int(2048 * scale)

    class Amp(torch.nn.Module):
        def __init__(self, m):
            super().__init__()
            self.m = m

        def forward(self, x):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return self.m(x.contiguous(memory_format=torch.channels_last)).float()

    for name, mdl, steps in (("bf16 autocast", Amp(model), 20),):
        t0 = time.perf_counter()
        ec, gs, st, mq, _ = evaluation.evaluate(mdl, "toric-code-v0", cfg, d // 2, "cuda", [0.1], num_of_episodes=n,
                                                num_of_steps=steps, seed=2, keep_failed=False)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        emit(config="c3 d=9 p=0.1 greedy NN_17 random-init (%s), %d-step cap" % (name, steps), syndromes=n,
             success=float(ec[0]), ground_state=float(gs[0]), mean_steps=float(st[0]), syndromes_per_s=n / wall,
             seconds=wall, note="weights unavailable in the reference checkout; network forward dominates")


def c4(scale):
    """configs[3]: d=7 low-error sweep with the reference's importance-sampled generator (q = d//2+1 fixed errors
    + depolarizing p on the rest, results/small_p_error_test.py) and Size_7_NN_11_random weights."""
    d = 7
    model = networks.load_state(networks.NN_11(d, 3), os.path.join(GOLD, "Size_7_NN_11_random_state.pt"), device="cuda")
    model = networks.pad_channels(model, 8).to(memory_format=torch.channels_last)
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    cfg = {"size": d, "min_qubit_errors": 0, "p_error": 0.1}
    ref = {0.0005: (0.999976, 3.877189105207568e-12, 4.0), 0.005: (0.9998975, 2.820680442648895e-08, 4.5),
           0.05: (0.978193, 1.5482072989164814e-05, 8.6)}       # data/evaluation_size_7_p_*.txt
    n = int((1 << 17) * scale)
    for p in (0.0005, 0.005, 0.05):
        t0 = time.perf_counter()
        ec, gs, st, mq, hist, nfail, pl, _ = evaluation.prediction_smart(
            model, "toric-code-v0", cfg, d // 2, "cuda", [p], num_of_episodes=n, num_of_steps=75,
            nbr_of_qubit_errors=d // 2 + 1, seed=4, keep_failed=False)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        g_ref, pl_ref, st_ref = ref[p]
        sigma = float(np.sqrt(max(g_ref * (1 - g_ref), 1e-12) / n))
        emit(config="c4 d=7 importance-sampled sweep (TF32 conv)", p_error=p, syndromes=n,
             ground_state=float(gs[0]), ground_state_reference=g_ref,
             sigmas_from_reference=(float(gs[0]) - g_ref) / sigma, P_l=float(pl[0]), P_l_reference=pl_ref,
             mean_steps=float(st[0]), mean_steps_reference=st_ref, mean_q=float(mq[0]),
             syndromes_per_s=n / wall, seconds=wall)


def c5(scale):
    """configs[4]: large-lattice stress.  Env state for the whole shard stays resident; the perspective tensor
    is produced in env chunks sized to the output buffer (the consumer reads it between chunks)."""
    only_d = os.environ.get("C5_ONLY_D")
    for d, p in ((15, 0.05), (15, 0.1), (15, 0.2), (25, 0.05), (25, 0.1), (25, 0.2)):
        if only_d and int(only_d) != d:
            continue
        n = int((1 << 17) * scale) if d == 15 else int((1 << 15) * scale)
        eng = pkg.ToricEngine(d, n, seed=9, p_error=p)
        budget = 24 << 30
        per_persp = 8 * d * d + 4
        eng.reset()
        eng.perspectives_count()
        pbar = eng.total_perspectives() / n
        chunk = max(256, min(n, int(budget / (per_persp * pbar * 1.5)) // 256 * 256))
        eng.ensure_capacity(int(chunk * pbar * 1.5))
        steps = 6

        def run():
            tot = 0
            for _ in range(steps):
                eng.actor_pre()                                   # step + auto-reset for every env of the shard
                for e0 in range(0, n, chunk):
                    e1 = min(n, e0 + chunk)
                    eng.perspectives(syn=eng.syn[e0:e1], n=e1 - e0)
            return tot

        eng.actor_pre()                                           # warm-up, with an overflow check per chunk
        for e0 in range(0, n, chunk):
            eng.perspectives(syn=eng.syn[e0:min(n, e0 + chunk)])
            assert int(eng.overflow.item()) == 0, "perspective buffer too small"
        c0 = eng.read_counters()
        _, dt = gpu_time(run)
        assert int(eng.overflow.item()) == 0
        c = eng.read_counters()
        sum_p = (c["perspectives"] - c0["perspectives"])
        nbytes = sum_p * per_persp
        emit(config="c5 large lattice, random actions", d=d, p_error=p, envs=n, env_chunk=chunk,
             mean_perspectives_per_env=sum_p / (n * steps), env_steps_per_s=n * steps / dt,
             perspective_GBps=nbytes / dt / 1e9, seconds=dt)
        del eng
        torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="c1,c2,c3,c4,c5")
    ap.add_argument("--scale", type=float, default=1.0)
    a = ap.parse_args()
    emit(gpu=torch.cuda.get_device_name(0), torch=torch.__version__)
    for name in a.only.split(","):
        try:
            {"c1": c1, "c2": c2, "c3": c3, "c4": c4, "c5": c5}[name](a.scale)
        except Exception as ex:  # keep going: each config is independent
            emit(config=name, error=repr(ex))


if __name__ == "__main__":
    main()
