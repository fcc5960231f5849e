#!/usr/bin/env python
"""BASELINE.json configs[3] at the reference's own sample size: d=7 low-error sweep with the importance-sampled
generator (results/small_p_error_test.py: q = d//2+1 forced errors + depolarizing p elsewhere), 2^21 syndromes per p
(the reference ran 2*10^6 per p, results/small_p_error_test.py:13) for p = 5e-4 and 5e-3, 2^19 for p = 5e-2.
Prints one JSON line per p with the failure counts next to both reference runs
(data/evaluation_size_7_p_*.txt and results/evaluation_size_7.txt) and a two-sample z-score.

With --d 5 the same sweep on the second lattice size the reference published (results/evaluation_size_5.txt,
network/converged/Size_5_NN_11_17_Mar_2020_22_33_59.pt, q = 3 forced errors).

  python scripts/c4_big.py [--d 7|5] [--scale 1.0] [--count 10000000]
  torchrun --nproc-per-node 8 ... scripts/c4_big.py --count 10000000      # episodes sharded over the ranks by global index,
                                                                      # sums all-reduced (evaluation._evaluate_one_p)
"""
import argparse
import json
import math
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from toric_rl_decoder_b200 import evaluation, networks, sharding  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
# ground-state rate, P_l, mean steps: (data/evaluation_size_7_p_<p>.txt:2, results/evaluation_size_7.txt) -- 2e6 episodes each
REF = {0.0005: ((0.999976, 3.877189105207568e-12, 4.0), 0.9999735),
       0.005: ((0.9998975, 2.820680442648895e-08, 4.5), 0.9998785),
       0.05: ((0.978193, 1.5482072989164814e-05, 8.6), 0.978138)}
# d=5: results/evaluation_size_5.txt (one run)
REF5 = {0.0005: ((0.998351, 3.594768387450828e-09, 3.0), None),
        0.005: ((0.996427, 3.1588607497089524e-06, 3.2), None),
        0.05: ((0.9159, 0.0007553728078117337, 5.3), None)}
N_REF = 2000000


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--d", type=int, default=7, choices=(5, 7))
    ap.add_argument("--count", type=int, default=0, help="syndromes per p (overrides the built-in sizes)")
    ap.add_argument("--count-high-p", type=int, default=0, help="syndromes at p = 0.05 (default: --count)")
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    rank, world, local = sharding.init_from_env("nccl")
    d = a.d
    weights = {7: "Size_7_NN_11_random_state.pt", 5: "Size_5_NN_11_converged_state.pt"}[d]
    ref_table = REF if d == 7 else REF5
    sizes = ((0.0005, 1 << 21), (0.005, 1 << 21), (0.05, 1 << 19)) if d == 7 else \
        ((0.0005, 1 << 20), (0.005, 1 << 20), (0.05, 1 << 19))
    model = networks.load_state(networks.NN_11(d, 3), os.path.join(GOLD, weights), device="cuda")
    model = networks.pad_channels(model, 8).to(memory_format=torch.channels_last)
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    cfg = {"size": d, "min_qubit_errors": 0, "p_error": 0.1}
    if a.count:
        sizes = tuple((p, (a.count_high_p or a.count) if p >= 0.05 else a.count) for p, _ in sizes)
    for p, n in sizes:
        n = max(1024, int(n * a.scale))
        t0 = time.perf_counter()
        ec, gs, st, mq, hist, nfail, pl, _ = evaluation.prediction_smart(
            model, "toric-code-v0", cfg, d // 2, "cuda", [p], num_of_episodes=n, num_of_steps=75,
            nbr_of_qubit_errors=d // 2 + 1, seed=11, keep_failed=False)
        torch.cuda.synchronize()
        wall = sharding.max_over_ranks(time.perf_counter() - t0, torch.device("cuda", local))
        if rank != 0:
            continue
        (g_ref, pl_ref, st_ref), g_ref2 = ref_table[p]
        f = 1.0 - float(gs[0])

        def z(g_other):
            fo = 1.0 - g_other
            pooled = (f * n + fo * N_REF) / (n + N_REF)
            return (f - fo) / math.sqrt(max(pooled * (1 - pooled), 1e-30) * (1.0 / n + 1.0 / N_REF))
        print(json.dumps({
            "config": "c4 d=%d importance-sampled sweep (NN_11 Size_%d converged weights, TF32 conv, padded channels)" % (d, d),
            "p_error": p, "n_gpus": world, "syndromes": n, "failures": int(round(f * n)), "failure_rate": f,
            "failure_rate_reference_data": 1 - g_ref,
            "failure_rate_reference_results": (1 - g_ref2) if g_ref2 is not None else None,
            "z_vs_reference_data": z(g_ref), "z_vs_reference_results": z(g_ref2) if g_ref2 is not None else None,
            "z_between_the_two_reference_runs": ((1 - g_ref) - (1 - g_ref2)) / math.sqrt(
                max(((1 - g_ref) + (1 - g_ref2)) / 2, 1e-30) * 2.0 / N_REF) if g_ref2 is not None else None,
            "solved": float(ec[0]), "P_l": float(pl[0]), "P_l_reference": pl_ref, "mean_steps": float(st[0]),
            "mean_steps_reference": st_ref, "mean_q": float(mq[0]), "failures_by_initial_weight": {
                str(int(k)): int(v) for k, v in zip(hist[0], hist[2]) if v > 0},
            "syndromes_per_s": n / wall, "seconds": wall}), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
