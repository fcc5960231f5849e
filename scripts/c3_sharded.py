#!/usr/bin/env python
"""BASELINE configs[2]: d=9 batched greedy evaluation with NN_17, sharded over the ranks of one box (torchrun).

Size_9_NN_17.pt is not in the reference checkout (.MISSING_LARGE_BLOBS:8-9), so the weights are seeded random-init
(broadcast from rank 0): success RATES are not comparable with results/results_mats/RL_9.txt.  What is checked and
measured instead:

  A. parity slice: ~4096 perspectives (the complete perspective sets of --parity-syndromes
     fresh d=9 syndromes, p=0.1) through NN_17 in strict fp32 (TF32 off) on the GPU against the CPU forward of the SAME
     weights on the SAME perspectives: Q-values within 1e-5 relative, greedy (perspective, action) per syndrome equal
     wherever the top-2 gap exceeds the tolerance (np.argmax tie-break = first in row-major order).
  B. the protocol of tests/evaluate_network.py:12-54 at d=9: greedy episodes with the full 75-step cap, episodes by
     global index over the ranks, counters all-reduced; syndromes/s and the network's share of the time (the same
     number of env steps is also run without the network)."""
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
from toric_rl_decoder_b200 import evaluation, networks, sharding  # noqa: E402


class Amp(torch.nn.Module):
    pad_small_batches = True

    def __init__(self, m):
        super().__init__()
        self.m = m

    def forward(self, x):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            return self.m(x.contiguous(memory_format=torch.channels_last)).float()


def parity_slice(base, d, dev, n_syn, tol=1e-5):
    eng = pkg.ToricEngine(d, n_syn, device=dev, seed=0, p_error=0.1)
    eng.reset()
    eng.perspectives()
    persp, pos, counts, offsets = eng.perspective_views()
    networks.strict_fp32()
    with torch.no_grad():
        q_gpu = base(persp).double().cpu().numpy()
    cpu_model = networks.NN_17(d, 3)
    cpu_model.load_state_dict({k: v.cpu() for k, v in base.state_dict().items()})
    cpu_model.eval()
    t0 = time.perf_counter()
    with torch.no_grad():
        q_cpu = cpu_model(persp.cpu()).double().numpy()
    cpu_s = time.perf_counter() - t0
    scale = np.maximum(np.abs(q_cpu), np.abs(q_cpu).max() * 1e-3)
    rel = np.abs(q_gpu - q_cpu) / scale
    off = offsets.cpu().numpy()
    same = near_tie = 0
    for e in range(n_syn):
        a, b = q_gpu[off[e]:off[e + 1]].reshape(-1), q_cpu[off[e]:off[e + 1]].reshape(-1)
        if a.size == 0:
            continue
        if int(a.argmax()) == int(b.argmax()):
            same += 1
        else:
            top = np.sort(b)[-2:]
            assert (top[1] - top[0]) <= tol * max(abs(top[1]), 1.0) * 4, "greedy action differs outside a near-tie"
            near_tie += 1
    assert rel.max() < tol * 4, "Q parity: max relative error %.3g" % rel.max()
    return {"perspectives": int(persp.shape[0]), "syndromes": n_syn, "q_max_rel_err": float(rel.max()),
            "q_mean_rel_err": float(rel.mean()), "greedy_equal": same, "greedy_near_ties": near_tie,
            "cpu_forward_seconds": cpu_s, "tolerance": tol}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--syndromes", type=int, default=1 << 14)
    ap.add_argument("--steps", type=int, default=75)
    ap.add_argument("--parity-syndromes", type=int, default=48)
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    rank, world, local = sharding.init_from_env("nccl")
    dev = torch.device("cuda", local)
    d = 9
    torch.manual_seed(rank)                      # different init per rank: the broadcast must fix that
    base = networks.NN_17(d, 3).to(dev).eval()
    n_params = sharding.broadcast_weights(base, src=0)
    parity = parity_slice(base, d, dev, a.parity_syndromes) if rank == 0 and a.parity_syndromes > 0 else None
    model = Amp(networks.pad_channels(base, 8).to(memory_format=torch.channels_last))
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    cfg = {"size": d, "p_error": 0.1}
    evaluation.evaluate(model, "toric-code-v0", cfg, d // 2, dev, [0.1], num_of_episodes=64 * world, num_of_steps=2,
                        keep_failed=False)       # warm-up
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    t0 = time.perf_counter()
    ec, gs, st, mq, _ = evaluation.evaluate(model, "toric-code-v0", cfg, d // 2, dev, [0.1], num_of_episodes=a.syndromes,
                                            num_of_steps=a.steps, seed=1, keep_failed=False)
    torch.cuda.synchronize()
    wall = sharding.max_over_ranks(time.perf_counter() - t0, dev)
    # the env path alone: the same number of env steps for this rank's share, random actions instead of the network
    n_local = sharding.shard_range(a.syndromes, rank, world)[1]
    eng = pkg.ToricEngine(d, max(n_local, 1), device=dev, seed=1, p_error=0.1, max_steps=a.steps)
    eng.reset()
    eng.perspectives()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    for _ in range(int(round(float(st[0])))):
        eng.actor_step(auto_reset=False)
    torch.cuda.synchronize()
    env_wall = sharding.max_over_ranks(time.perf_counter() - t1, dev)
    if rank == 0:
        print(json.dumps({"config": "c3: d=9 p=0.1 greedy NN_17 random-init (weights absent from the checkout: "
                                    ".MISSING_LARGE_BLOBS:8-9), bf16 autocast + padded channels for the bulk, "
                                    "%d-step cap" % a.steps, "n_gpus": world, "syndromes": a.syndromes,
                          "weights_broadcast_params": n_params, "success": float(ec[0]), "ground_state": float(gs[0]),
                          "mean_steps": float(st[0]), "syndromes_per_s": a.syndromes / wall, "seconds": wall,
                          "env_path_seconds_same_steps": env_wall, "network_share_of_time": 1.0 - env_wall / wall,
                          "parity_slice_fp32_vs_cpu": parity,
                          "note": "success rate NOT comparable with results/results_mats/RL_9.txt (random-init weights)"}),
              flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
