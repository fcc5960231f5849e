#!/usr/bin/env python
"""BASELINE configs[2] in shape: d=9 batched greedy evaluation with NN_17 sharded over the ranks of one box (torchrun).
Size_9_NN_17.pt is not in the reference checkout, so the weights are seeded random-init (broadcast from rank 0) and the
episode count is a bounded sample of the 2^24 the config names: what is shown is the sharded evaluation path
(episodes by global index, counters all-reduced) and its throughput, not a decoder success rate."""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from toric_rl_decoder_b200 import evaluation, networks, sharding  # noqa: E402


class Amp(torch.nn.Module):
    pad_small_batches = True

    def __init__(self, m):
        super().__init__()
        self.m = m

    def forward(self, x):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            return self.m(x.contiguous(memory_format=torch.channels_last)).float()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--syndromes", type=int, default=1 << 14)
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    rank, world, local = sharding.init_from_env("nccl")
    dev = torch.device("cuda", local)
    d = 9
    torch.manual_seed(rank)                      # different init per rank: the broadcast must fix that
    base = networks.NN_17(d, 3).to(dev).eval()
    n_params = sharding.broadcast_weights(base, src=0)
    model = Amp(networks.pad_channels(base, 8).to(memory_format=torch.channels_last))
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
    if rank == 0:
        print(json.dumps({"config": "c3 sharded: d=9 p=0.1 greedy NN_17 random-init (bf16 autocast, padded channels), "
                                    "%d-step cap" % a.steps, "n_gpus": world, "syndromes": a.syndromes,
                          "weights_broadcast_params": n_params, "success": float(ec[0]), "ground_state": float(gs[0]),
                          "mean_steps": float(st[0]), "syndromes_per_s": a.syndromes / wall, "seconds": wall}), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
