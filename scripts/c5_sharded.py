#!/usr/bin/env python
"""BASELINE configs[4] as stated: 2^20 envs of a large lattice sharded over the ranks of one box (torchrun), random
actions, env-chunked perspectives.  Prints one JSON line per (d, p) with the aggregate over all ranks (time = max)."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from toric_rl_decoder_b200 import sharding  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--global-envs", type=int, default=1 << 20)
    ap.add_argument("--cases", default="15:0.05,15:0.1,15:0.2,25:0.05,25:0.1,25:0.2")
    ap.add_argument("--steps", type=int, default=3)
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    rank, world, local = sharding.init_from_env("nccl")
    dev = torch.device("cuda", local)
    for case in a.cases.split(","):
        d, p = int(case.split(":")[0]), float(case.split(":")[1])
        eng = sharding.make_sharded_engine(d, a.global_envs, device=dev, seed=9, p_error=p)
        n = eng.n
        per_persp = 8 * d * d + 4
        eng.reset()
        eng.perspectives_count()
        pbar = eng.total_perspectives() / n
        chunk = max(256, min(n, int((24 << 30) / (per_persp * pbar * 1.5)) // 256 * 256))
        eng.ensure_capacity(int(chunk * pbar * 1.5))

        def one_step(check=False):
            eng.actor_pre()
            for e0 in range(0, n, chunk):
                eng.perspectives(syn=eng.syn[e0:min(n, e0 + chunk)])
                if check:
                    assert int(eng.overflow.item()) == 0, "perspective buffer too small"
        one_step(check=True)
        c0 = eng.counters.clone()
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            one_step()
        e1.record()
        torch.cuda.synchronize()
        assert int(eng.overflow.item()) == 0
        dt = sharding.max_over_ranks(e0.elapsed_time(e1) * 1e-3, dev)
        ctr = eng.counters - c0
        sharding.allreduce_counters(ctr)
        if rank == 0:
            sum_p, steps = int(ctr[5].item()), int(ctr[0].item())
            print(json.dumps({"config": "c5 sharded", "d": d, "p_error": p, "global_envs": a.global_envs, "n_gpus": world,
                              "envs_per_gpu": n, "env_chunk": chunk, "mean_perspectives_per_env": sum_p / steps,
                              "env_steps_per_s": steps / dt, "perspective_GBps_aggregate": sum_p * per_persp / dt / 1e9,
                              "perspective_GBps_per_gpu": sum_p * per_persp / dt / 1e9 / world, "seconds": dt}), flush=True)
        del eng
        torch.cuda.empty_cache()
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
