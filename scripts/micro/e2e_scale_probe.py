#!/usr/bin/env python
"""Why does the host-buffer (e2e) actor step stop scaling past 2 ranks?  Run under torchrun with N ranks:
times HostActorPipeline per rank with / without NUMA binding and with / without the state download."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import toric_rl_decoder_b200 as pkg                      # noqa: E402
from toric_rl_decoder_b200 import HostActorPipeline, sharding   # noqa: E402


def run(eng, states, steps=300, depth=3, n_buf=8):
    d, n = eng.d, eng.n
    rng = np.random.default_rng(5)
    acts = []
    for _ in range(n_buf):
        a = np.stack([rng.integers(0, 2, n), rng.integers(0, d, n), rng.integers(0, d, n), rng.integers(1, 4, n)],
                     1).astype(np.int32)
        acts.append(torch.from_numpy(a).pin_memory())
    pipe = HostActorPipeline(eng, depth=depth, states=states)

    def loop(k_steps):
        s = 0.0
        for k in range(k_steps):
            if pipe.submitted - pipe.collected >= depth:
                s += float(pipe.collect()["reward"][0])
            pipe.submit(acts[k % n_buf])
        while pipe.collected < pipe.submitted:
            s += float(pipe.collect()["reward"][0])
        return s
    loop(30)
    torch.distributed.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    loop(steps)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return 1e3 * dt / steps, pipe.d2h_bytes


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    rank, world, local = sharding.init_from_env("nccl")
    if world == 1:
        torch.distributed.init_process_group("gloo", init_method="tcp://127.0.0.1:29533", rank=0, world_size=1)
    if rank == 0:
        os.system("nvidia-smi topo -m; lscpu | grep -i -E 'numa|socket|^CPU\\(s\\)|model name'")
    eng = pkg.ToricEngine(7, 65536, device=torch.device("cuda", local), seed=0, env_base=rank * 65536, p_error=0.1)
    eng.ensure_capacity(eng.worst_case_capacity())
    eng.reset()
    res = {}
    res["states u8"] = run(eng, "u8")
    res["states rows"] = run(eng, "rows")
    res["states none"] = run(eng, None)
    ncpu = sharding.bind_to_gpu_numa(local)
    res["bound(%d cpus) states u8" % ncpu] = run(eng, "u8")
    for k, (ms, b) in res.items():
        t = torch.tensor([ms], device="cuda")
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        if rank == 0:
            print("world %d  %-28s max-over-ranks %.4f ms/step  (d2h %.2f MB/step/rank -> %.1f GB/s aggregate)"
                  % (world, k, t.item(), b / 1e6, world * b / (t.item() * 1e-3) / 1e9), flush=True)
    print("rank %d affinity now %s" % (rank, sorted(os.sched_getaffinity(0))[:4]), flush=True)
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
