#!/usr/bin/env python
"""Small batches are launch-bound: fused actor step eager vs captured in a CUDA graph (16 steps per replay)."""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import toric_rl_decoder_b200 as pkg  # noqa: E402

for d, n in ((7, 256), (7, 1024), (7, 4096), (7, 16384), (5, 1024)):
    eng = pkg.ToricEngine(d, n, seed=0, p_error=0.1)
    eng.reset()
    for _ in range(80):
        eng.actor_step()
    def timed(fn, reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    eager = timed(eng.actor_step, 320)
    g = eng.capture_actor_steps(16)
    g.replay()
    graph = timed(g.replay, 20) / 16
    print(json.dumps({"d": d, "envs": n, "eager_us_per_step": round(eager * 1e3, 2), "graph_us_per_step": round(graph * 1e3, 2),
                      "eager_env_steps_per_s": round(n / (eager * 1e-3)), "graph_env_steps_per_s": round(n / (graph * 1e-3))}), flush=True)
