#!/usr/bin/env python
"""Diagnostic for the single-pass step: per-env duration of the pre-phase (loads, step, store, publish) and of the
look-back wait, from a TC_FUSED_TIMING build (globaltimer stamps written into the workspace's unused local_off
region).  usage: TORIC_B200_LIB=<timing build> python scripts/micro/fused_timing.py"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import toric_rl_decoder_b200 as pkg  # noqa: E402

d, n = 7, 65536
eng = pkg.ToricEngine(d, n, seed=0, p_error=0.1, single_pass=True)
eng.ensure_capacity(eng.worst_case_capacity())
eng.reset()
for _ in range(30):
    eng.actor_step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
eng.actor_step()
e1.record()
torch.cuda.synchronize()
w = eng.workspace[:4 * n].view(torch.int32).cpu().numpy().astype(np.int64)
pre = (w & 0xffff) * 16 / 1000.0          # us
wait = ((w >> 16) & 0xffff) * 16 / 1000.0
pc = lambda x, q: float(np.percentile(x, q))  # noqa: E731
out = {"lib": os.environ.get("TORIC_B200_LIB"), "step_ms": e0.elapsed_time(e1),
       "pre_us": {"mean": float(pre.mean()), "p50": pc(pre, 50), "p99": pc(pre, 99)},
       "wait_us": {"mean": float(wait.mean()), "p50": pc(wait, 50), "p90": pc(wait, 90), "p99": pc(wait, 99),
                   "max": float(wait.max())},
       "wait_by_idx_in_group": [round(float(wait[i::32].mean()), 2) for i in range(0, 32, 4)] + [round(float(wait[31::32].mean()), 2)],
       "wait_by_position_in_grid": [round(float(x.mean()), 2) for x in np.array_split(wait, 16)],
       "pre_by_position_in_grid": [round(float(x.mean()), 2) for x in np.array_split(pre, 16)]}
print(json.dumps(out))
