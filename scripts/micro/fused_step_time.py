#!/usr/bin/env python
"""Time the fused actor step (tc_actor_step: three back-to-back launches, no events in between)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import toric_rl_decoder_b200 as pkg  # noqa: E402

eng = pkg.ToricEngine(7, 65536, seed=0, p_error=0.1)
eng.ensure_capacity(eng.worst_case_capacity())
eng.reset()
for _ in range(80):
    eng.actor_step()
torch.cuda.synchronize()
best = []
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        eng.actor_step()
    e1.record()
    torch.cuda.synchronize()
    best.append(e0.elapsed_time(e1) / 200)
print("%s fused step: %s ms" % (os.path.basename(os.environ.get("TORIC_B200_LIB", "default")),
                                " ".join("%.4f" % b for b in best)))
