#!/usr/bin/env python
"""tc_reset alone at 2^22 envs (the kernel-table size), for an ncu capture of reset_kernel."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import toric_rl_decoder_b200 as pkg  # noqa: E402

d = int(sys.argv[1]) if len(sys.argv) > 1 else 7
p = float(sys.argv[2]) if len(sys.argv) > 2 else 0.1
eng = pkg.ToricEngine(d, 1 << 22, seed=0, p_error=p)
for _ in range(3):
    eng.reset()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
eng.reset()
e1.record()
torch.cuda.synchronize()
print("reset d=%d p=%g 2^22 envs: %.4f ms" % (d, p, e0.elapsed_time(e1)))
