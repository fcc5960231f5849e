#!/usr/bin/env python
"""Q-network forward throughput (perspectives/s): reference channel schedule vs channels padded to a multiple of
8/16/32/64, fp32+TF32 and bf16 autocast, channels_last.  The network is not the hot path, but it bounds the greedy
evaluation configs."""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from toric_rl_decoder_b200 import networks  # noqa: E402

torch.backends.cudnn.allow_tf32 = True
torch.backends.cuda.matmul.allow_tf32 = True
torch.backends.cudnn.benchmark = True


def bench(model, x, amp):
    def run():
        with torch.no_grad():
            if amp:
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    return model(x)
            return model(x)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 5


for arch, d, B in (("NN_11", 7, 65536), ("NN_17", 9, 16384)):
    torch.manual_seed(0)
    base = getattr(networks, arch)(d, 3).cuda().eval()
    x = (torch.rand(B, 2, d, d, device="cuda") < 0.2).float().contiguous(memory_format=torch.channels_last)
    xb = x.to(torch.bfloat16)
    for mult in (1, 8, 16, 32, 64):
        m = base if mult == 1 else networks.pad_channels(base, mult)
        m = m.to(memory_format=torch.channels_last)
        for amp, inp, name in ((False, x, "tf32"), (True, x, "bf16 autocast, f32 in"), (True, xb, "bf16 autocast, bf16 in")):
            ms = bench(m, inp, amp)
            print(json.dumps({"arch": arch, "d": d, "batch": B, "pad_multiple": mult, "mode": name,
                              "ms": round(ms, 3), "perspectives_per_s": round(B / (ms * 1e-3))}), flush=True)
