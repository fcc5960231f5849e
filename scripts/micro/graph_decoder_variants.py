#!/usr/bin/env python
"""GPU time of one captured decoding step (d=5 and d=7, one env, NN_11, strict fp32) under cuDNN settings."""
import copy, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from toric_rl_decoder_b200 import networks, evaluation  # noqa: E402

networks.strict_fp32()


def gpu_ms(graph, reps=200):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    graph.replay(); torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        graph.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for d, w in ((5, "Size_5_NN_11_state.pt"), (7, "Size_7_NN_11_random_state.pt")):
    base = networks.load_state(networks.NN_11(d, 3), os.path.join(ROOT, "tests", "golden", w), device="cuda")
    x = torch.randint(0, 2, ((2 * d * d + 63) // 64 * 64, 2, d, d), device="cuda").float()
    with torch.no_grad():
        ref = base(x)
    for name, bench, cl, tf32 in (("default", False, False, False), ("cudnn.benchmark", True, False, False),
                                  ("channels_last", False, True, False), ("benchmark+channels_last", True, True, False),
                                  ("benchmark+tf32", True, False, True)):
        torch.backends.cudnn.benchmark = bench
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        m = copy.deepcopy(base)
        if cl:
            m = m.to(memory_format=torch.channels_last)
        dec = evaluation.GraphedGreedyDecoder(m, d, 1, torch.device("cuda", 0), 75)
        with torch.no_grad():
            err = ((m(x) - ref).abs().max() / ref.abs().max()).item()
        print("d=%d %-26s graph replay %.3f ms   max rel dQ vs default %.2e" % (d, name, gpu_ms(dec.graph), err))
    torch.backends.cudnn.benchmark = False
    networks.strict_fp32()
