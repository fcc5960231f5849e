#!/usr/bin/env python
"""Plain depolarizing evaluation sweep, the reference's tests/evaluate_network.py protocol: greedy evaluate() at
p = linspace(0.06, 0.20, 8), 75-step cap, against the rates the reference logged with 10 000 episodes per p
(runs/test_size_{5,7}_steps_10000 tfevents, decoded with tensorboard's EventAccumulator in the build container; the
run does not record which checkpoint it used -- the converged NN_11 state_dicts are the ones in the checkout).

  python scripts/c_depolarizing_sweep.py [--episodes 32768]
"""
import argparse
import json
import math
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from toric_rl_decoder_b200 import evaluation, networks  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
REF_GS = {5: (0.9854, 0.9561, 0.9043, 0.8319, 0.7443, 0.6411, 0.533, 0.4279),
          7: (0.9955, 0.9795, 0.947, 0.8705, 0.7779, 0.6558, 0.518, 0.3957)}
REF_OK = {5: (1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 0.9999, 0.9998),
          7: (1.0, 0.9999, 1.0, 0.9996, 0.9997, 0.9996, 0.9993, 0.9993)}
N_REF = 10000
WEIGHTS = {5: "Size_5_NN_11_converged_state.pt", 7: "Size_7_NN_11_random_state.pt"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--episodes", type=int, default=1 << 15)
    a = ap.parse_args()
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    ps = [float(x) for x in np.linspace(0.06, 0.2, 8, endpoint=True)]
    for d in (5, 7):
        model = networks.load_state(networks.NN_11(d, 3), os.path.join(GOLD, WEIGHTS[d]), device="cuda")
        model = networks.pad_channels(model, 8).to(memory_format=torch.channels_last)
        cfg = {"size": d, "min_qubit_errors": 0, "p_error": 0.1}
        t0 = time.perf_counter()
        ec, gs, st, mq, _ = evaluation.evaluate(model, "toric-code-v0", cfg, d // 2, "cuda", ps,
                                                num_of_episodes=a.episodes, num_of_steps=75, seed=21, keep_failed=False)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        z = []
        for g, r in zip(gs, REF_GS[d]):
            pooled = (g * a.episodes + r * N_REF) / (a.episodes + N_REF)
            z.append((g - r) / math.sqrt(pooled * (1 - pooled) * (1.0 / a.episodes + 1.0 / N_REF)))
        print(json.dumps({"config": "depolarizing sweep p=6..20 %%, d=%d, NN_11 converged weights (TF32 conv)" % d,
                          "episodes_per_p": a.episodes, "p_error": ps, "ground_state_rate": [float(x) for x in gs],
                          "ground_state_rate_reference_10000_episodes": REF_GS[d], "z": z,
                          "success_rate": [float(x) for x in ec], "success_rate_reference": REF_OK[d],
                          "mean_steps": [float(x) for x in st], "mean_q": [float(x) for x in mq],
                          "episodes_per_s": a.episodes * len(ps) / wall, "seconds": wall}), flush=True)


if __name__ == "__main__":
    main()
