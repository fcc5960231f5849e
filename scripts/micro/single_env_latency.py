#!/usr/bin/env python
"""Latency of the single-env drop-in path (BASELINE configs[0]: d=5 greedy rollout with NN_11): where does a
step's wall time go?"""
import os, sys, time, json
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import toric_rl_decoder_b200 as pkg  # noqa: E402
from toric_rl_decoder_b200 import networks, evaluation  # noqa: E402

networks.strict_fp32()
d = 5
model = networks.load_state(networks.NN_11(d, 3), os.path.join(ROOT, "tests", "golden", "Size_5_NN_11_state.pt"), device="cuda")


def t(fn, reps=50):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


res = {}
res["ToricEngine(d,1) construct ms"] = t(lambda: pkg.ToricEngine(d, 1), 20)
eng = pkg.ToricEngine(d, 1, seed=1)
eng.reset()
res["reset ms"] = t(lambda: eng.reset())
res["perspectives() ms"] = t(lambda: eng.perspectives())
res["total_perspectives() (sync) ms"] = t(lambda: eng.total_perspectives())
eng.perspectives()
total = eng.total_perspectives()
x = eng.out[:total]
with torch.no_grad():
    res["NN_11 forward (%d persp) ms" % total] = t(lambda: model(x))
    q = model(x).float().contiguous()
q_sel = torch.empty((1, 3), device="cuda")
res["select_action_device ms"] = t(lambda: pkg.select_action_device(q, eng.pos, eng.offsets, 1, actions=eng.actions, q_sel=q_sel))
res["actor_step(actions) ms"] = t(lambda: eng.actor_step(actions=eng.actions, auto_reset=False, freeze_done=True))
env = pkg.ToricCodeB200({"size": d, "p_error": 0.1})
env.reset()
res["ToricCodeB200.step ms"] = t(lambda: env.step([0, 1, 1, 1]))
res["ToricCodeB200.reset ms"] = t(lambda: env.reset())
st = env.state
res["generatePerspective(state) ms"] = t(lambda: pkg.generatePerspective(d // 2, d, st))
res["selectAction(...) ms"] = t(lambda: pkg.selectAction(3, 0.0, d // 2, d, st, model))
cfg = {"size": d, "p_error": 0.1}
res["evaluate(1 episode) ms"] = t(lambda: evaluation.evaluate(model, "toric-code-v0", cfg, d // 2, "cuda", [0.1], num_of_episodes=1, num_of_steps=75, keep_failed=False), 10)
for k, v in res.items():
    print("%-42s %8.3f" % (k, v))
