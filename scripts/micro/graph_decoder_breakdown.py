#!/usr/bin/env python
"""Where a single-env greedy episode's wall time goes with the graphed decoder (BASELINE configs[0], d=5, NN_11)."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from toric_rl_decoder_b200 import networks, evaluation  # noqa: E402

networks.strict_fp32()
d = int(os.environ.get("D", 5))
w = {5: "Size_5_NN_11_state.pt", 7: "Size_7_NN_11_random_state.pt"}[d]
model = networks.load_state(networks.NN_11(d, 3), os.path.join(ROOT, "tests", "golden", w), device="cuda")
cfg = {"size": d, "p_error": 0.1}
EP = 100


def sync():
    torch.cuda.synchronize()


def ev(seed, g):
    return evaluation.evaluate(model, "toric-code-v0", cfg, d // 2, "cuda", [0.1], num_of_episodes=1, num_of_steps=75,
                               seed=seed, keep_failed=False, use_graph=g)


for g in (False, True):
    for s in range(3):
        ev(1000 + s, g)
    sync(); t0 = time.perf_counter(); steps = 0.0
    for s in range(EP):
        steps += float(ev(s, g)[2][0])
    sync(); dt = time.perf_counter() - t0
    print("evaluate(1 episode) use_graph=%-5s  %.3f ms/episode  %.2f steps/episode  %.0f env-steps/s" % (g, 1e3 * dt / EP, steps / EP, steps / dt))

dec = evaluation.graphed_decoder(model, d, 1, "cuda", 75)
seg = {"reseed": 0.0, "reset+clone": 0.0, "run": 0.0, "stats": 0.0}
nsteps = 0
for s in range(EP):
    sync(); t0 = time.perf_counter()
    eng = dec.reseed(0, s, 0.1)
    sync(); t1 = time.perf_counter()
    eng.reset(); init = eng.qubits.clone()
    sync(); t2 = time.perf_counter()
    dec.run(75)
    sync(); t3 = time.perf_counter()
    solved = (eng.syn.reshape(1, -1) != 0).sum(dim=1) == 0
    eng.gs.fill_(1); gs = eng.ground_state().bool()
    out = torch.stack([solved.sum().double(), gs.sum().double(), eng.steps.sum().double()]).cpu()
    sync(); t4 = time.perf_counter()
    nsteps += int(out[2])
    for k, v in zip(seg, (t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
        seg[k] += v
for k, v in seg.items():
    print("%-14s %.3f ms/episode" % (k, 1e3 * v / EP))
print("steps/episode %.2f" % (nsteps / EP))
# one graph replay + the 8-byte read, GPU idle in between
eng = dec.reseed(0, 0, 0.1); eng.reset(); eng.perspectives(); sync()
t0 = time.perf_counter()
for _ in range(200):
    dec.graph.replay(); dec.unsolved.item()
dt = time.perf_counter() - t0
print("graph.replay() + unsolved.item()   %.3f ms" % (1e3 * dt / 200))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200):
    dec.graph.replay()
e1.record(); sync()
print("graph.replay() back to back (GPU) %.3f ms, nodes: see DESIGN" % (e0.elapsed_time(e1) / 200))
