"""Where does the host-buffer step spend its time?  (run on the GPU box)"""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import toric_rl_decoder_b200 as pkg

d, n, steps = 7, 65536, 100
eng = pkg.ToricEngine(d, n, seed=0, p_error=0.1)
eng.ensure_capacity(eng.worst_case_capacity())
eng.reset()
rng = np.random.default_rng(1)
acts = [torch.from_numpy(np.stack([rng.integers(0, 2, n), rng.integers(0, d, n), rng.integers(0, d, n),
                                   rng.integers(1, 4, n)], axis=1).astype(np.int32)).pin_memory() for _ in range(8)]
dacts = [a.cuda() for a in acts]
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def timed(name, fn, reps=steps):
    for k in range(10):
        fn(k)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ev0.record()
    for k in range(reps):
        fn(k)
    ev1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / reps * 1e3
    print("%-52s gpu %.4f ms/step   wall %.4f ms/step" % (name, ev0.elapsed_time(ev1) / reps, wall), flush=True)


timed("1 compute only (device actions)", lambda k: eng.actor_step(auto_reset=True, actions_buffer=dacts[k % 8]))
pipe = pkg.HostActorPipeline(eng, depth=2)
slot = pipe.slots[0]


def snap(k):
    eng.actor_step(auto_reset=True, actions_buffer=dacts[k % 8])
    slot["d_reward"].copy_(eng.reward, non_blocking=True)
    slot["d_counts"].copy_(eng.counts, non_blocking=True)
    slot["d_term"].copy_(eng.terminal, non_blocking=True)
    pkg._lib.check(eng.lib.tc_unpack_state(pkg.engine._ptr(eng.syn), pkg.engine._ptr(slot["d_state"]), d, n,
                                           torch.cuda.current_stream().cuda_stream), "x")


timed("2 compute + snapshot kernels", snap)
timed("3 D2H 7 MB only", lambda k: slot["h_stage"].copy_(slot["d_stage"], non_blocking=True))
timed("4 H2D 1 MB only", lambda k: slot["d_actions"].copy_(acts[k % 8], non_blocking=True))


def host_only(k):
    pipe.submit(acts[k % 8])
    pipe.collected = pipe.submitted      # never wait


timed("5 pipeline submit, never collect (queue depth unbounded)", host_only)
torch.cuda.synchronize()
for depth in (2, 3, 4):
    p = pkg.HostActorPipeline(eng, depth=depth)

    def run(k, p=p, depth=depth):
        if p.submitted - p.collected >= depth:
            p.collect()
        p.submit(acts[k % 8])

    timed("6 pipeline depth %d (collect oldest when full)" % depth, run)
    while p.collected < p.submitted:
        p.collect()
