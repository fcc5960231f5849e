"""Time the UNMODIFIED reference's own CPU path in the build container (needs /root/reference).

TEST / MEASUREMENT INFRASTRUCTURE ONLY.  The reference is Python + numba and cannot travel to the GPU box,
so `bench.py`'s CPU arm there times the C port (oracle/toric_oracle.c).  This script records what the
reference itself does on this container's cores for the same workload, so the port's number can be read
against the real thing ("CPU baseline"):

    per iteration, per process:  numba generatePerspectiveBatch (src/numba/util_actor.py:56-67)
                                 + np.concatenate + float32 cast (src/numba/util_actor.py:35-39)
                                 + random action from the perspective list
                                 + EnvSet.step (src/EnvSet.py:38-47)
                                 + resetTerminalEnvs on terminal / over-long envs (src/Actor_mp.py:171-183)
    no_envs = 100 per process (the value of the reference's recorded runs), one process per core
    (multiprocessing spawn, like Distributed_mp.py:199-211), numba warm-up excluded.

    python oracle/time_reference.py [--d 7] [--p 0.1] [--seconds 8] [--procs N] > profiles/r01_reference_python_cpu.json
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def worker(rank, d, p, no_envs, seconds, use_numba, q):
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    from oracle import ref_loader
    ref = ref_loader.load()
    np.random.seed(1234 + rank)
    rng = np.random.default_rng(rank)
    env = ref.gym.make("toric-code-v0", config={"size": d, "min_qbit_errors": 0, "p_error": p})
    envs = ref.EnvSet(env, no_envs)
    g = d // 2
    states = envs.resetAll()
    steps_per_env = np.zeros(no_envs, dtype=int)
    max_steps = 75                                               # Distributed_mp.py:44

    def iteration(states):
        if use_numba:
            per, pos, counts = ref.numba_util_actor.generatePerspectiveBatch(g, d, states)
            counts = np.array(list(counts))
            flat_pos = [x for e in pos for x in e]
            flat_per = [np.asarray(v) for e in per for v in e]
        else:
            flat_per, flat_pos, counts = [], [], []
            for s in states:
                v, ps = ref.util.generatePerspectiveOptimized(g, d, s)
                flat_per += v
                flat_pos += ps
                counts.append(len(v))
            counts = np.array(counts)
        splice = np.cumsum(counts)
        tensor = torch.from_numpy(np.concatenate(flat_per).reshape(-1, 2, d, d)).type("torch.Tensor")
        actions = np.zeros((no_envs, 4), dtype=int)
        for e in range(no_envs):
            lo = splice[e] - counts[e]
            m, i, j = flat_pos[lo + int(rng.integers(counts[e]))]
            actions[e] = (m, i, j, int(rng.integers(1, 4)))
        nxt, rew, term, _ = envs.step(actions)
        steps_per_env[:] += 1
        idx = np.argwhere(np.logical_or(term, steps_per_env > max_steps)).flatten()
        if len(idx):
            nxt[idx] = envs.resetTerminalEnvs(idx)
            steps_per_env[idx] = 0
        return nxt, int(tensor.shape[0])

    for _ in range(3):                                           # numba JIT + caches
        states, _ = iteration(states)
    t0 = time.perf_counter()
    n_it, n_p = 0, 0
    while time.perf_counter() - t0 < seconds:
        states, k = iteration(states)
        n_it += 1
        n_p += k
    dt = time.perf_counter() - t0
    q.put((rank, n_it * no_envs, n_p, dt))


def run(d, p, no_envs, seconds, procs, use_numba):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=worker, args=(r, d, p, no_envs, seconds, use_numba, q)) for r in range(procs)]
    for x in ps:
        x.start()
    res = [q.get() for _ in ps]
    for x in ps:
        x.join()
    steps = sum(r[1] for r in res)
    persp = sum(r[2] for r in res)
    wall = max(r[3] for r in res)
    return {"env_steps_per_sec": steps / wall, "per_core": steps / wall / procs, "procs": procs,
            "mean_perspectives_per_env": persp / max(steps, 1), "seconds": wall}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--d", type=int, default=7)
    ap.add_argument("--p", type=float, default=0.1)
    ap.add_argument("--no-envs", type=int, default=100)
    ap.add_argument("--seconds", type=float, default=8.0)
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    a = ap.parse_args()
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    os.environ.setdefault("MKL_NUM_THREADS", "1")
    out = {"what": "UNMODIFIED reference (Lindeby/toric-RL-decoder @ /root/reference) on the build container's CPU cores: "
                   "EnvSet.step + generatePerspectiveBatch + concatenate/float32 + random actions + reset of finished envs",
           "d": a.d, "p_error": a.p, "no_envs_per_process": a.no_envs, "cpu_count": os.cpu_count(), "unit": "env-steps/s",
           "numba_all_cores": run(a.d, a.p, a.no_envs, a.seconds, a.procs, True),
           "numba_one_core": run(a.d, a.p, a.no_envs, a.seconds, 1, True),
           "numpy_one_core": run(a.d, a.p, a.no_envs, a.seconds, 1, False)}
    print(json.dumps(out))
