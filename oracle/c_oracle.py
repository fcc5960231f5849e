"""ctypes wrapper of oracle/libtoric_oracle.so (the C restatement).  TEST INFRASTRUCTURE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libtoric_oracle.so")
_c = ctypes
_lib = None


def build():
    subprocess.run(["make", "-C", HERE, "-s"], check=True)
    return LIB


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(os.path.join(HERE, "toric_oracle.c")):
            build()
        lib = _c.CDLL(LIB)
        vp, i, i64, u32, u64 = _c.c_void_p, _c.c_int, _c.c_int64, _c.c_uint32, _c.c_uint64
        lib.to_syndrome.argtypes = [vp, vp, i]
        lib.to_syndrome.restype = None
        lib.to_step.argtypes = [vp, vp, vp, i, vp]
        lib.to_step.restype = i
        lib.to_ground_state.argtypes = [vp, i, i]
        lib.to_ground_state.restype = i
        lib.to_perspectives.argtypes = [vp, i, vp, vp]
        lib.to_perspectives.restype = i64
        lib.to_philox4x32_10.argtypes = [vp, vp, vp]
        lib.to_philox4x32_10.restype = None
        lib.to_reset.argtypes = [vp, vp, i, u64, u32, u32, i, i, u64]
        lib.to_reset.restype = u32
        lib.to_reset_batch.argtypes = [vp, vp, i, i64, u64, u32, u32, i, i, u64, vp]
        lib.to_reset_batch.restype = None
        lib.to_random_action.argtypes = [vp, i, u64, u32, u32, u64, vp, vp]
        lib.to_random_action.restype = None
        lib.to_actor_steps.argtypes = [vp, vp, vp, vp, vp, i64, i, i, i, u64, i64, vp, vp, i64, vp, vp]
        lib.to_actor_steps.restype = i
        _lib = lib
    return _lib


def _p(a):
    return a.ctypes.data_as(_c.c_void_p) if a is not None else None


def syndrome(q):
    q = np.ascontiguousarray(q, dtype=np.uint8)
    d = q.shape[-1]
    flat = q.reshape(-1, 2, d, d)
    s = np.zeros_like(flat)
    lib = load()
    for e in range(len(flat)):
        lib.to_syndrome(_p(flat[e]), _p(s[e]), d)
    return s.reshape(q.shape)


def step(q, s, action):
    """in-place on uint8 arrays q, s (2,d,d); returns (reward, terminal)"""
    d = q.shape[-1]
    a = np.ascontiguousarray(action, dtype=np.int32)
    r = _c.c_double(0)
    term = load().to_step(_p(q), _p(s), _p(a), d, _c.byref(r))
    return r.value, bool(term)


def ground_state(q, prev=True):
    q = np.ascontiguousarray(q, dtype=np.uint8)
    return bool(load().to_ground_state(_p(q), q.shape[-1], int(prev)))


def perspectives(s):
    s = np.ascontiguousarray(s, dtype=np.uint8)
    d = s.shape[-1]
    lib = load()
    P = lib.to_perspectives(_p(s), d, None, None)
    out = np.zeros((P, 2, d, d), np.float32)
    pos = np.zeros((P, 3), np.int32)
    lib.to_perspectives(_p(s), d, _p(out), _p(pos))
    return out, pos.astype(np.int64)


def philox(ctr, key):
    c = np.ascontiguousarray(ctr, dtype=np.uint32)
    k = np.ascontiguousarray(key, dtype=np.uint32)
    o = np.zeros(4, np.uint32)
    load().to_philox4x32_10(_p(c), _p(k), _p(o))
    return o


def reset(d, env_id, episode, thresh, seed, mode=0, n_fixed=0):
    q = np.zeros((2, d, d), np.uint8)
    s = np.zeros((2, d, d), np.uint8)
    att = load().to_reset(_p(q), _p(s), d, int(env_id), int(episode), int(thresh), mode, n_fixed, int(seed))
    return q, s, int(att)


def reset_batch(d, n, env_base, episode, thresh, seed, mode=0, n_fixed=0):
    """to_reset for envs env_base .. env_base+n-1: (qubits u8 (n,2,d,d), syndromes u8 (n,2,d,d), attempts u32 (n,))."""
    q = np.zeros((n, 2, d, d), np.uint8)
    s = np.zeros((n, 2, d, d), np.uint8)
    att = np.zeros(n, np.uint32)
    load().to_reset_batch(_p(q), _p(s), d, n, int(env_base), int(episode), int(thresh), mode, n_fixed, int(seed), _p(att))
    return q, s, att


def random_action(s, env_id, episode, step_no, seed):
    s = np.ascontiguousarray(s, dtype=np.uint8)
    d = s.shape[-1]
    act = np.zeros(4, np.int32)
    scratch = np.zeros(3 * 2 * d * d, np.int32)
    load().to_random_action(_p(s), d, int(env_id), int(episode), int(step_no), int(seed), _p(act), _p(scratch))
    return act


class ActorSlice:
    """One single-threaded actor over its own envs (mirrors one reference actor process)."""

    def __init__(self, d, n, p_error, seed=0, env_base=0, max_steps=75):
        self.lib = load()
        self.d, self.n, self.seed, self.env_base, self.max_steps = d, n, seed, env_base, max_steps
        self.q = np.zeros((n, 2, d, d), np.uint8)
        self.s = np.zeros((n, 2, d, d), np.uint8)
        self.steps = np.zeros(n, np.int32)
        self.episode = np.zeros(n, np.uint32)
        th = min(int(np.floor(p_error * 4294967296.0)), 4294967295)
        self.thresh = np.full(n, th, np.uint32)
        self.cap = n * 2 * d * d
        self.out = np.zeros((self.cap, 2, d, d), np.float32)
        self.pos = np.zeros((self.cap, 3), np.int32)
        self.offsets = np.zeros(n + 1, np.int64)
        self.counters = np.zeros(4, np.int64)
        self.lib.to_reset_batch(_p(self.q), _p(self.s), d, n, env_base, 0, th, 0, 0, seed, None)
        self.episode[:] = 1

    def run(self, steps):
        rc = self.lib.to_actor_steps(_p(self.q), _p(self.s), _p(self.steps), _p(self.episode), _p(self.thresh),
                                     self.n, self.d, steps, self.max_steps, self.seed, self.env_base, _p(self.out),
                                     _p(self.pos), self.cap, _p(self.offsets), _p(self.counters))
        if rc != 0:
            raise RuntimeError("to_actor_steps failed")


class ActorLoop:
    """The timed CPU arm: `nthreads` ActorSlices, one per host core, each on its own envs -- the
    reference's one-process-per-actor layout (Distributed_mp.py:199-211) with threads (the C call releases
    the GIL)."""

    def __init__(self, d, n, p_error, seed=0, env_base=0, max_steps=75, nthreads=0):
        self.nthreads = nthreads if nthreads > 0 else (os.cpu_count() or 1)
        self.n = n
        per = (n + self.nthreads - 1) // self.nthreads
        self.slices = []
        base = 0
        while base < n:
            k = min(per, n - base)
            self.slices.append(ActorSlice(d, k, p_error, seed, env_base + base, max_steps))
            base += k
        self._pool = None

    def run(self, steps):
        """Advance every slice by `steps` steps; the worker threads are created once and reused (a fresh
        thread per call showed up in sub-second timed regions)."""
        if self._pool is None:
            from concurrent.futures import ThreadPoolExecutor
            self._pool = ThreadPoolExecutor(max_workers=len(self.slices))
        for f in [self._pool.submit(s.run, steps) for s in self.slices]:
            f.result()

    @property
    def counters(self):
        return sum(s.counters for s in self.slices)
