"""Device-resident batch of toric-code environments driven through the C ABI.

`ToricEngine` owns the torch tensors (the library never allocates) and enqueues every call on
torch's current CUDA stream.  The reference-shaped classes in `envs.py` and the perspective
functions in `perspectives.py` are thin layers over it.
"""
import ctypes

import numpy as np
import torch

from . import _lib

SAMPLE_DEPOLARIZING = 0   # generateRandomError        (ToricCode.py:130-151)
SAMPLE_FIXED_WEIGHT = 1   # generateNRandomErrors      (ToricCode.py:153-166)
SAMPLE_FIXED_PLUS_P = 2   # generateNPlusQRandomErrors (results/small_p_error_test.py:43-52)
SAMPLE_ONE_DRAW = 16      # OR-ed into the mode: one generator draw, no rejection of trivial syndromes

ACTOR_RANDOM_ACTIONS = 1
ACTOR_AUTO_RESET = 2
ACTOR_FREEZE_DONE = 4
ACTOR_SINGLE_PASS = 8     # one kernel per step (step + look-back scan + writer) instead of three; d <= 16, float32
SINGLE_PASS_MAX_D = 16

STATUS_BAD_ACTION = 1
STATUS_RESET_GAVE_UP = 2

# element types of the perspective tensor (include/toric_b200.h TC_DTYPE_*)
OUT_DTYPES = {torch.float32: 0, torch.uint8: 1, torch.bfloat16: 2, torch.float16: 3}

COUNTER_NAMES = ("env_steps", "episodes", "solved", "ground_kept", "timeouts", "perspectives",
                 "err_x", "err_y", "err_z", "qubits_sampled", "reset_attempts", "bad_actions")


def p_to_thresh(p):
    """Error threshold used by the device sampler: a qubit errs iff its 32-bit draw < floor(p * 2^32)."""
    p = np.asarray(p, dtype=np.float64)
    return np.minimum(np.floor(p * 4294967296.0), 4294967295.0).astype(np.uint32)


class _Range:
    """NVTX range per phase (reset / step / perspectives / actor_step ...), visible in nsys / ncu timelines; the
    push/pop pair costs well under a microsecond when no profiler is attached.  ToricEngine.nvtx = False disables it."""
    __slots__ = ("on",)

    def __init__(self, name, on=True):
        self.on = on
        if on:
            torch.cuda.nvtx.range_push(name)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        if self.on:
            torch.cuda.nvtx.range_pop()
        return False


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _stream(device=None):
    """torch's current stream as a cudaStream_t.  The library launches on the CURRENT device, so an engine that
    lives on another device fails loudly here instead of launching on the wrong GPU (one process per GPU is the
    intended layout; otherwise wrap the call in `with torch.cuda.device(engine.device):`)."""
    if device is not None and device.index is not None and torch.cuda.current_device() != device.index:
        raise RuntimeError("engine lives on %s but the current CUDA device is cuda:%d; wrap the call in "
                           "torch.cuda.device(...)" % (device, torch.cuda.current_device()))
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class ToricEngine:
    """n independent d x d toric codes resident on one GPU.

    Device layout (include/toric_b200.h): qubits int32[n,4,d] (bit planes x0,x1,z0,z1, one row per word),
    syn int32[n,2,d] (vertex, plaquette planes).
    """

    nvtx = True        # NVTX ranges "toric/<phase>" around every enqueue (class-wide switch)
    # what state_dict() saves: with the counter-based sampler (Philox keyed by seed, global env id, episode, step)
    # these tensors + the scalars below are the WHOLE state of the env batch, so a resumed run is bit-identical
    _STATE_TENSORS = ("qubits", "syn", "gs", "step_gs", "status", "steps", "episode", "p_thresh", "counters",
                      "actions", "reward", "terminal")

    def __init__(self, d, n, device="cuda", seed=0, env_base=0, p_error=0.1, max_steps=75,
                 out_dtype=torch.float32, single_pass=False):
        """out_dtype: element type of the perspective tensor -- float32 (the reference's,
        src/numba/util_actor.py:39), or bfloat16 / float16 / uint8 (same 0/1 values, 2-4x less HBM traffic).
        single_pass: actor_step() issues the one-kernel form of the step (TC_ACTOR_SINGLE_PASS: step, cross-env
        scan by decoupled look-back and perspective writer fused) where it applies (d <= 16, float32, no
        `between` hook); results are identical bit for bit."""
        self.lib = _lib.load()
        if out_dtype not in OUT_DTYPES:
            raise ValueError("out_dtype must be one of %s" % (sorted(str(k) for k in OUT_DTYPES),))
        self.out_dtype = out_dtype
        self.single_pass = bool(single_pass)
        if not (3 <= int(d) <= 32):
            raise ValueError("lattice size d must be in [3, 32]")
        if not torch.cuda.is_available():
            raise RuntimeError("ToricEngine needs a CUDA device; there is no CPU fallback")
        self.d, self.n = int(d), int(n)
        self.device = torch.device(device)
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.seed, self.env_base, self.max_steps = int(seed), int(env_base), int(max_steps)
        dev, n_, d_ = self.device, self.n, self.d
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)  # noqa: E731
        self.qubits = z((n_, 4, d_), torch.int32)
        self.syn = z((n_, 2, d_), torch.int32)
        self.prev_syn = None
        self.gs = torch.ones(n_, dtype=torch.uint8, device=dev)        # sticky flag of evalGroundState (tc_ground_state)
        self.step_gs = torch.ones(n_, dtype=torch.uint8, device=dev)   # per-step parity result of the fused actor step
        self.status = z(n_, torch.uint8)
        self.steps = z(n_, torch.int32)
        self.episode = z(n_, torch.int32)
        self.p_thresh = z(n_, torch.int32)
        self.counters = z(16, torch.int64)
        self.actions = z((n_, 4), torch.int32)
        self.reward = z(n_, torch.float32)
        self.terminal = z(n_, torch.uint8)
        self.counts = z(n_, torch.int32)
        self.offsets = z(n_ + 1, torch.int64)
        self.overflow = z(1, torch.int32)
        nbytes = ctypes.c_size_t(0)
        _lib.check(self.lib.tc_workspace_bytes(d_, n_, ctypes.byref(nbytes)), "tc_workspace_bytes")
        self.workspace = z(max(int(nbytes.value), 256), torch.uint8)
        self.out = None       # out_dtype [cap, 2, d, d]
        self.pos = None       # uint8   [cap, 4]
        self.cap = 0
        self.set_p_error(p_error)

    # -- configuration -------------------------------------------------------------------------
    def set_p_error(self, p, idx=None):
        """Scalar or per-env error probability (ToricCode.reset's p_error override, ToricCode.py:97-98)."""
        if idx is None:
            th = np.broadcast_to(p_to_thresh(p), (self.n,)).copy()
            self.p_thresh.copy_(torch.from_numpy(th.view(np.int32)))
        else:
            idx = torch.as_tensor(idx, dtype=torch.int64, device=self.device)
            th = np.broadcast_to(p_to_thresh(p), (idx.numel(),)).copy()
            self.p_thresh[idx] = torch.from_numpy(th.view(np.int32)).to(self.device)

    def worst_case_capacity(self):
        return self.n * 2 * self.d * self.d

    def ensure_capacity(self, cap):
        cap = max(int(cap), 1)
        if cap > self.cap:
            self.out = None
            self.pos = None
            self.out = torch.empty((cap, 2, self.d, self.d), dtype=self.out_dtype, device=self.device)
            self.pos = torch.empty((cap, 4), dtype=torch.uint8, device=self.device)
            self.cap = cap

    # -- mask injection / read-back --------------------------------------------------------------
    def _as_u8(self, a, shape):
        t = torch.as_tensor(np.ascontiguousarray(a) if isinstance(a, np.ndarray) else a)
        t = t.to(device=self.device).to(torch.uint8).contiguous()
        if tuple(t.shape) != tuple(shape):
            raise ValueError("expected shape %s, got %s" % (tuple(shape), tuple(t.shape)))
        return t

    def set_qubits(self, codes):
        """Inject Pauli-code masks (n,2,d,d); the syndrome is recomputed and the sticky flag cleared, as
        the reference's scripts do by assigning env.qubit_matrix / env.state."""
        t = self._as_u8(codes, (self.n, 2, self.d, self.d))
        _lib.check(self.lib.tc_pack_qubits(_ptr(t), _ptr(self.qubits), self.d, self.n, _stream(self.device)), "tc_pack_qubits")
        self.syndrome()
        self.gs.fill_(1)

    def get_qubits(self):
        out = torch.empty((self.n, 2, self.d, self.d), dtype=torch.uint8, device=self.device)
        _lib.check(self.lib.tc_unpack_qubits(_ptr(self.qubits), _ptr(out), self.d, self.n, _stream(self.device)),
                   "tc_unpack_qubits")
        return out

    def set_state(self, state01):
        t = self._as_u8(state01, (self.n, 2, self.d, self.d))
        _lib.check(self.lib.tc_pack_state(_ptr(t), _ptr(self.syn), self.d, self.n, _stream(self.device)), "tc_pack_state")

    def get_state(self, syn=None):
        syn = self.syn if syn is None else syn
        out = torch.empty((syn.shape[0], 2, self.d, self.d), dtype=torch.uint8, device=self.device)
        _lib.check(self.lib.tc_unpack_state(_ptr(syn), _ptr(out), self.d, syn.shape[0], _stream(self.device)), "tc_unpack_state")
        return out

    # -- kernels ---------------------------------------------------------------------------------
    def syndrome(self):
        with _Range("toric/syndrome", self.nvtx):
            self._syndrome()

    def _syndrome(self):
        _lib.check(self.lib.tc_syndrome(_ptr(self.qubits), _ptr(self.syn), self.d, self.n, _stream(self.device)), "tc_syndrome")

    def reset(self, idx=None, mode=SAMPLE_DEPOLARIZING, n_fixed=0):
        """Resample the listed envs (all when idx is None) until each has a non-trivial syndrome."""
        if idx is None:
            idx_t, n_idx = None, self.n
        else:
            idx_t = torch.as_tensor(idx, dtype=torch.int64, device=self.device).contiguous()
            n_idx = idx_t.numel()
            self.steps[idx_t] = 0
        with _Range("toric/reset", self.nvtx):
            _lib.check(self.lib.tc_reset(_ptr(self.qubits), _ptr(self.syn), _ptr(self.gs), _ptr(self.status),
                                         _ptr(self.p_thresh), _ptr(self.episode), _ptr(idx_t), n_idx, int(mode),
                                         int(n_fixed), self.seed, self.env_base, self.d, _stream(self.device)),
                       "tc_reset")
        if idx is None:
            self.steps.zero_()

    def step(self, actions):
        """EnvSet.step: actions (n,4) int (matrix,row,col,op).  Returns (reward f32[n], terminal u8[n])."""
        a = torch.as_tensor(actions)
        if tuple(a.shape) != (self.n, 4):
            raise ValueError("actions must have shape (%d, 4)" % self.n)
        with _Range("toric/step", self.nvtx):
            self.actions.copy_(a.to(torch.int32), non_blocking=True)
            _lib.check(self.lib.tc_step(_ptr(self.qubits), _ptr(self.syn), _ptr(self.actions), _ptr(self.reward),
                                        _ptr(self.terminal), _ptr(self.status), self.d, self.n, _stream(self.device)),
                       "tc_step")
            self.steps += 1
        return self.reward, self.terminal

    def ground_state(self):
        _lib.check(self.lib.tc_ground_state(_ptr(self.qubits), _ptr(self.gs), self.d, self.n, _stream(self.device)),
                   "tc_ground_state")
        return self.gs

    def perspectives(self, syn=None, n=None):
        """Launch count + scan + write into the engine's buffers (asynchronous).  `syn` may be any packed
        state tensor int32[n',2,d] with n' <= n."""
        syn = self.syn if syn is None else syn
        n = syn.shape[0] if n is None else n
        self._last_syn = syn
        if self.out is None:
            self.ensure_capacity(self.worst_case_capacity())
        self.overflow.zero_()
        with _Range("toric/perspectives", self.nvtx):
            _lib.check(self.lib.tc_perspectives_as(_ptr(syn), _ptr(self.out), OUT_DTYPES[self.out_dtype],
                                                   _ptr(self.pos), _ptr(self.counts), _ptr(self.offsets),
                                                   _ptr(self.overflow), self.cap, _ptr(self.workspace),
                                                   self.workspace.numel(), self.d, n, _stream(self.device)),
                       "tc_perspectives_as")

    def perspectives_count(self, syn=None, n=None):
        """Count + scan only (tc_persp_count): fills counts and offsets[n] = total, writes no perspectives."""
        syn = self.syn if syn is None else syn
        n = syn.shape[0] if n is None else n
        _lib.check(self.lib.tc_persp_count(_ptr(syn), _ptr(self.counts), _ptr(self.offsets), _ptr(self.workspace),
                                           self.workspace.numel(), self.d, n, _stream(self.device)), "tc_persp_count")

    def actor_step(self, *args, **kwargs):
        with _Range("toric/actor_step", self.nvtx):
            res = self._actor_step(*args, **kwargs)
            if self._curriculum is not None and kwargs.get("auto_reset", True):
                self._curriculum_update(False)
            return res

    # -- p_error curriculum of the reference's actor (src/Actor_mp.py:43-47,171-181) ---------------------------------
    _curriculum = None

    def set_p_error_curriculum(self, start, final, delta, strategy="linear"):
        """Every auto-reset of `actor_step` raises the env's p_error ceiling by `delta` up to `final` and draws the new
        episode's p_error as the ceiling ('linear') or uniformly on [start, ceiling] ('random'), as the reference's
        actor does on the host (env_p_error_start / _final / _delta / _strategy, Distributed_mp.py:53-56).  Call after
        the initial reset; start=None switches it off.  The ceilings live in `self.p_cap` (thresholds)."""
        if start is None:
            self._curriculum = None
            return
        if strategy not in ("linear", "random"):
            raise ValueError("strategy must be 'linear' or 'random'")
        t = [int(p_to_thresh(float(x))) for x in (start, delta, final)]
        self._curriculum = (t[0], t[1], t[2], 1 if strategy == "random" else 0)
        self.p_cap = torch.from_numpy(np.full(self.n, t[0], dtype=np.uint32).view(np.int32)).to(self.device)
        self._curriculum_update(True)

    def _curriculum_update(self, all_envs):
        s, dl, f, rnd = self._curriculum
        _lib.check(self.lib.tc_p_error_curriculum(_ptr(self.steps), _ptr(self.episode), _ptr(self.p_cap),
                                                  _ptr(self.p_thresh), s, dl, f, rnd, 1 if all_envs else 0, self.seed,
                                                  self.env_base, self.n, _stream(self.device)),
                   "tc_p_error_curriculum")

    def _actor_step(self, actions=None, auto_reset=True, mode=SAMPLE_DEPOLARIZING, n_fixed=0, keep_prev=False,
                    freeze_done=False, actions_buffer=None, results=None, between=None):
        """One fused actor iteration (asynchronous).  actions=None draws them on device; passing
        self.actions itself (e.g. filled by select_action) skips the copy; actions_buffer (device int32
        (n,4), 16-byte aligned) is read in place instead of self.actions.  freeze_done leaves envs whose
        syndrome is already trivial untouched (batched evaluation).  results = (reward f32 (n,), terminal u8
        (n,), counts i32 (n,)) device tensors to receive this step's results instead of self.reward /
        self.terminal / self.counts (HostActorPipeline points them into its download block).  between: a
        callable run after the step / count kernels are enqueued and before the perspective writer is (the
        step's results and new states are final at that point; HostActorPipeline records an event there)."""
        flags = (ACTOR_AUTO_RESET if auto_reset else 0) | (ACTOR_FREEZE_DONE if freeze_done else 0)
        reward, terminal, counts = results if results is not None else (self.reward, self.terminal, self.counts)
        act_buf = self.actions
        if actions_buffer is not None:
            act_buf = actions_buffer
        elif actions is None:
            flags |= ACTOR_RANDOM_ACTIONS
        elif actions is not self.actions:
            self.actions.copy_(torch.as_tensor(actions).to(torch.int32), non_blocking=True)
        if keep_prev and self.prev_syn is None:
            self.prev_syn = torch.zeros_like(self.syn)
        if self.out is None:
            self.ensure_capacity(self.worst_case_capacity())
        self._last_syn = self.syn
        # `overflow` is sticky here (cleared by perspectives() / perspective_views()): no memset between the
        # writer of one step and the first kernel of the next, so the launches chain back to back
        if self.out_dtype is not torch.float32 or between is not None:   # the fused entry point's two halves
            _lib.check(self.lib.tc_actor_pre(
                _ptr(self.qubits), _ptr(self.syn), _ptr(self.prev_syn if keep_prev else None), _ptr(act_buf),
                _ptr(reward), _ptr(terminal), _ptr(self.step_gs), _ptr(self.status), _ptr(self.steps),
                _ptr(self.episode), _ptr(self.p_thresh), _ptr(self.counters), flags, self.max_steps, int(mode),
                int(n_fixed), self.seed, self.env_base, _ptr(counts), _ptr(self.offsets), _ptr(self.workspace),
                self.workspace.numel(), self.d, self.n, _stream(self.device)), "tc_actor_pre")
            if between is not None:
                between()
            self.persp_write()
            return
        if self.single_pass and self.d <= SINGLE_PASS_MAX_D:
            flags |= ACTOR_SINGLE_PASS
        _lib.check(self.lib.tc_actor_step(
            _ptr(self.qubits), _ptr(self.syn), _ptr(self.prev_syn if keep_prev else None), _ptr(act_buf),
            _ptr(reward), _ptr(terminal), _ptr(self.step_gs), _ptr(self.status), _ptr(self.steps),
            _ptr(self.episode), _ptr(self.p_thresh), _ptr(self.counters), flags, self.max_steps, int(mode),
            int(n_fixed), self.seed, self.env_base, _ptr(self.out), _ptr(self.pos), _ptr(counts),
            _ptr(self.offsets), _ptr(self.overflow), self.cap, _ptr(self.workspace), self.workspace.numel(),
            self.d, self.n, _stream(self.device)), "tc_actor_step")

    def capture_actor_steps(self, k=1, **kwargs):
        """Capture k fused actor steps (actor_step(**kwargs), device-drawn actions by default) into one CUDA
        graph and return it; `graph.replay()` then advances every env by k steps with a single launch -- for
        small batches, where three kernel launches per step are launch-bound.  Capturing runs nothing, so the
        env state is unchanged; the programmatic-dependent-launch edges between the kernels are kept."""
        if kwargs.get("actions") is not None and kwargs["actions"] is not self.actions:
            raise ValueError("a captured step reads self.actions (fill it in place) or draws actions on device")
        if self.out is None:
            self.ensure_capacity(self.worst_case_capacity())
        if kwargs.get("keep_prev") and self.prev_syn is None:
            self.prev_syn = torch.zeros_like(self.syn)
        self.perspectives()            # first-launch setup of the writer (function attributes) outside the capture
        torch.cuda.synchronize(self.device)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            for _ in range(int(k)):
                self.actor_step(**kwargs)
        return graph

    def actor_pre(self, actions=None, auto_reset=True, mode=SAMPLE_DEPOLARIZING, n_fixed=0):
        """First half of actor_step (step / bookkeeping / count / scan) -- used with persp_write() by
        callers that time the two halves separately (bench.py)."""
        flags = (ACTOR_AUTO_RESET if auto_reset else 0)
        if actions is None:
            flags |= ACTOR_RANDOM_ACTIONS
        else:
            self.actions.copy_(torch.as_tensor(actions).to(torch.int32), non_blocking=True)
        self._last_syn = self.syn
        _lib.check(self.lib.tc_actor_pre(
            _ptr(self.qubits), _ptr(self.syn), _ptr(None), _ptr(self.actions), _ptr(self.reward),
            _ptr(self.terminal), _ptr(self.step_gs), _ptr(self.status), _ptr(self.steps), _ptr(self.episode),
            _ptr(self.p_thresh), _ptr(self.counters), flags, self.max_steps, int(mode), int(n_fixed), self.seed,
            self.env_base, _ptr(self.counts), _ptr(self.offsets), _ptr(self.workspace), self.workspace.numel(),
            self.d, self.n, _stream(self.device)), "tc_actor_pre")

    def persp_write(self):
        """Second half: the streaming perspective write for the state left by actor_pre()."""
        if self.out is None:
            self.ensure_capacity(self.worst_case_capacity())
        _lib.check(self.lib.tc_persp_write_as(_ptr(self.syn), _ptr(self.out), OUT_DTYPES[self.out_dtype],
                                              _ptr(self.pos), _ptr(self.offsets), _ptr(self.overflow), self.cap,
                                              _ptr(self.workspace), self.workspace.numel(), self.d, self.n,
                                              _stream(self.device)), "tc_persp_write_as")

    # -- results ---------------------------------------------------------------------------------
    def total_perspectives(self, n=None):
        """Host read of offsets[n] (synchronises)."""
        n = self.n if n is None else n
        return int(self.offsets[n].item())

    def perspective_views(self, n=None):
        """(persp out_dtype (sumP,2,d,d), pos uint8 (sumP,4), counts int32 (n,), offsets int64 (n+1,)) as
        views of the engine buffers.  Synchronises to learn sumP; regenerates if the buffer overflowed."""
        n = self.n if n is None else n
        total = self.total_perspectives(n)
        if int(self.overflow.item()) != 0:
            self.ensure_capacity(total)
            self.perspectives(syn=getattr(self, "_last_syn", None), n=n)
            total = self.total_perspectives(n)
        return self.out[:total], self.pos[:total], self.counts[:n], self.offsets[:n + 1]

    # -- checkpoint / resume ---------------------------------------------------------------------
    def state_dict(self):
        """Everything a bit-exact resume needs (SURVEY.md section 5 "Checkpoint / resume"): the packed planes, the
        per-env episode / step clocks, flags, thresholds and counters (cloned, on the device) plus the scalar
        configuration.  The sampler is counter-based -- Philox keyed by (seed, global env id, episode, step) -- so there
        is no generator state beyond `episode` and `steps`.  The perspective buffer is not part of the state: it is a
        function of `syn` (perspectives() regenerates it)."""
        sd = {name: getattr(self, name).clone() for name in self._STATE_TENSORS}
        if self.prev_syn is not None:
            sd["prev_syn"] = self.prev_syn.clone()
        sd["meta"] = {"d": self.d, "n": self.n, "seed": self.seed, "env_base": self.env_base,
                      "max_steps": self.max_steps, "out_dtype": str(self.out_dtype).replace("torch.", ""),
                      "abi_version": _lib.ABI_VERSION}
        return sd

    def load_state_dict(self, sd):
        """Restore a state_dict() into this engine (same d and n; seed / env_base / max_steps are taken from the
        checkpoint so that the resumed trajectory is the saved one)."""
        meta = sd["meta"]
        if int(meta["d"]) != self.d or int(meta["n"]) != self.n:
            raise ValueError("checkpoint is for d=%s n=%s, engine is d=%d n=%d" % (meta["d"], meta["n"], self.d, self.n))
        if int(meta.get("abi_version", _lib.ABI_VERSION)) != _lib.ABI_VERSION:
            raise ValueError("checkpoint was written under sampler / ABI version %s, this build is %d: the draw "
                             "numbering differs, a resumed run would not continue the saved trajectory"
                             % (meta.get("abi_version"), _lib.ABI_VERSION))
        self.seed, self.env_base, self.max_steps = int(meta["seed"]), int(meta["env_base"]), int(meta["max_steps"])
        for name in self._STATE_TENSORS:
            getattr(self, name).copy_(sd[name])
        if "prev_syn" in sd:
            if self.prev_syn is None:
                self.prev_syn = torch.zeros_like(self.syn)
            self.prev_syn.copy_(sd["prev_syn"])

    def read_counters(self):
        c = self.counters.cpu().numpy()
        return {name: int(c[i]) for i, name in enumerate(COUNTER_NAMES)}
