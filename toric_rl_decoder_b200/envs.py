"""Reference-shaped environment classes on top of the CUDA engine.

`ToricCodeB200` mirrors gym_ToricCode's `ToricCode` (gym_ToricCode/gym_ToricCode/envs/ToricCode.py:8-344)
and `EnvSetB200` mirrors `EnvSet` (src/EnvSet.py:4-51): same method names, argument meaning and return
dtypes (int64 states, float64 rewards, bool terminals), so the reference's callers
(src/Actor_mp.py:48-62,116,171-183; src/evaluation.py:51-113) run unchanged.  All arithmetic happens in
libtoric_b200.so on the GPU; there is no CPU path.
"""
from collections import namedtuple

import numpy as np
import torch

from . import _lib
from .engine import (SAMPLE_DEPOLARIZING, SAMPLE_FIXED_WEIGHT, SAMPLE_ONE_DRAW, ToricEngine, _ptr, _stream)

_Box = namedtuple("Box", ["low", "high", "shape"])


def _config(config):
    size = config["size"] if "size" in config else 3
    # the env reads 'min_qbit_errors' (ToricCode.py:39); the reference's callers misspell it
    # 'min_qubit_errors' (Distributed_mp.py:73-76, evaluation.py:51) so the env silently uses 0: same here.
    min_err = config["min_qbit_errors"] if "min_qbit_errors" in config else 0
    p = config["p_error"] if "p_error" in config else 0.1
    return int(size), int(min_err), float(p)


class ToricCodeB200:
    """Single environment = engine batch of one.  API of ToricCode (ToricCode.py:24-88, :90-218, :289-337).

    `qubit_matrix` and `state` are read/write properties that unpack / pack the device bit planes; the
    setters are the mask-injection interface the reference's scripts use
    (results/small_p_error_test.py:119-120, results/start_from_state.py:36-38)."""

    metadata = {"render.modes": ["human"]}

    def __init__(self, config=None, device="cuda", seed=0, env_base=0):
        config = config or {}
        self.system_size, self.min_qbit_errors, self.p_error = _config(config)
        d = self.system_size
        self.action_space = _Box(np.array([0, 0, 0, 0]), np.array([1, d, d, 3]), (4,))
        self.observation_space = _Box(0, 1, (2, d, d))
        self.rule_table = np.array([[0, 1, 2, 3], [1, 0, 3, 2], [2, 3, 0, 1], [3, 2, 1, 0]], dtype=int)
        self._eng = ToricEngine(d, 1, device=device, seed=seed, env_base=env_base, p_error=self.p_error)
        self._scratch = None      # second engine for the pure functions (createSyndromOpt, generate*Error*)
        self._state_assigned = False   # `state` was assigned since the last step / reset (see the setter)
        self.next_state = self.state

    # -- attribute surface -------------------------------------------------------------------------
    @property
    def qubit_matrix(self):
        return self._eng.get_qubits()[0].cpu().numpy().astype(np.int64)

    @qubit_matrix.setter
    def qubit_matrix(self, q):
        """Mask injection.  The reference assigns `qubit_matrix` and `state` as two independent attributes
        (results/small_p_error_test.py:119-120) and leaves `ground_state` alone; here the syndrome additionally follows
        the qubits at once (the reference's `state` would stay stale until it is assigned or a step recomputes it) and
        `ground_state` is, like there, not touched."""
        e = self._eng
        t = e._as_u8(np.asarray(q).astype(np.uint8)[None], (1, 2, e.d, e.d))
        _lib.check(e.lib.tc_pack_qubits(_ptr(t), _ptr(e.qubits), e.d, 1, _stream(e.device)), "tc_pack_qubits")
        e.syndrome()

    @property
    def state(self):
        return self._eng.get_state()[0].cpu().numpy().astype(np.int64)

    @state.setter
    def state(self, s):
        """Assigning `state` may leave it INCONSISTENT with `qubit_matrix` (the reference allows it: two plain
        attributes).  The reference's next step then takes its reward from the assigned state but recomputes the new
        state from the qubits (ToricCode.py:83-86); the next step() here does the same instead of the incremental
        update."""
        self._eng.set_state((np.asarray(s) != 0).astype(np.uint8)[None])
        self._state_assigned = True

    @property
    def ground_state(self):
        return bool(self._eng.gs[0].item())

    @ground_state.setter
    def ground_state(self, v):
        self._eng.gs.fill_(1 if v else 0)

    # -- gym API -----------------------------------------------------------------------------------
    def reset(self, p_error=None):
        """ToricCode.reset (ToricCode.py:90-127)."""
        if p_error is not None:
            self.p_error = p_error
        self._state_assigned = False
        self._eng.set_p_error(self.p_error)
        if self.min_qbit_errors == 0:
            self._eng.reset(mode=SAMPLE_DEPOLARIZING)
        else:
            self._eng.reset(mode=SAMPLE_FIXED_WEIGHT, n_fixed=self.min_qbit_errors)
        return self.state

    def step(self, action):
        """ToricCode.step (ToricCode.py:62-88): returns (state, reward, terminal, {})."""
        a = np.asarray(action, dtype=np.int64).reshape(1, 4)
        if self._state_assigned:      # reward against the assigned state, new state recomputed from the qubits
            before = int(self.state.sum())
            self._eng.step(torch.from_numpy(a))
            self._eng.syndrome()
            self._state_assigned = False
            self.next_state = self.state
            after = int(self.next_state.sum())
            return self.next_state, (100 if after == 0 else before - after), after == 0, {}
        reward, terminal = self._eng.step(torch.from_numpy(a))
        r = reward[0].item()
        self.next_state = self.state
        return self.next_state, int(r), bool(terminal[0].item()), {}

    def _scratch_engine(self):
        if self._scratch is None:
            e = self._eng
            # its own Philox stream: same seed, env id far outside any batch
            self._scratch = ToricEngine(self.system_size, 1, device=e.device, seed=e.seed,
                                        env_base=e.env_base + (1 << 40))
        return self._scratch

    def createSyndromOpt(self, tcode):
        """createSyndromOpt (ToricCode.py:168-199) of an arbitrary qubit matrix (scratch engine state)."""
        eng = self._scratch_engine()
        eng.set_qubits(np.asarray(tcode).astype(np.uint8)[None])
        return eng.get_state()[0].cpu().numpy().astype(np.int64)

    def isTerminalState(self, state):
        return bool(np.all(np.asarray(state) == 0))

    def evalGroundState(self):
        """evalGroundState (ToricCode.py:289-337), sticky until reset."""
        return bool(self._eng.ground_state()[0].item())

    def generateRandomError(self, matrix, p_error):
        """generateRandomError (ToricCode.py:130-151): a fresh (2,d,d) matrix with independent errors of
        probability p_error and uniform Pauli codes 1..3 -- one draw of the device sampler, no rejection;
        like the reference it ignores the contents of `matrix` and leaves the env untouched."""
        eng = self._scratch_engine()
        eng.set_p_error(float(p_error))
        eng.reset(mode=SAMPLE_DEPOLARIZING | SAMPLE_ONE_DRAW)
        return eng.get_qubits()[0].cpu().numpy().astype(np.int64)

    def generateNRandomErrors(self, matrix, n):
        """generateNRandomErrors (ToricCode.py:153-166): exactly n errors on a uniform random n-subset of the
        2d^2 qubits, uniform codes 1..3; written into `matrix` in place and returned, as the reference does."""
        eng = self._scratch_engine()
        if int(n) <= 0:
            matrix[:, :, :] = 0
            return matrix
        eng.reset(mode=SAMPLE_FIXED_WEIGHT | SAMPLE_ONE_DRAW, n_fixed=int(n))
        matrix[:, :, :] = eng.get_qubits()[0].cpu().numpy()
        return matrix

    def plotToricCode(self, state, title):
        """ToricCode.plotToricCode (ToricCode.py:220-286) draws the lattice with matplotlib and saves a PDF: no
        graphics stack -- unavailable on this box.  Accepted and ignored, so that the reference's callers
        (src/evaluation.py:72,103 with plot_one_episode=True, results/start_from_state.py:40,58) keep running."""
        return None

    def render(self, mode="human"):
        pass

    def close(self):
        pass


class _EnvView:
    """`EnvSet.envs[i]` (src/EnvSet.py:14-16: the reference keeps a list of deep-copied envs): env i of the batch
    with ToricCode's attribute surface.  Every per-env buffer of the engine is env-major and contiguous, so row i is
    itself a valid batch of one: the kernels run on that slice, the other envs are not touched."""

    def __init__(self, envset, i):
        self._es, self._i = envset, int(i)
        d = self.system_size = envset.size
        self.min_qbit_errors = envset.min_qbit_errors
        self.action_space = _Box(np.array([0, 0, 0, 0]), np.array([1, d, d, 3]), (4,))
        self.observation_space = _Box(0, 1, (2, d, d))
        self.rule_table = np.array([[0, 1, 2, 3], [1, 0, 3, 2], [2, 3, 0, 1], [3, 2, 1, 0]], dtype=int)

    @property
    def next_state(self):
        """ToricCode.next_state (ToricCode.py:50,86): the state the env's last step produced (views are made on
        demand, so the batch keeps it); before any step it is the current state."""
        ns = self._es._next_states.get(self._i)
        return self.state if ns is None else ns

    @next_state.setter
    def next_state(self, s):
        self._es._next_states[self._i] = np.asarray(s)

    # the generators and the stencil as functions of their arguments (ToricCode.py:130-199): they do not touch this
    # env, so they run on the batch's one-env helper
    def createSyndromOpt(self, tcode):
        return self._es._helper().createSyndromOpt(tcode)

    def generateRandomError(self, matrix, p_error):
        return self._es._helper().generateRandomError(matrix, p_error)

    def generateNRandomErrors(self, matrix, n):
        return self._es._helper().generateNRandomErrors(matrix, n)

    def plotToricCode(self, state, title):
        """No graphics stack -- unavailable on this box (see ToricCodeB200.plotToricCode)."""
        return None

    def _rows(self):
        e, i = self._es.engine, self._i
        return e, e.qubits[i:i + 1], e.syn[i:i + 1]

    @property
    def p_error(self):
        th = int(self._es.engine.p_thresh[self._i].item()) & 0xffffffff
        return th / 4294967296.0

    @property
    def qubit_matrix(self):
        e, q, _ = self._rows()
        out = torch.empty((1, 2, e.d, e.d), dtype=torch.uint8, device=e.device)
        _lib.check(e.lib.tc_unpack_qubits(_ptr(q), _ptr(out), e.d, 1, _stream(e.device)), "tc_unpack_qubits")
        return out[0].cpu().numpy().astype(np.int64)

    @qubit_matrix.setter
    def qubit_matrix(self, codes):
        e, q, syn = self._rows()
        t = e._as_u8(np.asarray(codes).astype(np.uint8)[None], (1, 2, e.d, e.d))
        _lib.check(e.lib.tc_pack_qubits(_ptr(t), _ptr(q), e.d, 1, _stream(e.device)), "tc_pack_qubits")
        _lib.check(e.lib.tc_syndrome(_ptr(q), _ptr(syn), e.d, 1, _stream(e.device)), "tc_syndrome")
        # ground_state is left alone, as in the reference (plain attribute assignment)

    @property
    def state(self):
        e, _, syn = self._rows()
        return e.get_state(syn)[0].cpu().numpy().astype(np.int64)

    @state.setter
    def state(self, s):
        e, _, syn = self._rows()
        t = e._as_u8((np.asarray(s) != 0).astype(np.uint8)[None], (1, 2, e.d, e.d))
        _lib.check(e.lib.tc_pack_state(_ptr(t), _ptr(syn), e.d, 1, _stream(e.device)), "tc_pack_state")
        self._es._state_assigned.add(self._i)       # see ToricCodeB200.state: the next step recomputes from the qubits

    @property
    def ground_state(self):
        return bool(self._es.engine.gs[self._i].item())

    @ground_state.setter
    def ground_state(self, v):
        self._es.engine.gs[self._i] = 1 if v else 0

    def reset(self, p_error=None):
        """ToricCode.reset on this env only (what EnvSet.resetTerminalEnvs loops over, src/EnvSet.py:19-26)."""
        if p_error is not None:
            self._es.engine.set_p_error(float(p_error), idx=[self._i])
        self._es._reset(np.array([self._i], dtype=np.int64))
        return self.state

    def step(self, action):
        """ToricCode.step on this env only: (state, reward, terminal, {})."""
        e, q, syn = self._rows()
        i = self._i
        a = torch.as_tensor(np.asarray(action, dtype=np.int32).reshape(1, 4)).to(e.device)
        assigned = i in self._es._state_assigned
        before = int(self.state.sum()) if assigned else 0
        _lib.check(e.lib.tc_step(_ptr(q), _ptr(syn), _ptr(a), _ptr(e.reward[i:i + 1]), _ptr(e.terminal[i:i + 1]),
                                 _ptr(e.status[i:i + 1]), e.d, 1, _stream(e.device)), "tc_step")
        e.steps[i] += 1
        if assigned:                  # ToricCode.py:83-86: new state from the qubits, reward against the assigned one
            self._es._state_assigned.discard(i)
            _lib.check(e.lib.tc_syndrome(_ptr(q), _ptr(syn), e.d, 1, _stream(e.device)), "tc_syndrome")
            st = self.next_state = self.state
            after = int(st.sum())
            return st, (100 if after == 0 else before - after), after == 0, {}
        st = self.next_state = self.state
        return st, int(e.reward[i].item()), bool(e.terminal[i].item()), {}

    def evalGroundState(self):
        e, q, _ = self._rows()
        _lib.check(e.lib.tc_ground_state(_ptr(q), _ptr(e.gs[self._i:self._i + 1]), e.d, 1, _stream(e.device)),
                   "tc_ground_state")
        return self.ground_state

    def isTerminalState(self, state):
        return bool(np.all(np.asarray(state) == 0))


class _EnvList:
    """The `envs` list of the reference's EnvSet as a lazy sequence (a batch may hold millions of envs)."""

    def __init__(self, envset):
        self._es = envset

    def __len__(self):
        return self._es.no_envs

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [_EnvView(self._es, k) for k in range(*i.indices(len(self)))]
        i = int(i)
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError("env index out of range")
        return _EnvView(self._es, i)

    def __iter__(self):
        return (_EnvView(self._es, k) for k in range(len(self)))


class EnvSetB200:
    """EnvSet (src/EnvSet.py:4-51) over one engine batch.

    With numpy_compat=True (default) results come back as the reference's numpy dtypes; with False they
    stay on the device as torch tensors (uint8 states, float32 rewards, bool terminals) and nothing
    synchronises."""

    def __init__(self, env, no_envs, device="cuda", seed=0, env_base=0, numpy_compat=True):
        if isinstance(env, dict):
            size, min_err, p = _config(env)
        else:
            size, min_err, p = env.system_size, getattr(env, "min_qbit_errors", 0), env.p_error
        self.size = size
        self.no_envs = int(no_envs)
        self.min_qbit_errors = int(min_err)
        self.numpy_compat = numpy_compat
        self.engine = ToricEngine(size, self.no_envs, device=device, seed=seed, env_base=env_base, p_error=p)
        self.envs = _EnvList(self)
        self._state_assigned = set()     # envs whose `state` was assigned through envs[i] since their last step / reset
        self._next_states = {}           # envs[i].next_state of the envs stepped through their view
        self._one = None                 # one-env helper behind envs[i].createSyndromOpt / generate*Error*
        self.states = self._states()

    def _states(self):
        st = self.engine.get_state()
        return st.cpu().numpy().astype(np.int64) if self.numpy_compat else st

    def _helper(self):
        if self._one is None:
            e = self.engine
            self._one = ToricCodeB200({"size": self.size, "min_qbit_errors": self.min_qbit_errors}, device=e.device,
                                      seed=e.seed, env_base=e.env_base + (1 << 41))     # a Philox stream of its own
        return self._one

    def _reset(self, idx):
        if idx is None:
            self._state_assigned.clear()
        elif self._state_assigned:
            self._state_assigned.difference_update(int(i) for i in np.asarray(idx).reshape(-1))
        if self.min_qbit_errors == 0:
            self.engine.reset(idx, mode=SAMPLE_DEPOLARIZING)
        else:
            self.engine.reset(idx, mode=SAMPLE_FIXED_WEIGHT, n_fixed=self.min_qbit_errors)

    def resetAll(self, p_errors=None):
        """EnvSet.resetAll (src/EnvSet.py:28-36)."""
        if p_errors is not None:
            self.engine.set_p_error(np.asarray(p_errors, dtype=np.float64))
        self._reset(None)
        self.states = self._states()
        return self.states

    def resetTerminalEnvs(self, idx, p_errors=None):
        """EnvSet.resetTerminalEnvs (src/EnvSet.py:19-26): returns float64 states of the listed envs."""
        if torch.is_tensor(idx):                # numpy_compat=False callers keep the index list on the device
            it = idx.to(device=self.engine.device, dtype=torch.int64).reshape(-1)
            idx = it.cpu().numpy() if (p_errors is not None or self._state_assigned) else it
        else:
            idx = np.asarray(idx, dtype=np.int64).reshape(-1)
            it = torch.as_tensor(idx, device=self.engine.device)
        if p_errors is not None:
            self.engine.set_p_error(np.asarray(p_errors, dtype=np.float64), idx=idx)
        self._reset(idx)
        st = self.engine.get_state(self.engine.syn[it].contiguous())
        return st.cpu().numpy().astype(np.float64) if self.numpy_compat else st

    def step(self, actions):
        """EnvSet.step (src/EnvSet.py:38-47): (states, rewards, terminals, info)."""
        a = torch.as_tensor(np.asarray(actions, dtype=np.int64) if not torch.is_tensor(actions) else actions)
        e = self.engine
        assigned = sorted(self._state_assigned)
        if assigned:                  # defect counts of the states assigned through envs[i].state = ...
            it = torch.as_tensor(assigned, dtype=torch.int64, device=e.device)
            before = e.get_state(e.syn[it].contiguous()).flatten(1).sum(1, dtype=torch.float32)
        reward, terminal = e.step(a)
        if assigned:
            # ToricCode.py:83-86 for those envs, as ToricCodeB200.step / _EnvView.step do: the new state is recomputed
            # from the qubits (the incremental update ran on a state that need not match them), the reward is
            # taken against the assigned state
            for i in assigned:
                _lib.check(e.lib.tc_syndrome(_ptr(e.qubits[i:i + 1]), _ptr(e.syn[i:i + 1]), e.d, 1, _stream(e.device)),
                           "tc_syndrome")
            after = e.get_state(e.syn[it].contiguous()).flatten(1).sum(1, dtype=torch.float32)
            done = after == 0
            reward[it] = torch.where(done, torch.full_like(after, 100.0), before - after)
            terminal[it] = done.to(terminal.dtype)
            self._state_assigned.clear()
        st = e.get_state()
        if self.numpy_compat:
            return (st.cpu().numpy().astype(np.int64), reward.cpu().numpy().astype(np.float64),
                    terminal.cpu().numpy().astype(bool), {})
        return st, reward, terminal.bool(), {}


ENV_ID = "toric-code-b200-v0"


def make(env_id=ENV_ID, config=None, **kwargs):
    """gym.make('toric-code-v0', config=...) (gym_ToricCode/__init__.py:3-6) without needing gym: both the
    reference's id and ENV_ID give a ToricCodeB200."""
    if env_id not in (ENV_ID, "toric-code-v0"):
        raise ValueError("unknown environment id %r" % (env_id,))
    return ToricCodeB200(config, **kwargs)


def _register_with_gym():
    """If gym is importable, `gym.make(ENV_ID, config=...)` works like the reference's registration."""
    try:
        from gym.envs.registration import register
        register(id=ENV_ID, entry_point="toric_rl_decoder_b200.envs:ToricCodeB200")
        return True
    except Exception:
        return False


_register_with_gym()
