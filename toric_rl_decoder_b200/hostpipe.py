"""Host-buffer front end of the fused actor step: the reference-facing call with HOST arrays.

The reference's actor loop hands `EnvSet.step` a host array of actions and reads rewards, terminals and the
new states back as host arrays (src/Actor_mp.py:116-125, src/EnvSet.py:38-47).  `HostActorPipeline` keeps
that contract on the GPU engine and overlaps the PCIe copies with the kernels:

    h2d stream:      actions(k+1)
    compute stream:  step+count(k) | perspective writer(k)       | step+count(k+1) | writer(k+1) ...
    d2h stream:                    | unpack(k) | results(k)                        | unpack(k+1) | results(k+1)

(two copy streams: on a single one the upload of step k+1 would queue behind the download of step k, which
waits for compute k, and nothing would overlap; both have a HIGHER PRIORITY than the compute stream -- otherwise the
download's kernel work queues behind the writer's CTAs and the results leave the GPU only when the writer is done)

Per slot (ring of `depth`): a device action buffer, one contiguous device staging block
[reward f32 | counts i32 | terminal u8 | states] and its pinned host twin, so every step moves exactly one H2D and
one D2H memcpy; the whole step is enqueued by one native call (tc_host_actor_step).  `states` is the row-packed form
by default (written by the step kernel itself, seen on the host through a LazyStates view) or the reference's
one-byte-per-stabiliser array.  The perspective tensor stays on the device, where the reference also puts it
(`.to(device)`, src/numba/util_actor.py:39).  depth=1 is the closed loop of EnvSet.step's contract.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .engine import ACTOR_AUTO_RESET, _Range


class LazyStates:
    """The `states` array of EnvSet.step (src/EnvSet.py:39,47: (n,2,d,d) of 0/1) kept on the host in its
    row-packed download format and expanded only where a caller looks.

    Indexing follows numpy: the first index selects envs (int, slice, integer / boolean array) and is applied to
    the packed rows, the selected envs are expanded to uint8 0/1, the remaining indices are applied to that.
    `np.asarray(view)` / `view.astype(...)` expand every env.  The view aliases the pipeline's pinned download
    slot: it is valid until that slot is reused (`depth` submits later); `copy()` detaches it."""

    dtype = np.dtype(np.uint8)
    ndim = 4

    def __init__(self, rows, d, owned=False):
        self._rows, self._d = rows, int(d)
        self._owned = bool(owned)          # False: `rows` aliases a pipeline download slot and must not be written
        self.shape = (rows.shape[0], 2, self._d, self._d)

    def __len__(self):
        return self.shape[0]

    def _expand(self, rows):
        return np.unpackbits(rows, axis=-1, count=self._d, bitorder="little")

    def __getitem__(self, idx):
        first, rest = (idx[0], idx[1:]) if isinstance(idx, tuple) and len(idx) else (idx, ())
        if first is Ellipsis or (isinstance(idx, tuple) and not len(idx)):
            return np.asarray(self)[idx]
        sub = self._expand(self._rows[first])
        if not rest:
            return sub
        if isinstance(first, (int, np.integer)):
            return sub[rest]
        return sub[(slice(None),) + tuple(rest)]

    def __array__(self, dtype=None, copy=None):
        a = self._expand(self._rows)
        return a if dtype is None else a.astype(dtype, copy=False)

    def astype(self, dtype):
        return np.asarray(self).astype(dtype)

    def __setitem__(self, idx, value):
        """`next_state[idx] = reset_states` (src/Actor_mp.py:182) and numpy's other assignment forms.  The first
        write detaches the view from the pipeline's download slot (one copy of the packed rows, 1/7 of the expanded
        array at d=7); the assigned envs are stored packed again, any non-zero value counting as a defect."""
        if not self._owned:
            self._rows = self._rows.copy()
            self._owned = True
        first, rest = (idx[0], idx[1:]) if isinstance(idx, tuple) and len(idx) else (idx, ())
        if first is Ellipsis or (isinstance(idx, tuple) and not len(idx)):
            full = np.asarray(self)
            full[idx] = np.asarray(value) != 0
            self._rows[...] = np.packbits(full, axis=-1, bitorder="little")
            return
        if not rest:
            sub = np.empty(self._rows[first].shape[:-1] + (self._d,), dtype=np.uint8)
            sub[...] = np.asarray(value) != 0
        else:
            sub = self._expand(self._rows[first])
            key = tuple(rest) if isinstance(first, (int, np.integer)) else (slice(None),) + tuple(rest)
            sub[key] = np.asarray(value) != 0
        self._rows[first] = np.packbits(sub, axis=-1, bitorder="little")

    def copy(self):
        return LazyStates(self._rows.copy(), self._d, owned=True)

    @property
    def packed_rows(self):
        """uint8 (n,2,d,ceil(d/8)) = np.packbits(states, axis=-1, bitorder="little")."""
        return self._rows


class HostActorPipeline:
    def __init__(self, engine, depth=2, states="lazy"):
        """states: how each step's new states reach the host (rewards, terminals and perspective counts
        always do):
          "lazy" (default) the row-packed form crosses PCIe (ceil(d/8) bytes per lattice row, 7x fewer bytes than
                 one byte per stabiliser at d=7) and `collect()["states"]` is a `LazyStates` view with the
                 reference's (n,2,d,d) shape that expands rows -> 0/1 bytes on the host only for the envs a
                 caller indexes;
          "u8"   uint8 (n,2,d,d) of 0/1 -- the reference's array (src/EnvSet.py:39,47), one byte per stabiliser,
                 expanded on the device and downloaded whole (PCIe-bound beyond ~2 GPUs per host);
          "rows" uint8 (n,2,d,ceil(d/8)) = np.packbits(states, axis=-1, bitorder="little") as `state_rows`
                 (`unpack_rows` restores the (n,2,d,d) array);
          None   states stay on the device."""
        if states not in ("lazy", "u8", "rows", None):
            raise ValueError('states must be "lazy", "u8", "rows" or None')
        self.eng = engine
        self.depth = int(depth)
        self.states = states
        n, d = engine.n, engine.d
        dev = engine.device
        # high priority: the upload and the result download must not queue behind the perspective writer's CTAs
        self.h2d_stream = torch.cuda.Stream(device=dev, priority=-1)
        self.d2h_stream = torch.cuda.Stream(device=dev, priority=-1)
        self.o_reward = 0
        self.o_counts = 4 * n
        self.o_term = 8 * n
        self.o_state = (9 * n + 15) & ~15
        self.row_bytes = (d + 7) // 8
        self.state_bytes = {"u8": n * 2 * d * d, "rows": n * 2 * d * self.row_bytes,
                            "lazy": n * 2 * d * self.row_bytes, None: 0}[states]
        self.nbytes = self.o_state + self.state_bytes
        self.h2d_bytes = n * 16
        self.d2h_bytes = self.nbytes
        self.slots = []
        for _ in range(self.depth):
            slot = {
                "d_actions": torch.empty((n, 4), dtype=torch.int32, device=dev),
                "d_stage": torch.empty(self.nbytes, dtype=torch.uint8, device=dev),
                "h_stage": torch.empty(self.nbytes, dtype=torch.uint8).pin_memory(),
                "ev_actions": torch.cuda.Event(), "ev_compute": torch.cuda.Event(), "ev_result": torch.cuda.Event(),
                "ev_unpacked": torch.cuda.Event(),
                "busy": False,
            }
            for name in ("ev_actions", "ev_compute", "ev_result", "ev_unpacked"):
                slot[name].record(self.h2d_stream)        # torch creates the cudaEvent_t on first use
            hs = slot["h_stage"].numpy()
            slot["views"] = {
                "reward": hs[self.o_reward:self.o_reward + 4 * n].view(np.float32),
                "counts": hs[self.o_counts:self.o_counts + 4 * n].view(np.int32),
                "terminal": hs[self.o_term:self.o_term + n].view(np.bool_),
            }
            if states == "u8":
                slot["views"]["states"] = hs[self.o_state:self.o_state + self.state_bytes].reshape(n, 2, d, d)
            elif states in ("rows", "lazy"):
                rows = hs[self.o_state:self.o_state + self.state_bytes].reshape(n, 2, d, self.row_bytes)
                slot["views"]["state_rows"] = rows
                if states == "lazy":
                    slot["views"]["states"] = LazyStates(rows, d)
            ds = slot["d_stage"]
            slot["d_reward"] = ds[self.o_reward:self.o_reward + 4 * n].view(torch.float32)
            slot["d_counts"] = ds[self.o_counts:self.o_counts + 4 * n].view(torch.int32)
            slot["d_term"] = ds[self.o_term:self.o_term + n]
            slot["d_state"] = ds[self.o_state:self.o_state + self.state_bytes]
            self.slots.append(slot)
        self.submitted = 0
        self.collected = 0
        self._ev_unpacked = None

    def _native_block(self, slot):
        """The slot's tc_host_step argument block (include/toric_b200.h): everything that does not change from step to
        step is filled once; submit() only updates the action pointer, the flags and the event of the previous step."""
        from .engine import OUT_DTYPES
        eng = self.eng
        hs = _lib.HostStep()
        for name in ("qubits", "syn", "status", "steps", "episode", "p_thresh", "counters", "offsets", "overflow",
                     "out", "pos", "workspace"):
            setattr(hs, name, getattr(eng, name).data_ptr())
        hs.gs = eng.step_gs.data_ptr()
        hs.prev_syn = None
        hs.cap, hs.out_dtype = eng.cap, OUT_DTYPES[eng.out_dtype]
        hs.workspace_bytes = eng.workspace.numel()
        hs.max_steps, hs.mode, hs.n_fixed = eng.max_steps, 0, 0
        hs.seed, hs.env_base, hs.d, hs.n = eng.seed, eng.env_base, eng.d, eng.n
        hs.d_actions, hs.d_stage = slot["d_actions"].data_ptr(), slot["d_stage"].data_ptr()
        hs.h_stage, hs.stage_bytes = slot["h_stage"].data_ptr(), self.nbytes
        hs.o_reward, hs.o_counts, hs.o_term, hs.o_state = self.o_reward, self.o_counts, self.o_term, self.o_state
        hs.state_mode = {"u8": 1, "rows": 2, "lazy": 2, None: 0}[self.states]
        hs.h2d_stream, hs.d2h_stream = self.h2d_stream.cuda_stream, self.d2h_stream.cuda_stream
        for name in ("ev_actions", "ev_compute", "ev_unpacked", "ev_result"):
            setattr(hs, name, slot[name].cuda_event)
        return hs

    def submit(self, actions_host, auto_reset=True, mode=0, n_fixed=0):
        """Enqueue one step.  actions_host: pinned int32 tensor (n,4).  Never blocks the host unless the
        ring is full (then the oldest result must be collected first).

        The step's results (reward, terminal, counts, new states) are final once the step / count kernels have
        run, i.e. BEFORE the perspective writer: an event is recorded there, and the state kernel and the download
        run on the d2h stream behind that event, overlapping the writer.  The next step's first kernel waits for
        the state kernel (it rewrites the syndromes that kernel reads).  The whole sequence -- upload, event edges,
        three kernels, state kernel, download -- is enqueued by ONE native call (tc_host_actor_step): a closed-loop
        caller's turn-around is then the PCIe round trip, not a few hundred microseconds of Python."""
        if self.submitted - self.collected >= self.depth:
            raise RuntimeError("pipeline full: collect() before submitting again")
        eng = self.eng
        if actions_host.dtype != torch.int32 or tuple(actions_host.shape) != (eng.n, 4) or not actions_host.is_contiguous():
            raise ValueError("actions_host must be a contiguous int32 tensor of shape (%d, 4)" % eng.n)
        if eng.out is None:
            eng.ensure_capacity(eng.worst_case_capacity())
        slot = self.slots[self.submitted % self.depth]
        hs = slot.get("native")
        if hs is None or hs.cap != eng.cap or hs.out != eng.out.data_ptr():
            hs = slot["native"] = self._native_block(slot)
        hs.h_actions = actions_host.data_ptr()
        hs.flags = ACTOR_AUTO_RESET if auto_reset else 0
        hs.mode, hs.n_fixed = int(mode), int(n_fixed)
        hs.compute_stream = torch.cuda.current_stream(eng.device).cuda_stream
        hs.ev_unpacked_prev = self._ev_unpacked.cuda_event if self._ev_unpacked is not None else None
        eng._last_syn = eng.syn
        with _Range("toric/host_actor_step", eng.nvtx):
            _lib.check(eng.lib.tc_host_actor_step(ctypes.byref(hs)), "tc_host_actor_step")
        if self.states is not None:
            self._ev_unpacked = slot["ev_unpacked"]
        self.submitted += 1

    def collect(self):
        """Block until the oldest outstanding step's results are in host memory; returns numpy views
        {reward f32 (n,), counts i32 (n,), terminal bool (n,), states (n,2,d,d) -- a LazyStates view ("lazy") or
        a uint8 array ("u8") -- and/or state_rows u8 (n,2,d,ceil(d/8))} valid until the slot is reused (depth
        submits later)."""
        if self.collected >= self.submitted:
            raise RuntimeError("nothing to collect")
        slot = self.slots[self.collected % self.depth]
        slot["ev_result"].synchronize()
        self.collected += 1
        return slot["views"]

    def unpack_rows(self, state_rows):
        """(n,2,d,ceil(d/8)) row bitmasks -> the reference's (n,2,d,d) 0/1 array (host, numpy)."""
        return np.unpackbits(state_rows, axis=-1, bitorder="little")[..., :self.eng.d]

    def step(self, actions_host, auto_reset=True):
        """Unpipelined convenience form: submit + collect."""
        self.submit(actions_host, auto_reset=auto_reset)
        return self.collect()
