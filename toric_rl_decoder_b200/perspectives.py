"""generatePerspective / generatePerspectiveOptimized / generatePerspectiveBatch on the GPU.

Same names and argument order as src/util.py:46-150 and src/numba/util_actor.py:56-67.  The batch form
returns the select-action input layout of src/numba/util_actor.py:35-39 directly: one float32 tensor
(sumP,2,d,d) on the device, positions (sumP,3), counts (N,) and offsets (N+1,) with
offsets[1:] == np.cumsum(counts) == the reference's splice_idx.
"""
from collections import namedtuple

import numpy as np
import torch

from .engine import ToricEngine

Perspective = namedtuple("Perspective", ["perspective", "position"])   # src/util.py:12
Action = namedtuple("Action", ["position", "action"])                  # src/util.py:9
Transition = namedtuple("Transition", ["state", "action", "reward", "next_state", "terminal"])   # src/util.py:14-15
action_type = np.dtype([("position", (int, 3)), ("op", int)])          # src/util.py:10 (np.int is gone in numpy 2)


def transition_type(d):
    """The structured dtype the reference's actor builds for its transition records (src/Actor_mp.py:52-56):
    what generateTransitionParallel(..., trans_type=...) fills."""
    return np.dtype([("perspective", (int, (2, d, d))), ("action", action_type), ("reward", float),
                     ("next_perspective", (int, (2, d, d))), ("terminal", bool)])

_engines = {}


def _engine(d, n, device):
    """Scratch engine per (d, device): only its workspace / counts / offsets are used here (a few bytes per env);
    the perspective tensors handed to callers are allocated per call."""
    key = (d, str(device))
    eng = _engines.get(key)
    if eng is None or eng.n < n:
        eng = ToricEngine(d, max(n, 1), device=device)
        _engines[key] = eng
    return eng


def _as_defect_bytes(st, dev):
    """One byte per stabiliser on `dev`, any non-zero byte meaning "defect" (tc_pack_state tests != 0 itself).  A
    uint8 / bool tensor that already lives there -- what EnvSetB200(numpy_compat=False) hands out -- is used as it is;
    anything else is compared with 0 first (int64 / float64 arrays of the reference's dtypes: wider than a byte)."""
    if st.dtype in (torch.uint8, torch.bool) and st.device == dev and st.is_contiguous():
        return st.view(torch.uint8) if st.dtype is torch.bool else st
    if st.dtype in (torch.uint8, torch.bool):
        return st.to(device=dev).contiguous().view(torch.uint8)
    return (st != 0).to(device=dev, dtype=torch.uint8).contiguous()


def generatePerspectiveBatch(grid_shift, toric_size, states, device="cuda", raw_positions=False):
    """states: (N,2,d,d) array/tensor of 0/1.  Returns (perspectives f32 (sumP,2,d,d) cuda tensor,
    positions int64 (sumP,3) cuda tensor, counts int32 (N,), offsets int64 (N+1,)).  raw_positions=True keeps
    the positions as the device layout uint8 (sumP,4) = (m,i,j,0) that tc_select_action consumes.

    Like the reference (src/numba/util_actor.py:56-67 returns fresh lists) every call returns FRESH tensors
    sized to sumP: count + scan first, then one allocation, then the writer -- a later call (perspectives of
    `next_state` after `state`, target network after policy network) never overwrites an earlier result."""
    from . import _lib
    from .engine import OUT_DTYPES, _ptr, _stream
    d = int(toric_size)
    if int(grid_shift) != d // 2:
        raise ValueError("grid_shift must be toric_size // 2 (src/Actor_mp.py:59)")
    st = torch.as_tensor(np.ascontiguousarray(states) if isinstance(states, np.ndarray) else states)
    n = st.shape[0]
    eng = _engine(d, n, device)
    dev = eng.device
    syn = torch.empty((n, 2, d), dtype=torch.int32, device=dev)        # tc_pack_state writes every word
    u8 = _as_defect_bytes(st, dev)
    with torch.cuda.device(dev):
        _lib.check(eng.lib.tc_pack_state(_ptr(u8), _ptr(syn), d, n, _stream(dev)), "tc_pack_state")
        eng.perspectives_count(syn=syn, n=n)
        total = eng.total_perspectives(n)                       # one 8-byte host read
        persp = torch.empty((total, 2, d, d), dtype=torch.float32, device=dev)
        pos = torch.empty((total, 4), dtype=torch.uint8, device=dev)
        overflow = torch.zeros(1, dtype=torch.int32, device=dev)
        if n > 0:
            _lib.check(eng.lib.tc_persp_write_as(_ptr(syn), _ptr(persp), OUT_DTYPES[torch.float32], _ptr(pos),
                                                 _ptr(eng.offsets), _ptr(overflow), total, _ptr(eng.workspace),
                                                 eng.workspace.numel(), d, n, _stream(dev)), "tc_persp_write_as")
        counts, offsets = eng.counts[:n].clone(), eng.offsets[:n + 1].clone()
    if raw_positions:
        return persp, pos, counts, offsets
    return persp, pos[:, :3].to(torch.int64), counts, offsets


def generatePerspectiveOptimized(grid_shift, toric_size, state, device="cuda"):
    """src/util.py:106-150: returns (list of (2,d,d) arrays, list of (m,i,j) tuples), same dtype as input."""
    state = np.asarray(state)
    persp, pos, _, _ = generatePerspectiveBatch(grid_shift, toric_size, state[None], device=device)
    views = persp.cpu().numpy().astype(state.dtype)
    pos = pos.cpu().numpy()
    return [v for v in views], [tuple(int(x) for x in p) for p in pos]


def generatePerspective(grid_shift, toric_size, state, device="cuda"):
    """src/util.py:46-85: returns [Perspective(array, (m,i,j)), ...]."""
    views, pos = generatePerspectiveOptimized(grid_shift, toric_size, state, device=device)
    return [Perspective(v, p) for v, p in zip(views, pos)]
