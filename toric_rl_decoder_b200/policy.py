"""Policy-side mirror of the reference's callers of the env path, on the GPU.

Same names and argument meaning as the reference:
  selectActionBatch / selectAction          src/numba/util_actor.py:12-53, src/util_actor.py:7-66
  generateTransitionParallel                src/util_actor.py:223-264
  computePrioritiesParallel                 src/util_actor.py:268-287
  predictMaxOptimized                       src/util_learner.py:48-111
The Q-network forward stays in PyTorch (it is not the hot path); everything around it -- perspectives,
segmented argmax / epsilon-greedy, roll-and-rotate of the transition pair, TD priorities -- runs in
libtoric_b200.so.  The device-resident forms (`select_action_device`, `transitions_device`, ...) never
synchronise except to learn sumP for slicing the network input.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .engine import _ptr, _stream, p_to_thresh

DTYPE_F32, DTYPE_U8 = 0, 1


def select_action_device(q_values, pos, offsets, n, *, epsilon=0.0, eps_thresh=None, greedy=None, episode=None,
                         steps=None, seed=0, env_base=0, actions=None, q_sel=None, sel_idx=None, was_greedy=None):
    """tc_select_action on device tensors.  Returns (actions int32 (n,4), q_sel f32 (n,3))."""
    lib = _lib.load()
    dev = q_values.device
    if q_values.dtype != torch.float32 or not q_values.is_contiguous():
        q_values = q_values.float().contiguous()
    if actions is None:
        actions = torch.empty((n, 4), dtype=torch.int32, device=dev)
    if q_sel is None:
        q_sel = torch.empty((n, 3), dtype=torch.float32, device=dev)
    if greedy is not None:
        greedy = torch.as_tensor(greedy).to(device=dev, dtype=torch.uint8).contiguous()
    eps_scalar = int(p_to_thresh(float(epsilon)))
    _lib.check(lib.tc_select_action(_ptr(q_values), _ptr(pos), _ptr(offsets), _ptr(episode), _ptr(steps),
                                    _ptr(eps_thresh), ctypes.c_uint32(eps_scalar), _ptr(greedy), int(seed),
                                    int(env_base), _ptr(actions), _ptr(q_sel), _ptr(sel_idx), _ptr(was_greedy),
                                    int(n), _stream()), "tc_select_action")
    return actions, q_sel


def _forward(model, persp, chunk=1 << 16):
    """Q-network forward over the perspective tensor in chunks (PyTorch; not the hot path)."""
    if persp.shape[0] == 0:
        return torch.zeros((0, 3), dtype=torch.float32, device=persp.device)
    with torch.no_grad():
        n = persp.shape[0]
        if n <= 4096 and n % 64 and getattr(model, "pad_small_batches", False):
            # single-env / small-batch use: every step has a different number of perspectives, and cuDNN spends
            # milliseconds choosing kernels for each new batch size; 64 sizes instead of 4096
            padded = torch.zeros((n + 63) // 64 * 64, *persp.shape[1:], dtype=persp.dtype, device=persp.device)
            padded[:n] = persp
            return model(padded)[:n].float().contiguous()
        if persp.shape[0] <= chunk:
            return model(persp).float().contiguous()
        return torch.cat([model(persp[i:i + chunk]).float() for i in range(0, persp.shape[0], chunk)]).contiguous()


def selectActionBatch(number_of_actions, epsilon, grid_shift, toric_size, state, model, device="cuda", seed=0,
                      step=0):
    """selectActionBatch (src/numba/util_actor.py:12-53): state (N,2,d,d) -> (actions int64 (N,4),
    q_values (N,3)) as numpy arrays, like the reference."""
    from .perspectives import generatePerspectiveBatch
    model.eval()
    persp, pos3, counts, offsets = generatePerspectiveBatch(grid_shift, toric_size, state, device=device,
                                                            raw_positions=True)
    n = counts.shape[0]
    q = _forward(model, persp)
    steps = torch.full((n,), int(step), dtype=torch.int32, device=q.device)
    actions, q_sel = select_action_device(q, pos3, offsets, n, epsilon=epsilon, steps=steps, seed=seed)
    return actions.cpu().numpy().astype(np.int64), q_sel.cpu().numpy()


def selectAction(number_of_actions, epsilon, grid_shift, toric_size, state, model, device="cuda", seed=0, step=0):
    """selectAction (src/util_actor.py:7-66): single state -> (action list, q_values (3,))."""
    a, q = selectActionBatch(number_of_actions, epsilon, grid_shift, toric_size, np.asarray(state)[None], model,
                             device=device, seed=seed, step=step)
    return [int(x) for x in a[0]], q[0]


def transitions_device(syn_prev, syn_next, actions, d, out_dtype=torch.float32):
    """tc_transitions: packed syndromes int32 (n,2,d) x2 + actions int32 (n,4) -> (prev (n,2,d,d),
    next (n,2,d,d), action_out int32 (n,4) = (m,g,g,op))."""
    lib = _lib.load()
    n = actions.shape[0]
    dev = actions.device
    code = {torch.float32: DTYPE_F32, torch.uint8: DTYPE_U8}[out_dtype]
    op = torch.empty((n, 2, d, d), dtype=out_dtype, device=dev) if syn_prev is not None else None
    on = torch.empty((n, 2, d, d), dtype=out_dtype, device=dev) if syn_next is not None else None
    ao = torch.empty((n, 4), dtype=torch.int32, device=dev)
    _lib.check(lib.tc_transitions(_ptr(syn_prev), _ptr(syn_next), _ptr(actions), _ptr(op), _ptr(on), _ptr(ao),
                                  code, int(d), int(n), _stream()), "tc_transitions")
    return op, on, ao


def _pack(states, d, dev):
    lib = _lib.load()
    st = torch.as_tensor(np.ascontiguousarray(states) if isinstance(states, np.ndarray) else states)
    from .perspectives import _as_defect_bytes
    u8 = _as_defect_bytes(st, torch.device(dev))
    syn = torch.empty((u8.shape[0], 2, d), dtype=torch.int32, device=dev)
    _lib.check(lib.tc_pack_state(_ptr(u8), _ptr(syn), d, u8.shape[0], _stream()), "tc_pack_state")
    return syn


def rotate_state(state, device="cuda"):
    """rotate_state (src/util.py:87-94; numba twin src/numba/util.py:17-25): counter-clockwise rot90 of both
    planes, vertex plane then rolled by +1 along axis 0.  (2,d,d) in, int64 (2,d,d) out.  Computed by the
    transition kernel: acting on qubit (1,g,g) needs no shift, only the matrix-1 rotation."""
    st = np.asarray(state)
    d = int(st.shape[-1])
    dev = torch.device(device)
    syn = _pack(st[None], d, dev)
    act = torch.tensor([[1, d // 2, d // 2, 1]], dtype=torch.int32, device=dev)
    _, on, _ = transitions_device(None, syn, act, d, out_dtype=torch.uint8)
    return on[0].cpu().numpy().astype(np.int64)


def shift_state(row, col, previous_state, state, grid_shift, device="cuda"):
    """shift_state (src/util.py:97-102; numba twin src/numba/util.py:8-14): both states rolled by
    (grid_shift-row, grid_shift-col) so that (row,col) lands on (grid_shift,grid_shift)."""
    prev, st = np.asarray(previous_state), np.asarray(state)
    d = int(st.shape[-1])
    if int(grid_shift) != d // 2:
        raise ValueError("grid_shift must be toric_size // 2")
    dev = torch.device(device)
    act = torch.tensor([[0, int(row), int(col), 1]], dtype=torch.int32, device=dev)
    op, on, _ = transitions_device(_pack(prev[None], d, dev), _pack(st[None], d, dev), act, d,
                                   out_dtype=torch.uint8)
    return op[0].cpu().numpy().astype(np.int64), on[0].cpu().numpy().astype(np.int64)


def generateTransitionParallel(action, reward, state, next_state, terminal_state, grid_shift, trans_type=None,
                               device="cuda"):
    """generateTransitionParallel (src/util_actor.py:223-264).  With trans_type (the reference's structured
    dtype, src/Actor_mp.py:52-56) returns the same numpy record array; without it returns the device tensors
    (prev u8, action int32 (m,g,g,op), reward, next u8, terminal)."""
    next_state = np.asarray(next_state) if not torch.is_tensor(next_state) else next_state
    d = int(next_state.shape[-1])
    if int(grid_shift) != d // 2:
        raise ValueError("grid_shift must be toric_size // 2")
    dev = torch.device(device)
    sp, sn = _pack(state, d, dev), _pack(next_state, d, dev)
    act = action if torch.is_tensor(action) else torch.as_tensor(np.asarray(action, dtype=np.int64))
    act = act.to(device=dev, dtype=torch.int32).contiguous()
    op, on, ao = transitions_device(sp, sn, act, d, out_dtype=torch.uint8)
    if trans_type is None:
        return op, ao, reward, on, terminal_state
    n = act.shape[0]
    res = np.empty(n, dtype=trans_type)
    names = trans_type.names
    res[names[0]] = op.cpu().numpy()
    aon = ao.cpu().numpy()
    res[names[1]]['position'] = aon[:, :3]
    res[names[1]]['op'] = aon[:, 3]
    res[names[2]] = reward.cpu().numpy() if torch.is_tensor(reward) else np.asarray(reward)
    res[names[3]] = on.cpu().numpy()
    res[names[4]] = terminal_state.cpu().numpy() if torch.is_tensor(terminal_state) else np.asarray(terminal_state)
    return res


def computePrioritiesParallel(A, R, Q, Qns, discount, device="cuda"):
    """computePrioritiesParallel (src/util_actor.py:268-287): A (envs,T,4), R (envs,T), Q/Qns (envs,T,3)
    -> |R + discount * max Qns - Q[action]| (envs,T), float32 arithmetic on the device."""
    lib = _lib.load()
    dev = torch.device(device)
    shape = np.asarray(R).shape if not torch.is_tensor(R) else tuple(R.shape)
    a = torch.as_tensor(np.asarray(A) if not torch.is_tensor(A) else A).to(device=dev, dtype=torch.int32).reshape(-1, 4).contiguous()
    r = torch.as_tensor(np.asarray(R) if not torch.is_tensor(R) else R).to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
    q = torch.as_tensor(np.asarray(Q) if not torch.is_tensor(Q) else Q).to(device=dev, dtype=torch.float32).reshape(-1, 3).contiguous()
    qn = torch.as_tensor(np.asarray(Qns) if not torch.is_tensor(Qns) else Qns).to(device=dev, dtype=torch.float32).reshape(-1, 3).contiguous()
    out = torch.empty(r.shape[0], dtype=torch.float32, device=dev)
    _lib.check(lib.tc_td_priorities(_ptr(a), _ptr(r), _ptr(q), _ptr(qn), ctypes.c_float(float(discount)), _ptr(out),
                                    r.shape[0], _stream()), "tc_td_priorities")
    return out.reshape(shape)


def predictMaxOptimized(model, batch_state, grid_shift, system_size, device="cuda"):
    """predictMaxOptimized (src/util_learner.py:48-111): max Q-value of every state of the batch (float32
    tensor on the CPU like the reference), including the reference's zero-padding of shorter segments."""
    from .perspectives import generatePerspectiveBatch
    lib = _lib.load()
    d = int(system_size)
    if isinstance(batch_state, (list, tuple)):
        batch_state = torch.stack([torch.as_tensor(s) for s in batch_state])
    model.eval()
    persp, _, counts, offsets = generatePerspectiveBatch(int(d / 2), d, batch_state, device=device,
                                                         raw_positions=True)
    n = counts.shape[0]
    q = _forward(model, persp)
    pad = counts.max().to(torch.int32).reshape(1) if n else torch.zeros(1, dtype=torch.int32, device=q.device)
    out = torch.empty(n, dtype=torch.float32, device=q.device)
    qq = q if q.shape[0] else torch.zeros((1, 3), dtype=torch.float32, device=out.device)
    _lib.check(lib.tc_segmented_max(_ptr(qq), _ptr(offsets), 0, _ptr(pad), _ptr(out), n, _stream()),
               "tc_segmented_max")
    return out.cpu()


# ---- the single-env / "Parallel" spellings of src/util_actor.py, on the same device entry points ----------------

def selectActionParallel(number_of_actions, epsilon, grid_shift, toric_size, state, model, device="cuda", seed=0,
                         step=0):
    """selectActionParallel (src/util_actor.py:132-175): the numpy-era spelling of selectActionBatch, same
    arguments and returns ((N,4) actions, (N,3) Q-value rows of the chosen perspectives)."""
    return selectActionBatch(number_of_actions, epsilon, grid_shift, toric_size, state, model, device=device,
                             seed=seed, step=step)


def generatePerspectiveParallel(grid_shift, toric_size, states, device="cuda"):
    """generatePerspectiveParallel (src/util_actor.py:178-186): per env the list of (2,d,d) perspectives and the
    list of (m,i,j) positions -- one batched tc_perspectives call, split per env on the host.  (The reference
    wraps the two ragged lists in np.array, which numpy >= 1.24 refuses; lists are returned.)"""
    from .perspectives import generatePerspectiveBatch
    states = np.asarray(states)
    persp, pos, counts, offsets = generatePerspectiveBatch(grid_shift, toric_size, states, device=device)
    persp = persp.cpu().numpy().astype(states.dtype)
    pos, off = pos.cpu().numpy(), offsets.cpu().numpy()
    per_result = [[persp[k] for k in range(off[e], off[e + 1])] for e in range(len(states))]
    pos_result = [[tuple(int(x) for x in pos[k]) for k in range(off[e], off[e + 1])] for e in range(len(states))]
    return per_result, pos_result


def generateTransition(action, reward, grid_shift, previous_state, state, terminal_state, device="cuda"):
    """generateTransition (src/util_actor.py:67-103): one Transition(previous_perspective, Action((m,g,g), op),
    reward, perspective, terminal) -- tc_transitions on a batch of one."""
    from .perspectives import Action, Transition
    state = np.asarray(state)
    d = int(state.shape[-1])
    if int(grid_shift) != d // 2:
        raise ValueError("grid_shift must be toric_size // 2")
    dev = torch.device(device)
    act = torch.as_tensor(np.asarray(action, dtype=np.int64).reshape(1, 4)).to(device=dev, dtype=torch.int32)
    op, on, ao = transitions_device(_pack(np.asarray(previous_state)[None], d, dev), _pack(state[None], d, dev),
                                    act.contiguous(), d, out_dtype=torch.uint8)
    a = ao[0].cpu().numpy()
    return Transition(op[0].cpu().numpy().astype(state.dtype), Action((int(a[0]), int(a[1]), int(a[2])), int(a[3])),
                      reward, on[0].cpu().numpy().astype(state.dtype), terminal_state)


def computePriorities(local_buffer_trans, local_buffer_qs, local_buffer_qs_ns, discount_factor, device="cuda"):
    """computePriorities (src/util_actor.py:105-124): |r + discount * max Q(s') - Q(s)[a]| for a list of Transition
    records and their Q-value rows (tc_td_priorities; float32 arithmetic on the device)."""
    n = len(local_buffer_trans)
    A = np.zeros((1, n, 4), dtype=np.int64)
    R = np.zeros((1, n), dtype=np.float64)
    for i, t in enumerate(local_buffer_trans):
        A[0, i, :3] = t.action.position
        A[0, i, 3] = t.action.action
        R[0, i] = t.reward
    Q = np.asarray(local_buffer_qs, dtype=np.float32).reshape(1, n, 3)
    Qn = np.asarray(local_buffer_qs_ns, dtype=np.float32).reshape(1, n, 3)
    return computePrioritiesParallel(A, R, Q, Qn, discount_factor, device=device)[0].cpu().numpy()


def calculateEpsilon(e, alpha, num):
    """calculateEpsilon (src/util_actor.py:294-312): the per-actor exploration schedule e_i = e^(1 + i/(n-1) * alpha)
    (host-side configuration arithmetic, one value per env; feeds tc_select_action's eps_thresh)."""
    i = np.arange(int(num), dtype=np.float64)
    return np.power(float(e), 1.0 + (i / max(1, int(num) - 1)) * float(alpha))
