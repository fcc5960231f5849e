"""CPU oracle for the toric-code environment hot path (numpy restatement).

TEST INFRASTRUCTURE ONLY.  Nothing in the product package
(`toric_rl_decoder_b200/`) may import this module; only `tests/`,
`__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of
`bench.py` use it, and only as the checker / the timed CPU arm.

Parity status: PINNED.  Every function below is checked against the
reference's own Python implementation, imported unmodified in the build
container by `oracle/gen_golden.py`, through the golden vectors committed in
`tests/golden/*.npz` (see `tests/test_oracle_golden.py`).  The reference has no
stored golden vectors for this path other than
`output_speed_test/transitions_0.npy`; that known-answer set is part of the
golden files too (`transitions_d9.npz`).

All `file:line` citations are relative to the reference checkout
(`/root/reference`).  The formulas are restated as closed forms over index
arithmetic (mod d) instead of the reference's roll/where temporaries.

Conventions: `d` lattice size, `g = d // 2` (grid_shift, src/Actor_mp.py:59),
qubit (m, i, j), Pauli code c in {0:I, 1:X, 2:Y, 3:Z}
(gym_ToricCode/gym_ToricCode/envs/ToricCode.py:57).
state[0] = vertex plane V, state[1] = plaquette plane Pl (ToricCode.py:199).
"""
import numpy as np

# ---------------------------------------------------------------------------
# syndrome, step, reward, terminal, logical check
# ---------------------------------------------------------------------------


def x_part(q):
    """x-component of Pauli codes: X (1) and Y (2) flip plaquettes (ToricCode.py:189-191)."""
    return ((q == 1) | (q == 2)).astype(np.int64)


def z_part(q):
    """z-component of Pauli codes: Y (2) and Z (3) flip vertices (ToricCode.py:178-180)."""
    return ((q == 2) | (q == 3)).astype(np.int64)


def syndrome(q):
    """createSyndromOpt (ToricCode.py:168-199).

    V[i,j]  = z0[i,j] ^ z0[i-1,j] ^ z1[i,j] ^ z1[i,j-1]      (:180-186)
    Pl[i,j] = x0[i,j] ^ x0[i,j+1] ^ x1[i,j] ^ x1[i+1,j]      (:191-197)
    `q` is (..., 2, d, d) int; returns (..., 2, d, d) int64 of 0/1.
    """
    q = np.asarray(q)
    x, z = x_part(q), z_part(q)
    z0, z1 = z[..., 0, :, :], z[..., 1, :, :]
    x0, x1 = x[..., 0, :, :], x[..., 1, :, :]
    v = z0 ^ np.roll(z0, 1, axis=-2) ^ z1 ^ np.roll(z1, 1, axis=-1)
    pl = x0 ^ np.roll(x0, -1, axis=-1) ^ x1 ^ np.roll(x1, -1, axis=-2)
    return np.stack((v, pl), axis=-3).astype(np.int64)


def apply_action(q, action):
    """Pauli product of ToricCode.step (ToricCode.py:74-81): rule_table[c, op] == c ^ op."""
    m, i, j, op = (int(a) for a in action)
    q = np.array(q, dtype=np.int64, copy=True)
    q[m, i, j] = q[m, i, j] ^ op
    return q


def is_terminal(state):
    """isTerminalState (ToricCode.py:202-207)."""
    return bool(np.all(np.asarray(state) == 0))


def reward(state, next_state):
    """getReward (ToricCode.py:209-218): 100 if next_state is empty else defect difference."""
    if is_terminal(next_state):
        return 100
    return int(np.sum(state)) - int(np.sum(next_state))


def step(q, state, action):
    """ToricCode.step (ToricCode.py:62-88) on explicit state.

    Returns (q', next_state, reward, terminal).  No bounds check, no auto-reset.
    """
    q2 = apply_action(q, action)
    ns = syndrome(q2)
    return q2, ns, reward(state, ns), is_terminal(ns)


def ground_state(q, previous=True):
    """evalGroundState (ToricCode.py:289-337): sticky False if any of the four
    whole-plane popcounts popcount(x0), popcount(x1), popcount(z0), popcount(z1) is odd."""
    q = np.asarray(q)
    x, z = x_part(q), z_part(q)
    odd = (x[0].sum() % 2 == 1) or (x[1].sum() % 2 == 1) or \
          (z[0].sum() % 2 == 1) or (z[1].sum() % 2 == 1)
    return False if odd else bool(previous)


# ---------------------------------------------------------------------------
# perspectives
# ---------------------------------------------------------------------------


def perspective_masks(state):
    """Which qubits get a perspective (src/util.py:68-69, :77-78; :128-131, :140-143).

    mask[0,i,j] = V[i,j] | V[i+1,j] | Pl[i,j] | Pl[i,j-1]
    mask[1,i,j] = V[i,j] | V[i,j+1] | Pl[i,j] | Pl[i-1,j]
    """
    s = np.asarray(state)
    v, pl = s[0] != 0, s[1] != 0
    m0 = v | np.roll(v, -1, axis=0) | pl | np.roll(pl, 1, axis=1)
    m1 = v | np.roll(v, -1, axis=1) | pl | np.roll(pl, 1, axis=0)
    return np.stack((m0, m1), axis=0)


def perspective_positions(state):
    """(P,3) int64 positions (m,i,j): all m=0 row-major, then all m=1 row-major
    (src/util.py:66-83, :132-148)."""
    return np.argwhere(perspective_masks(state)).astype(np.int64).reshape(-1, 3)


def perspective_view(state, m, i, j):
    """One defect-centred view (src/util.py:70-71 for m=0; :79-81 + rotate_state :87-94 for m=1).

    m=0: out[c,a,b] = S[c,(a+i-g)%d,(b+j-g)%d]
    m=1: out[0,a,b] = V [(b+i-g)%d, ((d-a)%d + j-g)%d]
         out[1,a,b] = Pl[(b+i-g)%d, (d-1-a  + j-g)%d]
    """
    s = np.asarray(state)
    d = s.shape[-1]
    g = d // 2
    a = np.arange(d)[:, None]
    b = np.arange(d)[None, :]
    out = np.empty_like(s)
    if m == 0:
        r = (a + i - g) % d
        c = (b + j - g) % d
        out[0] = s[0][r, c]
        out[1] = s[1][r, c]
    else:
        r = (b + i - g) % d
        out[0] = s[0][r, ((d - a) % d + j - g) % d]
        out[1] = s[1][r, (d - 1 - a + j - g) % d]
    return out


def perspectives(state):
    """generatePerspective / generatePerspectiveOptimized (src/util.py:46-85, :106-150;
    numba twin src/numba/util.py:28-76).  Returns (views (P,2,d,d) same dtype, positions (P,3))."""
    s = np.asarray(state)
    pos = perspective_positions(s)
    views = np.empty((len(pos),) + s.shape, dtype=s.dtype)
    for k, (m, i, j) in enumerate(pos):
        views[k] = perspective_view(s, int(m), int(i), int(j))
    return views, pos


def perspectives_batch(states):
    """generatePerspectiveBatch + the concatenate / float32 cast of selectActionBatch
    (src/numba/util_actor.py:56-67 and :33-39).

    Returns (persp float32 (sumP,2,d,d), positions int64 (sumP,3), counts int64 (N,),
    offsets int64 (N+1,)); cumsum(counts) == offsets[1:] is the reference's splice_idx.
    """
    states = np.asarray(states)
    views, poss, counts = [], [], []
    for s in states:
        v, p = perspectives(s)
        views.append(v)
        poss.append(p)
        counts.append(len(p))
    d = states.shape[-1]
    persp = np.concatenate(views).astype(np.float32) if views else np.zeros((0, 2, d, d), np.float32)
    pos = np.concatenate(poss) if poss else np.zeros((0, 3), np.int64)
    counts = np.asarray(counts, dtype=np.int64)
    offsets = np.concatenate(([0], np.cumsum(counts))).astype(np.int64)
    return persp, pos, counts, offsets


def rotate_state(state):
    """rotate_state (src/util.py:87-94): CCW rot90 of both planes, vertex plane rolled +1 on axis 0.
    out[0,a,b] = V[b, (d-a)%d];  out[1,a,b] = Pl[b, d-1-a]."""
    s = np.asarray(state)
    d = s.shape[-1]
    a = np.arange(d)[:, None]
    b = np.arange(d)[None, :]
    out = np.empty_like(s)
    out[0] = s[0][b, (d - a) % d]
    out[1] = s[1][b, d - 1 - a]
    return out


def shift_state(row, col, previous_state, state, grid_shift):
    """shift_state (src/util.py:97-102): roll both states so (row,col) lands on (g,g)."""
    def sh(s):
        return np.roll(np.roll(s, grid_shift - row, axis=1), grid_shift - col, axis=2)
    return sh(np.asarray(previous_state)), sh(np.asarray(state))


def transition_views(action, state, next_state):
    """The (state, next_state) pair of generateTransitionParallel (src/util_actor.py:223-264):
    both centred on the acted qubit, rotated when it lives in matrix 1; the stored action
    position becomes (m, g, g)."""
    m, i, j, op = (int(a) for a in action)
    d = np.asarray(state).shape[-1]
    g = d // 2
    ps, ns = shift_state(i, j, state, next_state, g)
    if m == 1:
        ps, ns = rotate_state(ps), rotate_state(ns)
    return ps, (m, g, g, op), ns


# --------------------------------------------------------------------------------------------------
# policy-side restatements
# --------------------------------------------------------------------------------------------------
def select_action_prime(q_values_table, offsets, positions, greedy, rand_p=None, rand_a=None):
    """_selectActionBatch_prime (src/numba/util_actor.py:69-107): per env the first maximum of its
    Q-value rows in row-major (perspective, action) order, or the supplied random pick; returns
    (actions int64 (n,4), q_v float (n,3), chosen global row (n,)).  Envs without rows get the identity
    action (0,0,0,0) (the reference would fail on an empty slice; the device path defines this)."""
    q = np.asarray(q_values_table)
    n = len(offsets) - 1
    actions = np.zeros((n, 4), np.int64)
    q_v = np.zeros((n, 3), q.dtype)
    rows = np.full(n, -1, np.int64)
    for e in range(n):
        a0, a1 = int(offsets[e]), int(offsets[e + 1])
        if a1 == a0:
            continue
        seg = q[a0:a1]
        if greedy[e]:
            p, a = np.unravel_index(np.argmax(seg, axis=None), seg.shape)   # src/util_actor.py:51
        else:
            p, a = int(rand_p[e]), int(rand_a[e])
        q_v[e] = seg[p]
        actions[e] = (positions[a0 + p][0], positions[a0 + p][1], positions[a0 + p][2], a + 1)
        rows[e] = a0 + p
    return actions, q_v, rows


def td_priorities(actions, reward, q, q_next, discount):
    """computePrioritiesParallel (src/util_actor.py:268-287) on flat (n,...) arrays, float32 arithmetic."""
    q = np.asarray(q, np.float32)
    qn = np.asarray(q_next, np.float32).max(axis=-1)
    a = np.asarray(actions)[..., 3] - 1
    qv = np.take_along_axis(q, a[..., None], axis=-1)[..., 0]
    return np.abs(np.asarray(reward, np.float32) + np.float32(discount) * qn - qv)


def predict_max(q_values_table, offsets):
    """The reduction of predictMaxOptimized (src/util_learner.py:83-111): terminal states count as one
    (zero) perspective for the padding width and yield 0; shorter segments are zero-padded before argmax."""
    q = np.asarray(q_values_table, np.float32)
    n = len(offsets) - 1
    counts = np.diff(np.asarray(offsets))
    largest = int(max(1, counts.max())) if n else 0
    out = np.zeros(n, np.float32)
    for e in range(n):
        if counts[e] == 0:
            continue
        seg = q[int(offsets[e]):int(offsets[e + 1])].reshape(-1)
        if counts[e] < largest:
            seg = np.concatenate((seg, np.zeros(1, np.float32)))
        out[e] = seg.max()
    return out


def p_error_curriculum(steps, episode, p_cap, p_thresh, start_t, delta_t, final_t, random, all_envs, seed, env_base=0):
    """The reference actor's p_error schedule (src/Actor_mp.py:171-181) in threshold units floor(p * 2^32), as
    csrc/policy_kernels.cuh:p_error_curriculum_kernel applies it after a fused step: the envs that were just reset
    (steps == 0; all_envs: every env) get ceiling <- min(final, ceiling + delta) (:176) and the p of their next
    reset = the ceiling ('linear', :180) or a uniform integer on [min(start, ceiling), ceiling] ('random', :178) drawn
    as lo + mulhi(Philox(env id, episode, 0xA0000000).x, ceiling - lo + 1).  Returns new (p_cap, p_thresh) uint32."""
    from . import c_oracle as C
    cap = np.asarray(p_cap, dtype=np.uint64).copy()
    thr = np.asarray(p_thresh, dtype=np.uint64).copy()
    who = np.arange(len(cap)) if all_envs else np.nonzero(np.asarray(steps) == 0)[0]
    key = (int(seed) & 0xffffffff, int(seed) >> 32)
    for e in who:
        c = min(int(final_t), int(cap[e]) + int(delta_t))
        cap[e] = c
        p = c
        if random:
            lo = min(int(start_t), c)
            g = int(env_base) + int(e)
            u = int(C.philox((g & 0xffffffff, g >> 32, int(episode[e]) & 0xffffffff, 0xA0000000), key)[0])
            p = lo + ((u * (c - lo + 1)) >> 32)
        thr[e] = p
    return cap.astype(np.uint32), thr.astype(np.uint32)
