"""Generate the golden fixtures in tests/golden/ from the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python oracle/gen_golden.py

Every output array below is produced by the reference's own code
(gym_ToricCode ToricCode, src/EnvSet.py, src/util.py, src/numba/util*.py),
imported through oracle/ref_loader.py.  Error masks are injected the way the
reference's own scripts do it, by attribute assignment
(results/small_p_error_test.py:119-120, results/start_from_state.py:36-38).

Files (all small, compressed):
  golden_d{3,4,5,6,7,9,15,25}.npz  random masks from the reference sampler, their syndromes,
                                   perspectives, logical flags, and random action sequences
  golden_exhaustive.npz            all single-qubit masks at d=3,5; all two-qubit masks at d=3;
                                   a sample of two-qubit masks at d=5 (wrap-around included)
  golden_batch_d7.npz              EnvSet.step + numba generatePerspectiveBatch + the
                                   concatenate/float32 layout of selectActionBatch
  transitions_d9.npz               the reference's stored known-answer transitions
                                   (output_speed_test/transitions_0.npy)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
P_LIST = (0.005, 0.05, 0.1, 0.2)
T_STEPS = 12


def make_env(ref, d, p=0.1, min_errors=0):
    return ref.gym.make("toric-code-v0", config={"size": d, "min_qbit_errors": min_errors, "p_error": p})


def inject(env, q):
    env.qubit_matrix = np.array(q, dtype=int)
    env.state = env.createSyndromOpt(env.qubit_matrix)
    env.next_state = env.state
    env.ground_state = True


def persp_pack(ref, d, states):
    """counts, positions (sumP,3) u8, packed bits of the (sumP,2,d,d) views (numpy reference path)."""
    g = d // 2
    counts, poss, views = [], [], []
    for s in states:
        v, p = ref.util.generatePerspectiveOptimized(g, d, np.array(s, dtype=int))
        # the loop form must agree with the optimized form (tests/time_test_genPersp.py:34-39 pattern)
        v2 = ref.util.generatePerspective(g, d, np.array(s, dtype=int))
        assert len(v2) == len(v)
        for a, b, c in zip(v2, v, p):
            assert (a.perspective == b).all() and tuple(a.position) == tuple(c)
        counts.append(len(v))
        if len(v):
            poss.append(np.array(p, dtype=np.uint8).reshape(-1, 3))
            views.append(np.array(v, dtype=np.uint8).reshape(-1, 2, d, d))
    pos = np.concatenate(poss) if poss else np.zeros((0, 3), np.uint8)
    vw = np.concatenate(views) if views else np.zeros((0, 2, d, d), np.uint8)
    return np.array(counts, np.int32), pos, np.packbits(vw.reshape(-1)), vw.shape[0]


def random_action(rng, d, state, ref):
    """Mostly a qubit from the perspective list (what an agent can pick), sometimes any qubit,
    sometimes forced onto the wrap-around border."""
    r = rng.random()
    op = int(rng.integers(1, 4))
    if r < 0.7:
        _, pos = ref.util.generatePerspectiveOptimized(d // 2, d, state)
        if len(pos):
            m, i, j = pos[int(rng.integers(len(pos)))]
            return [int(m), int(i), int(j), op]
    if r < 0.85:
        edge = [0, d - 1]
        return [int(rng.integers(2)), edge[int(rng.integers(2))], edge[int(rng.integers(2))], op]
    return [int(rng.integers(2)), int(rng.integers(d)), int(rng.integers(d)), op]


def gen_d(ref, d, k, seed):
    rng = np.random.default_rng(seed)
    np.random.seed(seed)
    env = make_env(ref, d)
    g = d // 2
    Q = np.zeros((k, 2, d, d), np.uint8)
    S = np.zeros((k, 2, d, d), np.uint8)
    GS = np.zeros(k, bool)
    A = np.zeros((k, T_STEPS, 4), np.uint8)
    SQ = np.zeros((k, T_STEPS, 2, d, d), np.uint8)
    SS = np.zeros((k, T_STEPS, 2, d, d), np.uint8)
    SR = np.zeros((k, T_STEPS), np.float64)
    ST = np.zeros((k, T_STEPS), bool)
    SG = np.zeros((k, T_STEPS), bool)
    TP = np.zeros((k, T_STEPS, 2, d, d), np.uint8)
    TN = np.zeros((k, T_STEPS, 2, d, d), np.uint8)
    for e in range(k):
        p = P_LIST[e % len(P_LIST)]
        if e % 7 == 6:  # fixed-weight sampler of the reference (ToricCode.py:153-166)
            env.min_qbit_errors = 1 + e % 5
            s = env.reset(p_error=p)
            env.min_qbit_errors = 0
        else:
            s = env.reset(p_error=p)
        Q[e] = env.qubit_matrix
        S[e] = s
        assert (env.createSyndromOpt(env.qubit_matrix) == s).all()
        GS[e] = env.evalGroundState()
        inject(env, Q[e])  # evalGroundState is sticky: start the sequence from a clean flag
        for t in range(T_STEPS):
            prev = np.array(env.state)
            left = np.argwhere(env.qubit_matrix != 0)
            if e % 4 == 1 and len(left):
                # "undo" policy: re-apply each error's own Pauli -> reaches the terminal state
                # (reward 100), after which the env keeps being stepped (legal in the reference)
                m_, i_, j_ = left[0]
                a = [int(m_), int(i_), int(j_), int(env.qubit_matrix[m_, i_, j_])]
            else:
                a = random_action(rng, d, prev, ref)
            ns, r, term, _ = env.step(a)
            A[e, t] = a
            SQ[e, t] = env.qubit_matrix
            SS[e, t] = ns
            SR[e, t] = r
            ST[e, t] = term
            SG[e, t] = env.evalGroundState()
            tr = ref.util_actor.generateTransition(a, r, g, prev, np.array(ns), term)
            TP[e, t] = tr.state
            TN[e, t] = tr.next_state
            assert tuple(tr.action.position) == (a[0], g, g) and tr.action.action == a[3]
    counts, pos, bits, total = persp_pack(ref, d, S)
    # perspectives of the last post-step state too (denser syndromes than fresh resets)
    counts2, pos2, bits2, total2 = persp_pack(ref, d, SS[:, -1])
    np.savez_compressed(
        os.path.join(OUT, "golden_d%d.npz" % d), d=d, qubits=Q, state=S, gs=GS,
        counts=counts, positions=pos, persp_bits=bits, persp_total=total,
        actions=A, step_qubits=SQ, step_state=SS, step_reward=SR, step_terminal=ST, step_gs=SG,
        tr_prev=TP, tr_next=TN,
        counts_last=counts2, positions_last=pos2, persp_bits_last=bits2, persp_total_last=total2)
    print("d=%d: %d envs, sumP=%d (+%d), terminals=%d" % (d, k, total, total2, ST.sum()))


def gen_exhaustive(ref):
    out = {}
    rng = np.random.default_rng(1234)
    for d, pair_limit in ((3, None), (5, 400)):
        env = make_env(ref, d)
        n = 2 * d * d
        masks = []
        for q in range(n):
            for c in (1, 2, 3):
                m = np.zeros(n, np.uint8)
                m[q] = c
                masks.append(m)
        pairs = [(a, b) for a in range(n) for b in range(a + 1, n)]
        if pair_limit is not None:
            sel = rng.choice(len(pairs), size=pair_limit, replace=False)
            pairs = [pairs[i] for i in sorted(sel)]
        for (a, b) in pairs:
            for ca in (1, 2, 3):
                for cb in (1, 2, 3):
                    if pair_limit is not None and rng.random() > 0.34:
                        continue
                    m = np.zeros(n, np.uint8)
                    m[a], m[b] = ca, cb
                    masks.append(m)
        Q = np.array(masks, np.uint8).reshape(-1, 2, d, d)
        S = np.zeros_like(Q)
        GS = np.zeros(len(Q), bool)
        for e in range(len(Q)):
            inject(env, Q[e])
            S[e] = env.state
            GS[e] = env.evalGroundState()
        counts, pos, bits, total = persp_pack(ref, d, S)
        out.update({"q%d" % d: Q, "s%d" % d: S, "gs%d" % d: GS, "counts%d" % d: counts,
                    "pos%d" % d: pos, "bits%d" % d: bits, "total%d" % d: total})
        print("exhaustive d=%d: %d masks, sumP=%d" % (d, len(Q), total))
    np.savez_compressed(os.path.join(OUT, "golden_exhaustive.npz"), **out)


def gen_batch(ref, d=7, n=48, seed=7):
    """EnvSet façade + numba batch generator + the select-action input layout."""
    np.random.seed(seed)
    rng = np.random.default_rng(seed)
    env = make_env(ref, d)
    es = ref.EnvSet(env, n)
    p_errors = rng.choice(P_LIST, size=n)
    states0 = np.array(es.resetAll(p_errors=p_errors))
    Q0 = np.array([e.qubit_matrix for e in es.envs], np.uint8)
    T = 6
    A = np.zeros((T, n, 4), np.int64)
    S = np.zeros((T, n, 2, d, d), np.uint8)
    R = np.zeros((T, n), np.float64)
    TM = np.zeros((T, n), bool)
    cur = states0
    for t in range(T):
        A[t] = [random_action(rng, d, np.array(cur[i]), ref) for i in range(n)]
        st, rw, tm, _ = es.step(A[t])
        assert st.dtype == np.int64 and rw.dtype == np.float64 and tm.dtype == np.bool_
        S[t], R[t], TM[t] = st, rw, tm
        cur = st
    # numba generatePerspectiveBatch (src/numba/util_actor.py:56-67) and the layout lines :33-39
    per, pos, cnt = ref.numba_util_actor.generatePerspectiveBatch(d // 2, d, np.array(cur, dtype=np.int64))
    keep = [i for i in range(n) if cnt[i] > 0]
    splice_idx = np.cumsum(np.array(list(cnt)))
    positions = np.concatenate([np.array(list(pos[i])).reshape(-1, 3) for i in keep])
    import torch
    perspectives = np.concatenate([np.array(list(per[i])) for i in keep])
    tensor = torch.from_numpy(perspectives).type('torch.Tensor')
    assert tensor.dtype == torch.float32
    np.savez_compressed(
        os.path.join(OUT, "golden_batch_d%d.npz" % d), d=d, p_errors=p_errors, qubits0=Q0,
        states0=states0.astype(np.uint8), actions=A, states=S, rewards=R, terminals=TM,
        splice_idx=splice_idx.astype(np.int64), positions=positions.astype(np.int64),
        persp_bits=np.packbits(tensor.numpy().astype(np.uint8).reshape(-1)),
        persp_total=tensor.shape[0])
    print("batch d=%d: n=%d sumP=%d" % (d, n, tensor.shape[0]))


def gen_transitions(ref):
    path = os.path.join(ref_loader.REF_ROOT, "output_speed_test", "transitions_0.npy")
    raw = np.load(path, allow_pickle=True)
    st = np.array([r[0] for r in raw], np.uint8)
    pos = np.array([r[1].position for r in raw], np.int64)
    op = np.array([int(r[1].action) for r in raw], np.int64)
    rew = np.array([float(r[2]) for r in raw], np.float64)
    ns = np.array([r[3] for r in raw], np.uint8)
    term = np.array([bool(r[4]) for r in raw], bool)
    np.savez_compressed(os.path.join(OUT, "transitions_d9.npz"), state=st, position=pos, op=op,
                        reward=rew, next_state=ns, terminal=term)
    print("transitions: %d entries, d=%d" % (len(raw), st.shape[-1]))


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = ref_loader.load()
    for d, k in ((3, 64), (4, 16), (5, 64), (6, 16), (7, 64), (9, 32), (15, 8), (25, 4)):
        gen_d(ref, d, k, seed=1000 + d)
    gen_exhaustive(ref)
    gen_batch(ref)
    gen_transitions(ref)


if __name__ == "__main__":
    main()
