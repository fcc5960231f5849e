"""Golden fixtures for the policy-side functions, produced by the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python oracle/gen_golden_policy.py

Writes tests/golden/golden_policy.npz:
  sel_*   src/numba/util_actor.py:_selectActionBatch_prime on Q tables with deliberate ties (greedy) --
          the reference's np.where(q == amax)[0] tie-break -- and src/util_actor.py:selectAction's
          np.argmax form on the same tables
  tr_*    src/util_actor.py:generateTransitionParallel on random (state, next_state, action) triples
  pr_*    src/util_actor.py:computePrioritiesParallel
  pm_*    src/util_learner.py:predictMaxOptimized with an integer-weight linear stand-in model
          (exact in float32 whatever the summation order), terminal states included
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def random_states(rng, n, d, density):
    return (rng.random((n, 2, d, d)) < density).astype(np.int64)


def gen_select(ref, out):
    rng = np.random.default_rng(31)
    prime = ref.numba_util_actor._selectActionBatch_prime
    for d in (5, 7):
        n = 40
        states = random_states(rng, n, d, 0.12)
        states[0, 0, 0, 0] = 1                      # never empty: the reference fails on an empty slice
        for e in range(n):
            if not states[e].any():
                states[e, 1, d - 1, d - 1] = 1
        per, pos, cnt = ref.numba_util_actor.generatePerspectiveBatch(d // 2, d, states)
        counts = np.array(list(cnt), np.int64)
        splice = np.cumsum(counts)
        positions = np.concatenate([np.array(list(p)).reshape(-1, 3) for p in pos])
        # Q values on a coarse grid so that maxima repeat inside many segments
        q = rng.integers(-4, 5, size=(int(splice[-1]), 3)).astype(np.float32) * 0.25
        actions, q_v = prime(q, splice, positions, np.ones(n, dtype=np.bool_))
        # the single-env form's argmax (src/util_actor.py:51) on the same segments
        first = 0
        am = np.zeros((n, 2), np.int64)
        for e in range(n):
            seg = q[first:splice[e]]
            am[e] = np.unravel_index(np.argmax(seg, axis=None), seg.shape)
            first = splice[e]
        out.update({"sel_states%d" % d: states.astype(np.uint8), "sel_q%d" % d: q, "sel_counts%d" % d: counts,
                    "sel_positions%d" % d: positions.astype(np.int64),
                    "sel_actions%d" % d: np.asarray(actions).astype(np.int64),
                    "sel_qv%d" % d: np.asarray(q_v).astype(np.float32), "sel_argmax%d" % d: am})
        print("select d=%d: n=%d sumP=%d" % (d, n, splice[-1]))


def gen_transitions(ref, out):
    rng = np.random.default_rng(32)
    ua = ref.util_actor
    from src.util import action_type                                  # reference dtype of the stored action
    for d in (4, 5, 7, 9):
        n = 64
        g = int(d / 2)
        st = random_states(rng, n, d, 0.15)
        ns = random_states(rng, n, d, 0.15)
        act = np.stack([rng.integers(0, 2, n), rng.integers(0, d, n), rng.integers(0, d, n),
                        rng.integers(1, 4, n)], axis=1).astype(np.int64)
        act[:4, 1:3] = [[0, 0], [d - 1, d - 1], [0, d - 1], [d - 1, 0]]   # wrap-around corners
        rew = rng.integers(-2, 3, n).astype(np.float64)
        term = rng.random(n) < 0.2
        ttype = np.dtype([('perspective', (int, (2, d, d))), ('action', action_type), ('reward', float),
                          ('next_perspective', (int, (2, d, d))), ('terminal', bool)])
        res = ua.generateTransitionParallel(act, rew, st, ns, term, g, ttype)
        out.update({"tr_state%d" % d: st.astype(np.uint8), "tr_next%d" % d: ns.astype(np.uint8),
                    "tr_action%d" % d: act, "tr_out_prev%d" % d: res['perspective'].astype(np.uint8),
                    "tr_out_next%d" % d: res['next_perspective'].astype(np.uint8),
                    "tr_out_pos%d" % d: np.array([list(r['action']['position']) for r in res], np.int64),
                    "tr_out_op%d" % d: np.array([int(r['action']['op']) for r in res], np.int64),
                    "tr_reward%d" % d: res['reward'].astype(np.float64), "tr_terminal%d" % d: res['terminal']})
        print("transitions d=%d: n=%d" % (d, n))


def gen_priorities(ref, out):
    rng = np.random.default_rng(33)
    envs, T = 12, 9
    A = np.zeros((envs, T, 4), np.int64)
    A[..., 3] = rng.integers(1, 4, (envs, T))
    R = rng.integers(-2, 3, (envs, T)).astype(np.float64)
    R[rng.random((envs, T)) < 0.1] = 100.0
    Q = (rng.integers(-400, 400, (envs, T, 3)) / 8.0).astype(np.float64)
    Qns = (rng.integers(-400, 400, (envs, T, 3)) / 8.0).astype(np.float64)
    pr = ref.util_actor.computePrioritiesParallel(A, R, Q, Qns, 0.5)
    out.update({"pr_A": A, "pr_R": R, "pr_Q": Q, "pr_Qns": Qns, "pr_out": np.asarray(pr, np.float64),
                "pr_discount": 0.5})
    print("priorities: %s" % (pr.shape,))


def gen_predict_max(ref, out):
    import torch
    import src.util_learner as ul
    rng = np.random.default_rng(34)
    for d in (5, 7):
        n = 24
        st = random_states(rng, n, d, 0.08)
        st[3] = 0                                    # terminal states (src/util_learner.py:74-76)
        st[11] = 0
        W = rng.integers(-3, 4, size=(2 * d * d, 3)).astype(np.float32)
        bias = np.array([-1.0, 0.5, -2.0], np.float32)

        class Lin(torch.nn.Module):
            def forward(self, x):
                return x.reshape(x.shape[0], -1) @ torch.from_numpy(W) + torch.from_numpy(bias)

        batch_state = [torch.from_numpy(s) for s in st]
        res = ul.predictMaxOptimized(Lin(), batch_state, int(d / 2), d, 'cpu')
        out.update({"pm_states%d" % d: st.astype(np.uint8), "pm_W%d" % d: W, "pm_bias%d" % d: bias,
                    "pm_out%d" % d: res.numpy().astype(np.float32)})
        print("predictMax d=%d: n=%d" % (d, n))


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = ref_loader.load()
    out = {}
    gen_select(ref, out)
    gen_transitions(ref, out)
    gen_priorities(ref, out)
    gen_predict_max(ref, out)
    np.savez_compressed(os.path.join(OUT, "golden_policy.npz"), **out)


if __name__ == "__main__":
    main()
