"""CPU: the numpy oracle's policy-side restatements (oracle/toric_oracle.py: select_action_prime,
transition_views, td_priorities, predict_max) against fixtures produced by the reference's own
_selectActionBatch_prime, generateTransitionParallel, computePrioritiesParallel and predictMaxOptimized
(oracle/gen_golden_policy.py)."""
import numpy as np
import pytest

from oracle import toric_oracle as O
import golden as G


@pytest.fixture(scope="module")
def pol():
    return G.load("golden_policy.npz")


@pytest.mark.parametrize("d", (5, 7))
def test_select_action_prime(pol, d):
    states = pol["sel_states%d" % d].astype(np.int64)
    _, pos, counts, offsets = O.perspectives_batch(states)
    assert (counts == pol["sel_counts%d" % d]).all() and (pos == pol["sel_positions%d" % d]).all()
    q = pol["sel_q%d" % d]
    actions, q_v, rows = O.select_action_prime(q, offsets, pos, np.ones(len(counts), bool))
    assert (actions == pol["sel_actions%d" % d]).all()
    assert (q_v == pol["sel_qv%d" % d]).all()
    am = pol["sel_argmax%d" % d]
    assert (rows - offsets[:-1] == am[:, 0]).all() and (actions[:, 3] - 1 == am[:, 1]).all()


@pytest.mark.parametrize("d", (4, 5, 7, 9))
def test_transition_views(pol, d):
    st, ns, act = pol["tr_state%d" % d], pol["tr_next%d" % d], pol["tr_action%d" % d]
    for e in range(len(act)):
        ps, a, nx = O.transition_views(act[e], st[e], ns[e])
        assert (ps == pol["tr_out_prev%d" % d][e]).all() and (nx == pol["tr_out_next%d" % d][e]).all()
        assert tuple(a[:3]) == tuple(pol["tr_out_pos%d" % d][e]) and a[3] == pol["tr_out_op%d" % d][e]


def test_td_priorities(pol):
    out = O.td_priorities(pol["pr_A"], pol["pr_R"], pol["pr_Q"], pol["pr_Qns"], float(pol["pr_discount"]))
    assert np.allclose(out, pol["pr_out"], rtol=1e-6, atol=1e-6)   # values are exact eighths: equal in f32


@pytest.mark.parametrize("d", (5, 7))
def test_predict_max(pol, d):
    st = pol["pm_states%d" % d].astype(np.int64)
    persp, _, counts, offsets = O.perspectives_batch(st)
    q = persp.reshape(len(persp), -1) @ pol["pm_W%d" % d] + pol["pm_bias%d" % d]
    out = O.predict_max(q.astype(np.float32), offsets)
    assert (out == pol["pm_out%d" % d]).all()
