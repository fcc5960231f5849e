"""Pins the numpy oracle (oracle/toric_oracle.py) to the reference through the golden fixtures."""
import numpy as np
import pytest

import golden
from oracle import toric_oracle as O


@pytest.mark.parametrize("d", golden.ALL_D)
def test_syndrome_and_logical(d):
    g = golden.per_d(d)
    assert (O.syndrome(g["qubits"]) == g["state"]).all()
    for q, gs in zip(g["qubits"], g["gs"]):
        assert O.ground_state(q) == bool(gs)


@pytest.mark.parametrize("d", golden.ALL_D)
def test_perspectives(d):
    g = golden.per_d(d)
    for states, counts, pos, views in ((g["state"], g["counts"], g["positions"], g["views"]),
                                       (g["step_state"][:, -1], g["counts_last"], g["positions_last"],
                                        g["views_last"])):
        persp, p, c, off = O.perspectives_batch(states)
        assert persp.dtype == np.float32
        assert (c == counts).all()
        assert (p == pos).all()
        assert (persp == views.astype(np.float32)).all()
        assert off[-1] == len(views)


@pytest.mark.parametrize("d", golden.ALL_D)
def test_step_sequences(d):
    g = golden.per_d(d)
    for e in range(len(g["qubits"])):
        q = g["qubits"][e].astype(np.int64)
        s = g["state"][e].astype(np.int64)
        gs = True
        for t in range(g["actions"].shape[1]):
            a = g["actions"][e, t]
            q, ns, r, term = O.step(q, s, a)
            assert (q == g["step_qubits"][e, t]).all()
            assert (ns == g["step_state"][e, t]).all()
            assert r == g["step_reward"][e, t]
            assert term == bool(g["step_terminal"][e, t])
            gs = O.ground_state(q, gs)
            assert gs == bool(g["step_gs"][e, t])
            ps, act, nv = O.transition_views(a, s, ns)
            assert (ps == g["tr_prev"][e, t]).all() and (nv == g["tr_next"][e, t]).all()
            s = ns


@pytest.mark.parametrize("d", (3, 5))
def test_exhaustive(d):
    g = golden.exhaustive(d)
    assert (O.syndrome(g["qubits"]) == g["state"]).all()
    persp, p, c, off = O.perspectives_batch(g["state"])
    assert (c == g["counts"]).all() and (p == g["positions"]).all()
    assert (persp == g["views"]).all()
    for q, gs in zip(g["qubits"], g["gs"]):
        assert O.ground_state(q) == bool(gs)


def test_batch_layout():
    g = golden.batch_d7()
    q = g["qubits0"].astype(np.int64)
    s = g["states0"].astype(np.int64)
    assert (O.syndrome(q) == s).all()
    for t in range(g["actions"].shape[0]):
        for e in range(len(q)):
            q[e], ns, r, term = O.step(q[e], s[e], g["actions"][t, e])
            assert (ns == g["states"][t, e]).all()
            assert r == g["rewards"][t, e] and term == g["terminals"][t, e]
            s[e] = ns
    persp, pos, counts, off = O.perspectives_batch(s)
    assert (np.cumsum(counts) == g["splice_idx"]).all()
    assert (pos == g["positions"]).all()
    assert persp.dtype == np.float32 and (persp == g["views"]).all()


def test_known_answer_transitions():
    """The reference's stored transitions (output_speed_test/transitions_0.npy, d=9).  The stored
    state is the action-centred (and, for matrix 1, rotated) view, so applying the incremental rule
    of a matrix-0 qubit at (g,g) to it gives the stored next state.  The file was written by an older
    actor whose records are not all single-step (several entries differ in unrelated cells, or not at
    all); the entries that ARE single-step must match our rule exactly, entry 0 (the one quoted in
    SURVEY.md section 4) among them."""
    g = golden.load("transitions_d9.npz")
    d = g["state"].shape[-1]
    c = d // 2
    consistent = []
    for k, (s, pos, op, r, ns, term) in enumerate(zip(g["state"], g["position"], g["op"], g["reward"],
                                                      g["next_state"], g["terminal"])):
        assert tuple(pos[1:]) == (c, c)
        exp = s.astype(np.int64).copy()
        if op in (2, 3):      # z-part flips the two vertices of a matrix-0 qubit
            exp[0, c, c] ^= 1
            exp[0, (c + 1) % d, c] ^= 1
        if op in (1, 2):      # x-part flips its two plaquettes
            exp[1, c, c] ^= 1
            exp[1, c, (c - 1) % d] ^= 1
        if (exp == ns).all():
            consistent.append(k)
            assert O.reward(s, ns) == r and O.is_terminal(ns) == bool(term)
    assert 0 in consistent and len(consistent) >= 5
