"""Pins the C oracle (oracle/toric_oracle.c) to the reference through the golden fixtures, and its
Philox / sampler restatement to published known-answer vectors and to the numpy oracle."""
import numpy as np
import pytest

import golden
from oracle import c_oracle as C
from oracle import toric_oracle as O


@pytest.mark.parametrize("d", golden.ALL_D)
def test_c_syndrome_perspectives_logical(d):
    g = golden.per_d(d)
    assert (C.syndrome(g["qubits"]) == g["state"]).all()
    off = 0
    for e, s in enumerate(g["state"]):
        out, pos = C.perspectives(s)
        n = int(g["counts"][e])
        assert len(out) == n
        assert (out == g["views"][off:off + n]).all() and (pos == g["positions"][off:off + n]).all()
        off += n
        assert C.ground_state(g["qubits"][e]) == bool(g["gs"][e])


@pytest.mark.parametrize("d", golden.ALL_D)
def test_c_step_sequences(d):
    g = golden.per_d(d)
    for e in range(len(g["qubits"])):
        q = g["qubits"][e].copy()
        s = g["state"][e].copy()
        gs = True
        for t in range(g["actions"].shape[1]):
            r, term = C.step(q, s, g["actions"][e, t])
            assert (q == g["step_qubits"][e, t]).all() and (s == g["step_state"][e, t]).all()
            assert r == g["step_reward"][e, t] and term == bool(g["step_terminal"][e, t])
            gs = C.ground_state(q, gs)
            assert gs == bool(g["step_gs"][e, t])


def test_philox_known_answers():
    """Random123 kat_vectors for philox4x32-10."""
    assert [int(x) for x in C.philox([0, 0, 0, 0], [0, 0])] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert [int(x) for x in C.philox([0xffffffff] * 4, [0xffffffff] * 2)] == \
        [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert [int(x) for x in C.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0])] == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


@pytest.mark.parametrize("mode,n_fixed", ((0, 0), (1, 3), (2, 4)))
def test_sampler_properties(mode, n_fixed):
    d = 7
    th = int(0.05 * 2 ** 32)
    weights = []
    for env in range(200):
        q, s, att = C.reset(d, env, 3, th, seed=99, mode=mode, n_fixed=n_fixed)
        assert (O.syndrome(q) == s).all() and s.any() and att >= 1
        weights.append(int((q != 0).sum()))
        q2, s2, _ = C.reset(d, env, 3, th, seed=99, mode=mode, n_fixed=n_fixed)
        assert (q == q2).all()                       # counter-based: same (env, episode, seed) -> same draw
    if mode == 1:
        assert set(weights) == {n_fixed}
    if mode == 2:
        assert min(weights) >= n_fixed


def test_depolarizing_rates_cpu():
    """The sampler spec itself: P(X)=P(Y)=P(Z)=p/3 (ToricCode.py:141-149), chi-square on 3 categories."""
    d, p = 9, 0.1
    th = int(np.floor(p * 2 ** 32))
    cnt = np.zeros(4)
    for env in range(1500):
        q, _, _ = C.reset(d, env, 0, th, seed=7)
        cnt += np.bincount(q.reshape(-1), minlength=4)
    n = cnt.sum()
    assert abs(cnt[1:].sum() / n - p) < 4 * np.sqrt(p * (1 - p) / n)
    exp = cnt[1:].sum() / 3
    chi2 = ((cnt[1:] - exp) ** 2 / exp).sum()
    assert chi2 < 13.8     # chi-square(2 dof) at 0.001


def test_random_action_is_from_list():
    rng = np.random.default_rng(0)
    d = 5
    for env in range(100):
        q = ((rng.random((2, d, d)) < 0.1) * rng.integers(1, 4, (2, d, d))).astype(np.uint8)
        s = O.syndrome(q).astype(np.uint8)
        a = C.random_action(s, env, 1, 2, seed=5)
        pos = O.perspective_positions(s)
        if len(pos):
            assert any((a[:3] == p).all() for p in pos) and 1 <= a[3] <= 3
        else:
            assert tuple(a[:3]) == (0, 0, 0)


@pytest.mark.parametrize("d", (8, 10, 11, 13, 16, 20, 31, 32))
def test_c_oracle_equals_numpy_oracle_at_sizes_without_fixtures(d):
    """The two restatements are independent (numpy rolls vs explicit C index loops); at lattice sizes the golden
    fixtures do not cover they must still agree with each other -- syndrome, step, reward, terminal, logical flag
    and perspectives, including the largest size the engine accepts."""
    rng = np.random.default_rng(500 + d)
    for trial in range(3):
        p = (0.01, 0.08, 0.3)[trial]
        q = ((rng.random((2, d, d)) < p) * rng.integers(1, 4, (2, d, d))).astype(np.int64)
        s = O.syndrome(q)
        assert (C.syndrome(q) == s).all()
        assert C.ground_state(q) == O.ground_state(q)
        views, pos = O.perspectives(s)
        cv, cp = C.perspectives(s)
        assert (cv == views).all() and (cp == pos).all()
        qc, sc = q.astype(np.uint8), s.astype(np.uint8)
        qo, so = q, s
        for t in range(6):
            a = [int(rng.integers(2)), int(rng.integers(d)), int(rng.integers(d)), int(rng.integers(1, 4))]
            if t % 2 == 0:
                a[1], a[2] = (0, d - 1)[int(rng.integers(2))], (0, d - 1)[int(rng.integers(2))]
            qo, so2, r, term = O.step(qo, so, a)
            rc, tc = C.step(qc, sc, a)
            assert (qc == qo).all() and (sc == so2).all() and rc == r and tc == term
            so = so2
