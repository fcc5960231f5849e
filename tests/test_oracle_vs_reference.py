"""Live cross-check of both oracles against the UNMODIFIED reference, imported from /root/reference.

Runs only where the reference checkout exists (the build container); skipped on the GPU box, where the
committed golden fixtures (tests/golden/, made by oracle/gen_golden*.py from the same import) carry the
pin.  Inputs here are fresh random masks and action sequences, not the ones the fixtures hold, and include
even d (where evalGroundState is blind, ToricCode.py:294) and the numba twins of the perspective path.
"""
import numpy as np
import pytest

from oracle import c_oracle as C
from oracle import ref_loader
from oracle import toric_oracle as O

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference checkout not present")


@pytest.fixture(scope="module")
def ref():
    return ref_loader.load()


def _env(ref, d, p=0.1):
    return ref.gym.make("toric-code-v0", config={"size": d, "min_qbit_errors": 0, "p_error": p})


def _inject(env, q):
    # the reference's own scripts inject masks by attribute assignment (results/small_p_error_test.py:119-120)
    env.qubit_matrix = np.array(q, dtype=int)
    env.state = env.createSyndromOpt(env.qubit_matrix)
    env.next_state = env.state
    env.ground_state = True


@pytest.mark.parametrize("d", (3, 4, 5, 6, 7, 8, 9, 12, 15, 25, 32))
def test_syndrome_step_logical_vs_reference(ref, d):
    rng = np.random.default_rng(1000 + d)
    env = _env(ref, d)
    for trial in range(12):
        p = (0.02, 0.1, 0.3)[trial % 3]
        q = (rng.random((2, d, d)) < p) * rng.integers(1, 4, (2, d, d))
        _inject(env, q)
        s = np.array(env.state)
        assert (O.syndrome(q) == s).all() and (C.syndrome(q) == s).all()
        env.evalGroundState()
        assert O.ground_state(q) == bool(env.ground_state) == C.ground_state(q)
        qo, so, gs = np.array(q, dtype=np.int64), s.astype(np.int64), bool(env.ground_state)
        qc, sc = qo.astype(np.uint8), so.astype(np.uint8)     # C.step works in place on uint8
        for t in range(10):
            a = [int(rng.integers(2)), int(rng.integers(d)), int(rng.integers(d)), int(rng.integers(1, 4))]
            if t % 3 == 0:                      # wrap-around border
                a[1], a[2] = (0, d - 1)[int(rng.integers(2))], (0, d - 1)[int(rng.integers(2))]
            st, r, term, _ = env.step(np.array(a))
            qo, so2, ro, to = O.step(qo, so, a)
            rc, tc = C.step(qc, sc, a)
            assert (np.array(env.qubit_matrix) == qo).all() and (qc == qo).all()
            assert (np.array(st) == so2).all() and (sc == so2).all()
            assert r == ro == rc and bool(term) == to == tc
            assert bool(env.isTerminalState(st)) == O.is_terminal(so2)
            so = so2
            env.evalGroundState()
            gs = O.ground_state(qo, gs)
            assert gs == bool(env.ground_state)


@pytest.mark.parametrize("d", (3, 4, 5, 6, 7, 9, 15, 25, 32))
def test_perspectives_vs_reference_numpy_and_numba(ref, d):
    rng = np.random.default_rng(2000 + d)
    g = d // 2
    for trial in range(6 if d <= 15 else 2):
        p = (0.01, 0.1, 0.4)[trial % 3]
        q = (rng.random((2, d, d)) < p) * rng.integers(1, 4, (2, d, d))
        s = O.syndrome(q)
        views, pos = O.perspectives(s)
        rv, rp = ref.util.generatePerspectiveOptimized(g, d, np.array(s, dtype=int))
        loop = ref.util.generatePerspective(g, d, np.array(s, dtype=int))
        assert len(rv) == len(views) == len(loop)
        for k in range(len(rv)):
            assert (np.array(rv[k]) == views[k]).all() and tuple(rp[k]) == tuple(pos[k])
            assert (loop[k].perspective == views[k]).all() and tuple(loop[k].position) == tuple(pos[k])
        cv, cp = C.perspectives(s)
        assert (cv == views).all() and (cp == pos).all()
        if ref.numba_util is not None:
            nv, npos = ref.numba_util.generatePerspectiveOptimized(g, d, np.array(s, dtype=np.int64))
            assert len(nv) == len(views)
            for k in range(len(nv)):
                assert (np.asarray(nv[k]) == views[k]).all() and tuple(npos[k]) == tuple(pos[k])
    # an all-zero syndrome has no perspectives on either side
    z = np.zeros((2, d, d), dtype=int)
    assert len(ref.util.generatePerspectiveOptimized(g, d, z)[0]) == 0 == len(O.perspectives(z)[1])


@pytest.mark.parametrize("d", (5, 6, 7, 9))
def test_rotate_shift_vs_reference(ref, d):
    rng = np.random.default_rng(3000 + d)
    g = d // 2
    for _ in range(8):
        a = rng.integers(0, 2, (2, d, d))
        b = rng.integers(0, 2, (2, d, d))
        assert (ref.util.rotate_state(a.copy()) == O.rotate_state(a)).all()
        i, j = int(rng.integers(d)), int(rng.integers(d))
        ra, rb = ref.util.shift_state(i, j, a.copy(), b.copy(), g)
        oa, ob = O.shift_state(i, j, a, b, g)
        assert (np.array(ra) == oa).all() and (np.array(rb) == ob).all()


def test_envset_batch_layout_vs_reference(ref):
    """EnvSet.step + numba generatePerspectiveBatch + the concatenate of selectActionBatch
    (src/EnvSet.py:38-47, src/numba/util_actor.py:35-39,56-67) against O.step / O.perspectives_batch."""
    if ref.numba_util_actor is None:
        pytest.skip("numba twins not importable: %s" % getattr(ref, "numba_error", ""))
    d, n = 7, 6
    rng = np.random.default_rng(77)
    np.random.seed(77)
    es = ref.EnvSet(_env(ref, d), n)
    states = np.array(es.resetAll())
    qs = np.stack([np.array(e.qubit_matrix) for e in es.envs]).astype(np.int64)
    assert (O.syndrome(qs) == states).all()
    for t in range(5):
        per, pos, counts = ref.numba_util_actor.generatePerspectiveBatch(d // 2, d, states.astype(np.int64))
        op, opos, oc, ooff = O.perspectives_batch(states)
        assert list(counts) == list(oc) and (np.cumsum(np.array(list(counts))) == ooff[1:]).all()
        flat = np.concatenate([np.asarray(v) for env in per for v in env]).reshape(-1, 2, d, d)
        assert (flat.astype(np.float32) == op).all()
        assert (np.array([tuple(x) for env in pos for x in env]).reshape(-1, 3) == opos).all()
        acts = np.zeros((n, 4), dtype=int)
        for e in range(n):
            k = int(rng.integers(oc[e])) if oc[e] else 0
            row = opos[ooff[e] + k] if oc[e] else (0, 0, 0)
            acts[e] = (row[0], row[1], row[2], int(rng.integers(1, 4)))
        nxt, rew, term, _ = es.step(acts)
        for e in range(n):
            qs[e], ns, r, tm = O.step(qs[e], states[e], acts[e])
            assert (np.array(nxt[e]) == ns).all() and rew[e] == r and bool(term[e]) == tm
        states = np.array(nxt)
