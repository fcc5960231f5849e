"""Size-independent properties of the path, checked on the numpy oracle (CPU) at lattice sizes up to the engine's
maximum -- the same properties tests/test_gpu_actor.py::test_full_size_properties checks on the CUDA path at
BASELINE.json's full batch sizes.  They hold for the reference by construction (ToricCode.py:62-88,168-199;
src/util.py:46-150) and do not depend on any fixture."""
import numpy as np
import pytest

from oracle import toric_oracle as O


def _masks(rng, d, p, k):
    return ((rng.random((k, 2, d, d)) < p) * rng.integers(1, 4, (k, 2, d, d))).astype(np.int64)


@pytest.mark.parametrize("d", (3, 4, 7, 9, 16, 25, 32))
def test_syndrome_is_linear_and_translation_covariant(d):
    rng = np.random.default_rng(d)
    a, b = _masks(rng, d, 0.15, 8), _masks(rng, d, 0.15, 8)
    # Pauli product = XOR of the 2-bit codes (the rule table of ToricCode.py:57); the syndrome is linear in it
    assert (O.syndrome(a ^ b) == (O.syndrome(a) ^ O.syndrome(b))).all()
    r, c = int(rng.integers(d)), int(rng.integers(d))
    shifted = np.roll(np.roll(a, r, axis=-2), c, axis=-1)
    assert (O.syndrome(shifted) == np.roll(np.roll(O.syndrome(a), r, axis=-2), c, axis=-1)).all()


@pytest.mark.parametrize("d", (3, 5, 8, 15))
def test_step_is_an_involution_and_reward_is_the_defect_difference(d):
    rng = np.random.default_rng(10 + d)
    for q in _masks(rng, d, 0.1, 6):
        s = O.syndrome(q)
        a = [int(rng.integers(2)), int(rng.integers(d)), int(rng.integers(d)), int(rng.integers(1, 4))]
        q1, s1, r1, t1 = O.step(q, s, a)
        q2, s2, r2, t2 = O.step(q1, s1, a)
        assert (q2 == q).all() and (s2 == s).all()
        assert (s1 ^ s).sum() in (2, 4)                    # X / Z flips two stabilisers, Y flips four
        if not t1:
            assert r1 == s.sum() - s1.sum()
        if not t1 and not t2:
            assert r1 == -r2
        assert t1 == (not s1.any()) and (r1 == 100) == t1


@pytest.mark.parametrize("d", (3, 4, 5, 7, 12, 25))
def test_perspectives_count_and_translation_invariance(d):
    """One view per qubit next to a defect; the views are centred on their qubit, so translating the syndrome only
    permutes them (positions shift, contents do not change)."""
    rng = np.random.default_rng(20 + d)
    for q in _masks(rng, d, 0.05, 4):
        s = O.syndrome(q)
        views, pos = O.perspectives(s)
        mask = O.perspective_masks(s)
        assert len(views) == int(mask.sum()) <= 2 * d * d
        assert all(mask[m, i, j] for m, i, j in pos)
        assert (np.diff(pos[:, 0]) >= 0).all()             # all matrix-0 qubits first, then matrix 1
        r, c = int(rng.integers(d)), int(rng.integers(d))
        s2 = np.roll(np.roll(s, r, axis=-2), c, axis=-1)
        v2, p2 = O.perspectives(s2)
        key = {(int(m), int(i), int(j)): k for k, (m, i, j) in enumerate(pos)}
        assert len(v2) == len(views)
        for k2, (m, i, j) in enumerate(p2):
            k = key[(int(m), (int(i) - r) % d, (int(j) - c) % d)]
            assert (v2[k2] == views[k]).all()
        # odd d (the reference's domain): every view, rotated ones included, shows a defect next to its centre qubit
        # (0, g, g) -- that is why it was generated.  (For even d the rotation of rotate_state, src/util.py:87-94,
        # lands matrix-1 views one row higher.)
        g = d // 2
        if d % 2 == 1:
            for v in views:
                assert v[0, g, g] or v[0, (g + 1) % d, g] or v[1, g, g] or v[1, g, (g - 1) % d]


@pytest.mark.parametrize("d", (3, 5, 7, 9))
def test_logical_flag_is_invariant_under_stabilisers_for_odd_d(d):
    """evalGroundState (ToricCode.py:289-337) reads plane parities: applying a plaquette or a vertex stabiliser
    (X on a vertex' star or Z around a plaquette, trivial syndrome) flips two qubits in each matrix plane, so the flag cannot change."""
    rng = np.random.default_rng(30 + d)
    for q in _masks(rng, d, 0.1, 5):
        i, j = int(rng.integers(d)), int(rng.integers(d))
        # V reads the z-parts of the four qubits of a vertex' star (ToricCode.py:180-186), so the vertex stabiliser is
        # X on that star; Pl reads the x-parts around a plaquette (:191-197), so the plaquette stabiliser is Z there
        x_star = np.zeros_like(q)
        x_star[0, i, j] = x_star[0, (i - 1) % d, j] = 1
        x_star[1, i, j] = x_star[1, i, (j - 1) % d] = 1
        assert not O.syndrome(x_star).any()
        z_plaq = np.zeros_like(q)
        z_plaq[0, i, j] = z_plaq[0, i, (j + 1) % d] = 3
        z_plaq[1, i, j] = z_plaq[1, (i + 1) % d, j] = 3
        assert not O.syndrome(z_plaq).any()
        for stab in (x_star, z_plaq):
            assert O.ground_state(q ^ stab) == O.ground_state(q)
            assert (O.syndrome(q ^ stab) == O.syndrome(q)).all()
