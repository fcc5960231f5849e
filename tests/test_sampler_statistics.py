"""The sampler's DISTRIBUTION, checked on the C oracle (oracle/toric_oracle.c:to_reset restates the device's draw
numbering, csrc/sampler.cuh, so this is the spec the GPU tests then compare the kernels with bit for bit and
statistically).  All three modes of the reference: depolarizing with rejection (ToricCode.py:108-117, :130-151),
fixed weight (:153-166), n plus q (results/small_p_error_test.py:43-52)."""
import numpy as np
import pytest

import sampler_checks as S
from oracle import c_oracle as C


def _thresh(p):
    return int(min(np.floor(p * 4294967296.0), 4294967295))


@pytest.mark.parametrize("d,p,n", ((7, 0.1, 60000), (7, 0.005, 200000), (7, 0.0005, 200000), (5, 0.0005, 200000),
                                   (9, 0.02, 60000), (7, 0.0812, 100000), (7, 0.0820, 100000), (25, 0.05, 600)))
def test_depolarizing_law(d, p, n):
    """p = 0.0812 / 0.0820 straddle p * 2d^2 = 8 at d=7: the conditioned sampler below, the plain rejection loop
    above -- both must give the same law."""
    q, s, att = C.reset_batch(d, n, 1 << 33, 2, _thresh(p), 99)
    assert (s.reshape(n, -1).sum(1) > 0).all() and (att < (1 << 20)).all()
    nq = 2 * d * d
    if _thresh(p) * nq < (8 << 32):
        assert att.mean() < 1.001                     # conditioned on >= 1 error: (almost) never rejected
    S.check_depolarizing(q, p)
    if n >= 20000:
        S.check_first_error_position(q, p)


def test_depolarizing_p_zero_gives_up_at_once():
    q, s, att = C.reset_batch(5, 4, 0, 0, 0, 1)
    assert not q.any() and not s.any() and (att == (0x80000000 | 1)).all()


@pytest.mark.parametrize("d,k,n", ((5, 3, 120000), (3, 2, 40000), (7, 1, 30000)))
def test_fixed_weight_law(d, k, n):
    q, s, att = C.reset_batch(d, n, 7, 0, _thresh(0.1), 5, mode=1, n_fixed=k)
    assert (s.reshape(n, -1).sum(1) > 0).all()
    S.check_fixed_weight(q, k)


def test_fixed_weight_one_draw_keeps_trivial_syndromes():
    """generateNRandomErrors on its own (ToricCode.py:153-166) does not reject anything."""
    q, s, att = C.reset_batch(3, 20000, 0, 0, _thresh(0.1), 5, mode=1 | 16, n_fixed=4)
    assert (att == 1).all()
    S.check_fixed_weight(q, 4)


@pytest.mark.parametrize("d,k,p,n", ((7, 4, 0.005, 100000), (5, 3, 0.05, 60000)))
def test_n_plus_q_law(d, k, p, n):
    q, s, att = C.reset_batch(d, n, 0, 1, _thresh(p), 3, mode=2, n_fixed=k)
    S.check_n_plus_q(q, k, p)
