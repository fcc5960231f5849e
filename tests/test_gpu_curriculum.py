"""tc_p_error_curriculum: the p_error schedule of the reference's actor (src/Actor_mp.py:43-47,171-181) applied on the
device after every fused step, against the numpy restatement (oracle/toric_oracle.py:p_error_curriculum) bit for bit
along a run with auto-resets, plus the law of the 'random' draw."""
import numpy as np
import pytest
import torch

from oracle import toric_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg(cuda_device):
    import toric_rl_decoder_b200 as pkg
    return pkg


def _u32(t):
    return t.cpu().numpy().view(np.uint32).copy()


@pytest.mark.parametrize("strategy", ("linear", "random"))
def test_curriculum_follows_the_reference_schedule(pkg, strategy):
    d, n, seed, base = 5, 3000, 31, 1000
    start, final, delta = 0.03, 0.11, 0.02
    eng = pkg.ToricEngine(d, n, seed=seed, env_base=base, p_error=start, max_steps=6)
    eng.reset()
    eng.set_p_error_curriculum(start, final, delta, strategy)
    ts, td, tf = (int(pkg.p_to_thresh(x)) for x in (start, delta, final))
    rnd = strategy == "random"
    # priming: every env's first auto-reset will use ceiling start + delta (Actor_mp.py:47,176)
    cap = np.full(n, ts, dtype=np.uint32)
    thr = np.full(n, ts, dtype=np.uint32)
    cap, thr = O.p_error_curriculum(None, _u32(eng.episode), cap, thr, ts, td, tf, rnd, True, seed, base)
    assert (_u32(eng.p_cap) == cap).all() and (_u32(eng.p_thresh) == thr).all()
    assert (cap == min(tf, ts + td)).all()
    resets = np.zeros(n, dtype=np.int64)
    used = []                                   # (ceiling, p) pairs the auto-resets actually sampled with
    for k in range(40):
        thr_before, cap_before = thr.copy(), cap.copy()
        eng.actor_step()                        # random actions, auto-reset, then the curriculum kernel
        torch.cuda.synchronize()
        steps = eng.steps.cpu().numpy()
        just_reset = steps == 0
        resets += just_reset
        used.append((cap_before[just_reset], thr_before[just_reset]))
        cap, thr = O.p_error_curriculum(steps, _u32(eng.episode), cap, thr, ts, td, tf, rnd, False, seed, base)
        assert (_u32(eng.p_cap) == cap).all() and (_u32(eng.p_thresh) == thr).all(), k
    assert resets.min() >= 3                    # max_steps = 6: every env was reset several times
    # the ceiling an env's k-th reset used is min(final, start + k * delta), exactly the reference's
    expect = np.minimum(tf, ts + (resets + 1) * td)          # ceiling waiting for reset number resets + 1
    assert (cap == expect).all() and cap.max() == tf
    caps = np.concatenate([c for c, _ in used]).astype(np.float64)
    ps = np.concatenate([p for _, p in used]).astype(np.float64)
    assert len(ps) > 5 * n
    if not rnd:
        assert (ps == caps).all()
    else:
        lo = np.minimum(ts, caps)
        assert (ps >= lo).all() and (ps <= caps).all()
        u = (ps - lo) / (caps - lo)             # uniform on [0, 1]: mean 1/2, variance 1/12, flat histogram
        m = len(u)
        assert abs(u.mean() - 0.5) <= 5 * np.sqrt(1.0 / 12.0 / m)
        hist = np.histogram(u, bins=16, range=(0.0, 1.0))[0]
        chi2 = ((hist - m / 16.0) ** 2 / (m / 16.0)).sum()
        assert chi2 < 60.0                      # 15 degrees of freedom: the 1e-6 quantile is 52


def test_curriculum_argument_errors(pkg):
    from toric_rl_decoder_b200 import _lib
    lib = _lib.load()
    assert lib.tc_p_error_curriculum(None, None, None, None, 1, 1, 1, 0, 0, 0, 0, 4, None) == -1
    assert lib.tc_p_error_curriculum(None, None, None, None, 1, 1, 1, 0, 0, 0, 0, 0, None) == 0
    assert lib.tc_p_error_curriculum(None, None, None, None, 1, 1, 1, 0, 0, 0, 0, -1, None) == -1
    eng = pkg.ToricEngine(5, 8, p_error=0.1)
    with pytest.raises(ValueError):
        eng.set_p_error_curriculum(0.1, 0.2, 0.01, "quadratic")
