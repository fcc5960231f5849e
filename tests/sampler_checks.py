"""Distribution checks for the error samplers, shared by the CPU suite (the C oracle = the sampler's spec) and
the GPU suite (the device kernels).  Each check takes sampled qubit matrices (n,2,d,d) of Pauli codes 0..3 and
compares them with the law of the REFERENCE's sampler:

  depolarizing  ToricCode.reset (ToricCode.py:108-117, :130-151): every qubit errs independently with probability p,
                Pauli ~ U{1,2,3}; draws with a trivial syndrome are rejected
  fixed weight  generateNRandomErrors (ToricCode.py:153-166): a uniform random n-subset of the 2d^2 qubits
  n plus q      generateNPlusQRandomErrors (results/small_p_error_test.py:43-52): q fixed-weight errors, then the
                depolarizing field on the qubits that are still free

All thresholds are chi-square quantiles at 1e-6 (scipy): a correct sampler fails a check once in 10^6 runs, a biased one (a wrong
threshold bit, a skipped slot, correlated picks) overshoots them by orders of magnitude at these sample sizes.
"""
import numpy as np
from scipy import stats

ALPHA = 1e-6


def _chi2_ok(observed, expected, what, ddof=0):
    observed = np.asarray(observed, dtype=np.float64)
    expected = np.asarray(expected, dtype=np.float64)
    assert expected.min() >= 5.0, "%s: expected counts too small for a chi-square test: %s" % (what, expected.min())
    chi = float(((observed - expected) ** 2 / expected).sum())
    dof = observed.size - 1 - ddof
    limit = float(stats.chi2.ppf(1.0 - ALPHA, dof))
    assert chi < limit, "%s: chi2 %.1f >= %.1f (dof %d)" % (what, chi, limit, dof)
    return chi


def _weights(q):
    return (q.reshape(q.shape[0], -1) != 0).sum(axis=1)


def check_codes_uniform(q, what="Pauli codes"):
    cnt = np.array([(q == c).sum() for c in (1, 2, 3)], dtype=np.float64)
    _chi2_ok(cnt, np.full(3, cnt.sum() / 3.0), what)


def check_sites_uniform(q, what="error sites"):
    per_site = (q != 0).reshape(q.shape[0], -1).sum(axis=0).astype(np.float64)
    _chi2_ok(per_site, np.full(per_site.size, per_site.mean()), what)


def binom_pmf(n, p, kmax):
    return stats.binom.pmf(np.arange(kmax + 1), n, p)


def check_depolarizing(q, p, what="depolarizing"):
    """Weight histogram against Binomial(2d^2, p) conditioned on a non-trivial syndrome.  A non-empty error set with a
    trivial syndrome is a product of stabilisers / logical loops (weight >= 4 here, probability O(p^4) and below
    1e-7 of the accepted mass at the p used), so the accepted law is Binomial given weight >= 1 up to that."""
    n, nq = q.shape[0], q[0].size
    w = _weights(q)
    assert w.min() >= 1
    kmax = int(max(3, stats.binom.ppf(1 - 1e-3, nq, p) + 1))
    pmf = binom_pmf(nq, p, kmax)
    cond = pmf[1:] / (1.0 - pmf[0])
    # merge the tails so that every expected count is comfortable
    lo = 1
    while lo < kmax and n * cond[:lo].sum() < 20:
        lo += 1
    hi = kmax
    while hi > lo and n * (1.0 - cond[:hi - 1].sum()) < 20:
        hi -= 1
    edges = list(range(lo, hi))                      # bins: [1..lo], lo+1, ..., hi-1, [hi..]
    obs = [float((w <= lo).sum())] + [float((w == k).sum()) for k in range(lo + 1, hi)] + [float((w >= hi).sum())]
    exp = [n * cond[:lo].sum()] + [n * cond[k - 1] for k in range(lo + 1, hi)] + [n * (1.0 - cond[:hi - 1].sum())]
    if len(obs) >= 2:
        _chi2_ok(obs, exp, what + ": weight histogram (bins %s)" % (edges,))
    check_codes_uniform(q, what + ": codes")
    check_sites_uniform(q, what + ": sites")


def check_first_error_position(q, p, bins=14, what="first error"):
    """Linear index (m, i, j order) of the first error against the truncated geometric law
    P(K = k) = (1-p)^k p / (1 - (1-p)^N) -- the law the conditioned low-p sampler draws from and the one the plain
    rejection loop implies."""
    n, nq = q.shape[0], q[0].size
    first = (q.reshape(n, -1) != 0).argmax(axis=1)
    k = np.arange(nq)
    pk = (1 - p) ** k * p / (1 - (1 - p) ** nq)
    groups = [list(g) for g in np.array_split(k, bins)]
    while len(groups) > 2 and n * pk[groups[-1]].sum() < 20:       # merge the thin tail (large p: K is small)
        tail = groups.pop()
        groups[-1] = groups[-1] + tail
    obs = [float(np.isin(first, g).sum()) for g in groups]
    exp = [n * pk[g].sum() for g in groups]
    _chi2_ok(obs, exp, what)


def check_fixed_weight(q, n_fixed, what="fixed weight"):
    """Exactly n errors; every site equally likely; every unordered PAIR of sites equally likely (a uniform random
    n-subset has uniform 2-subsets -- catches a sampler whose later picks depend on the earlier ones); codes uniform
    and independent of the site."""
    n, nq = q.shape[0], q[0].size
    flat = (q.reshape(n, -1) != 0)
    assert (flat.sum(axis=1) == n_fixed).all()
    check_sites_uniform(q, what + ": sites")
    check_codes_uniform(q, what + ": codes")
    if n_fixed >= 2:
        co = flat.astype(np.float64).T @ flat.astype(np.float64)          # co-occurrence counts
        iu = np.triu_indices(nq, 1)
        pairs = co[iu]
        exp = n * n_fixed * (n_fixed - 1) / 2.0 / pairs.size
        _chi2_ok(pairs, np.full(pairs.size, exp), what + ": site pairs")


def check_n_plus_q(q, n_fixed, p, what="n plus q"):
    """weight - q ~ Binomial(2d^2 - q, p) (the field only lands on free qubits); sites and codes uniform."""
    n, nq = q.shape[0], q[0].size
    extra = _weights(q) - n_fixed
    assert extra.min() >= 0
    kmax = int(max(2, stats.binom.ppf(1 - 1e-3, nq - n_fixed, p) + 1))
    pmf = binom_pmf(nq - n_fixed, p, kmax)
    hi = kmax
    while hi > 1 and n * (1.0 - pmf[:hi].sum()) < 20:
        hi -= 1
    obs = [float((extra == k).sum()) for k in range(hi)] + [float((extra >= hi).sum())]
    exp = [n * pmf[k] for k in range(hi)] + [n * (1.0 - pmf[:hi].sum())]
    _chi2_ok(obs, exp, what + ": extra-weight histogram")
    check_sites_uniform(q, what + ": sites")
    check_codes_uniform(q, what + ": codes")
