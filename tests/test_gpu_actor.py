"""GPU: the fused actor step, the on-device sampler and full-size properties.

The device sampler's draw numbering is restated by the C oracle (oracle/toric_oracle.c: to_reset,
to_random_action, to_actor_steps), so whole trajectories -- random actions, steps, time-outs, resets,
perspectives -- are compared bit for bit.  The sampler's DISTRIBUTION is checked statistically against
what the reference's sampler produces (ToricCode.py:130-151: P(X)=P(Y)=P(Z)=p/3, conditioned on a
non-trivial syndrome)."""
import numpy as np
import pytest
import torch

from oracle import c_oracle as C
import golden
import sampler_checks as S
from oracle import toric_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg(cuda_device):
    import toric_rl_decoder_b200 as pkg
    return pkg


@pytest.mark.parametrize("d,mode,n_fixed,p", ((7, 0, 0, 0.1), (5, 0, 0, 0.003), (9, 1, 5, 0.1), (7, 2, 4, 0.05),
                                               (16, 0, 0, 0.02), (25, 0, 0, 0.05), (3, 1, 18, 0.1), (32, 2, 7, 0.01),
                                               (7, 0, 0, 0.0005), (7, 0, 0, 0.0103), (7, 0, 0, 0.0812), (7, 0, 0, 0.082), (9, 0, 0, 0.02), (15, 0, 0, 0.005), (32, 0, 0, 0.2), (3, 0, 0, 0.3),
                                               (25, 0, 0, 0.0001), (8, 0, 0, 0.1), (4, 2, 2, 0.3), (7, 0, 0, 0.99)))
def test_reset_matches_sampler_spec(pkg, d, mode, n_fixed, p):
    n, seed, base = 96, 4242, 1000
    eng = pkg.ToricEngine(d, n, seed=seed, env_base=base, p_error=p)
    eng.episode.fill_(3)
    eng.reset(mode=mode, n_fixed=n_fixed)
    q = eng.get_qubits().cpu().numpy()
    s = eng.get_state().cpu().numpy()
    th = int(pkg.p_to_thresh(p))
    for e in range(n):
        cq, cs, att = C.reset(d, base + e, 3, th, seed, mode=mode, n_fixed=n_fixed)
        assert (q[e] == cq).all() and (s[e] == cs).all(), (e, att)
    assert (eng.episode.cpu().numpy() == 4).all() and (eng.gs.cpu().numpy() == 1).all()
    # subset reset (resetTerminalEnvs): only the listed envs change
    idx = np.array([5, 17, 18, 90])
    eng.reset(idx=idx, mode=mode, n_fixed=n_fixed)
    q2 = eng.get_qubits().cpu().numpy()
    for e in range(n):
        if e in idx:
            cq, _, _ = C.reset(d, base + e, 4, th, seed, mode=mode, n_fixed=n_fixed)
            assert (q2[e] == cq).all()
        else:
            assert (q2[e] == q[e]).all()


def test_bare_generators(pkg):
    """generateRandomError / generateNRandomErrors called on their own (ToricCode.py:130-166): one draw of the
    device sampler without reset()'s rejection -- bit-exact vs the C oracle's one-draw mode, statistics of
    the reference's distribution (trivial syndromes included), env state untouched."""
    d, seed = 5, 77
    env = pkg.ToricCodeB200({"size": d, "p_error": 0.1}, seed=seed, env_base=9)
    s0 = env.reset()
    p = 0.004                                     # most draws have no error at all: never rejected
    th = int(pkg.p_to_thresh(p))
    n_empty = 0
    for k in range(40):
        m = env.generateRandomError(np.zeros((2, d, d), dtype=int), p)
        cq, _, att = C.reset(d, 9 + (1 << 40), k, th, seed, mode=16)
        assert att == 1 and m.shape == (2, d, d) and (m == cq).all()
        n_empty += int(not m.any())
    assert n_empty > 20
    buf = np.zeros((2, d, d), dtype=int)
    for k in range(10):
        out = env.generateNRandomErrors(buf, 3)
        cq, _, att = C.reset(d, 9 + (1 << 40), 40 + k, th, seed, mode=1 | 16, n_fixed=3)
        assert out is buf and (buf == cq).all() and (buf != 0).sum() == 3
    assert (env.generateNRandomErrors(buf, 0) == 0).all()
    assert (env.state == s0).all()
    # raw ABI: one-draw over a batch at p=0.1 reproduces the unconditional per-qubit rate
    n = 20000
    eng = pkg.ToricEngine(7, n, seed=1, p_error=0.1)
    eng.reset(mode=16)
    q = eng.get_qubits().cpu().numpy()
    rate = (q != 0).mean()
    assert abs(rate - 0.1) < 5 * np.sqrt(0.1 * 0.9 / q.size)
    for c in (1, 2, 3):
        assert abs((q == c).mean() - 0.1 / 3) < 5 * np.sqrt(0.0333 * 0.9667 / q.size)


@pytest.mark.parametrize("single_pass", (False, True))
@pytest.mark.parametrize("d,n,steps,max_steps,p", ((7, 512, 100, 75, 0.1), (5, 384, 60, 4, 0.02), (9, 128, 30, 10, 0.1),
                                                    (4, 64, 40, 6, 0.1), (17, 48, 12, 5, 0.05), (7, 1000, 30, 5, 0.003),
                                                    (3, 77, 20, 3, 0.2), (16, 33, 8, 3, 0.1), (8, 2051, 10, 4, 0.1)))
def test_fused_actor_trajectory_bit_exact(pkg, d, n, steps, max_steps, p, single_pass):
    """tc_actor_step (random actions, auto-reset) == the C oracle's actor loop, every step -- in its three-kernel
    form and in the single-pass form (TC_ACTOR_SINGLE_PASS: one kernel, offsets by decoupled look-back)."""
    seed, base = 11, 7 * n
    if single_pass and d > 16:
        pytest.skip("single-pass step: d <= 16")
    eng = pkg.ToricEngine(d, n, seed=seed, env_base=base, p_error=p, max_steps=max_steps, single_pass=single_pass)
    eng.reset()
    ref = C.ActorSlice(d, n, p, seed=seed, env_base=base, max_steps=max_steps)
    assert (eng.get_qubits().cpu().numpy() == ref.q).all()
    for t in range(steps):
        eng.actor_step(actions=None, auto_reset=True)
        ref.run(1)
        if t % 7 == 0 or t == steps - 1:
            assert (eng.get_qubits().cpu().numpy() == ref.q).all(), t
            assert (eng.get_state().cpu().numpy() == ref.s).all(), t
            assert (eng.steps.cpu().numpy() == ref.steps).all() and (eng.episode.cpu().numpy() == ref.episode).all()
            persp, pos, counts, offsets = eng.perspective_views()
            tot = int(ref.offsets[n])
            assert (offsets.cpu().numpy() == ref.offsets).all()
            assert (persp.cpu().numpy() == ref.out[:tot]).all()
            assert (pos.cpu().numpy()[:, :3] == ref.pos[:tot]).all()
    c = eng.read_counters()
    assert c["env_steps"] == n * steps == ref.counters[0]
    assert c["episodes"] == ref.counters[1] and c["solved"] == ref.counters[2]
    assert c["perspectives"] == ref.counters[3]
    assert c["episodes"] > 0 and c["timeouts"] + c["solved"] == c["episodes"]


def test_actor_step_host_policy_and_prev_state(pkg):
    """Host-supplied actions + keep_prev: the pre-reset post-step syndrome is what a transition stores."""
    d, n = 7, 256
    rng = np.random.default_rng(3)
    eng = pkg.ToricEngine(d, n, seed=5, p_error=0.1, max_steps=3)
    eng.reset()
    q = eng.get_qubits().cpu().numpy().astype(np.int64)
    s = eng.get_state().cpu().numpy().astype(np.int64)
    for t in range(8):
        acts = np.stack([rng.integers(0, 2, n), rng.integers(0, d, n), rng.integers(0, d, n), rng.integers(1, 4, n)], 1)
        steps_before = eng.steps.cpu().numpy().copy()
        eng.actor_step(actions=torch.from_numpy(acts), auto_reset=True, keep_prev=True)
        prev = eng.get_state(eng.prev_syn).cpu().numpy()
        rw, tm, gs = eng.reward.cpu().numpy(), eng.terminal.cpu().numpy(), eng.step_gs.cpu().numpy()
        q_now = eng.get_qubits().cpu().numpy().astype(np.int64)
        s_now = eng.get_state().cpu().numpy().astype(np.int64)
        for e in range(n):
            q1, ns, r, term = O.step(q[e], s[e], acts[e])
            assert (prev[e] == ns).all() and rw[e] == r and bool(tm[e]) == term
            assert bool(gs[e]) == O.ground_state(q1)
            done = term or steps_before[e] + 1 > 3
            if not done:
                assert (q_now[e] == q1).all() and (s_now[e] == ns).all()
            else:
                assert s_now[e].any() and (O.syndrome(q_now[e]) == s_now[e]).all()
        q, s = q_now, s_now


def test_sharding_invariance(pkg):
    """Counter-based RNG keyed by the GLOBAL env id: one engine over 0..n-1 == two engines over halves."""
    d, n, steps = 7, 512, 30
    full = pkg.ToricEngine(d, n, seed=9, env_base=0, p_error=0.1, max_steps=6)
    lo = pkg.ToricEngine(d, n // 2, seed=9, env_base=0, p_error=0.1, max_steps=6)
    hi = pkg.ToricEngine(d, n // 2, seed=9, env_base=n // 2, p_error=0.1, max_steps=6)
    for e in (full, lo, hi):
        e.reset()
    for _ in range(steps):
        for e in (full, lo, hi):
            e.actor_step()
    qf = full.get_qubits().cpu().numpy()
    assert (qf[:n // 2] == lo.get_qubits().cpu().numpy()).all()
    assert (qf[n // 2:] == hi.get_qubits().cpu().numpy()).all()
    cf, cl, ch = full.read_counters(), lo.read_counters(), hi.read_counters()
    for k in cf:
        assert cf[k] == cl[k] + ch[k], k                      # what the counter all-reduce computes
    pf = full.perspective_views()[0].cpu().numpy()
    pl_, ph = lo.perspective_views()[0].cpu().numpy(), hi.perspective_views()[0].cpu().numpy()
    assert (pf == np.concatenate([pl_, ph])).all()


@pytest.mark.parametrize("p", (0.1, 0.02))
def test_device_rng_statistics(pkg, p):
    """Per-Pauli rates and chi-square against the reference sampler's distribution.  At d=9 the
    rejection of zero-syndrome draws changes the marginal error rate by < 1e-4 relative at p=0.1
    (P(zero syndrome) ~ 0.9^162) and by ~3.8% of configurations at p=0.02, where the exact conditional
    expectation is used: E[errors | syndrome != 0] ~= (nq*p - 0) / (1 - P0) with P0 = (1-p)^nq up to
    O(p^4) stabiliser configurations."""
    d, n = 9, 100000
    eng = pkg.ToricEngine(d, n, seed=2024, p_error=p)
    eng.reset()
    q = eng.get_qubits()
    cnt = torch.bincount(q.reshape(-1).to(torch.int64), minlength=4).cpu().numpy().astype(np.float64)
    nq = 2 * d * d
    total = float(n * nq)
    p0 = (1 - p) ** nq
    exp_rate = p / (1 - p0)           # zero-error draws are rejected; they contribute no errors
    rate = cnt[1:].sum() / total
    sigma = np.sqrt(p * (1 - p) / total)
    assert abs(rate - exp_rate) < 5 * sigma + 1e-6, (rate, exp_rate)
    e3 = cnt[1:].sum() / 3
    chi2 = ((cnt[1:] - e3) ** 2 / e3).sum()
    assert chi2 < 13.8                # chi-square, 2 dof, 0.001
    # per-site uniformity: chi-square over the nq sites
    per_site = (q != 0).to(torch.float64).sum(dim=0).reshape(-1).cpu().numpy()
    es = per_site.mean()
    chi_site = ((per_site - es) ** 2 / es).sum()
    assert chi_site < nq + 5 * np.sqrt(2 * nq)
    # every env has a non-trivial syndrome and a consistent one
    s = eng.get_state()
    assert bool((s.reshape(n, -1).sum(dim=1) > 0).all())
    c = eng.read_counters()
    assert c["env_steps"] == 0


@pytest.mark.parametrize("d,p,n", ((7, 0.1, 400000), (7, 0.005, 1 << 20), (7, 0.0005, 1 << 20), (5, 0.0005, 1 << 20),
                                   (7, 0.0812, 400000), (7, 0.0820, 400000), (7, 0.02, 400000), (15, 0.05, 40000), (12, 0.01, 100000)))
def test_device_depolarizing_law(pkg, d, p, n):
    """The device sampler against the law of ToricCode.reset (ToricCode.py:108-117, :130-151) at the low error rates of
    BASELINE configs[3] too: weight histogram = Binomial(2d^2, p) given a non-trivial syndrome, uniform sites and
    codes, truncated-geometric first-error position; p = 0.0812 / 0.0820 straddle the switch between the sampler
    conditioned on >= 1 error and the plain rejection loop (csrc/sampler.cuh)."""
    eng = pkg.ToricEngine(d, n, seed=31, p_error=p)
    eng.reset()
    q = eng.get_qubits().cpu().numpy()
    s = eng.get_state()
    assert bool((s.reshape(n, -1).sum(dim=1) > 0).all())
    assert int((eng.status & 2).sum().item()) == 0
    S.check_depolarizing(q, p)
    S.check_first_error_position(q, p)


@pytest.mark.parametrize("d,k,n", ((5, 3, 400000), (3, 2, 100000), (7, 1, 100000)))
def test_device_fixed_weight_law(pkg, d, k, n):
    """generateNRandomErrors (ToricCode.py:153-166): a uniform random k-subset -- exact count, site uniformity,
    site-PAIR uniformity, code uniformity."""
    eng = pkg.ToricEngine(d, n, seed=32, p_error=0.1)
    eng.reset(mode=pkg.SAMPLE_FIXED_WEIGHT, n_fixed=k)
    S.check_fixed_weight(eng.get_qubits().cpu().numpy(), k)


@pytest.mark.parametrize("d,k,p,n", ((7, 4, 0.005, 400000), (5, 3, 0.05, 200000)))
def test_device_n_plus_q_law(pkg, d, k, p, n):
    """generateNPlusQRandomErrors (results/small_p_error_test.py:43-52): q fixed errors + Binomial(2d^2 - q, p)."""
    eng = pkg.ToricEngine(d, n, seed=33, p_error=p)
    eng.reset(mode=pkg.SAMPLE_FIXED_PLUS_P, n_fixed=k)
    S.check_n_plus_q(eng.get_qubits().cpu().numpy(), k, p)


def test_reset_p_zero_gives_up_at_once(pkg):
    eng = pkg.ToricEngine(5, 64, seed=1, p_error=0.0)
    eng.reset()
    assert not eng.get_qubits().any() and bool((eng.status & 2).all())


def _torch_reference_perspectives(state, pos, env_of, d):
    """Independent full-size check in torch: gather each listed perspective from the unpacked state."""
    g = d // 2
    dev = state.device
    a = torch.arange(d, device=dev).view(1, d, 1)
    b = torch.arange(d, device=dev).view(1, 1, d)
    m = pos[:, 0].long().view(-1, 1, 1)
    i = pos[:, 1].long().view(-1, 1, 1)
    j = pos[:, 2].long().view(-1, 1, 1)
    r0, c0 = (a + i - g) % d, (b + j - g) % d                       # m = 0
    r1 = (b + i - g) % d                                            # m = 1
    cv, cp = ((d - a) % d + j - g) % d, (d - 1 - a + j - g) % d
    rv = torch.where(m == 0, r0.expand(-1, d, d), r1.expand(-1, d, d))
    colv = torch.where(m == 0, c0.expand(-1, d, d), cv.expand(-1, d, d))
    colp = torch.where(m == 0, c0.expand(-1, d, d), cp.expand(-1, d, d))
    e = env_of.view(-1, 1, 1).expand(-1, d, d)
    out0 = state[e, 0, rv, colv]
    out1 = state[e, 1, rv, colp]
    return torch.stack((out0, out1), dim=1).float()


@pytest.mark.parametrize("d,n", ((7, 65536), (5, 30000), (8, 4096), (4, 8192), (3, 5000), (3, 300001), (9, 8192), (15, 4096),
                                 (25, 1024), (32, 300), (4, 530001)))      # the last two d <= 4 sizes: multi-CTA scan
def test_full_size_properties(pkg, d, n):
    """BASELINE configs[1] size (d=7, 65536 envs) and the other lattice sizes / both writer code paths: after a
    few fused steps, every perspective equals an independent torch gather; counts equal the popcount of the
    selection masks; offsets are their scan."""
    eng = pkg.ToricEngine(d, n, seed=1, p_error=0.1)
    eng.reset()
    for _ in range(5):
        eng.actor_step()
    persp, pos, counts, offsets = eng.perspective_views()
    state = eng.get_state().long()
    v, pl = state[:, 0] != 0, state[:, 1] != 0
    m0 = v | torch.roll(v, -1, 1) | pl | torch.roll(pl, 1, 2)
    m1 = v | torch.roll(v, -1, 2) | pl | torch.roll(pl, 1, 1)
    masks = torch.stack((m0, m1), 1)
    assert (masks.reshape(n, -1).sum(1).int() == counts).all()
    assert offsets[0].item() == 0 and (offsets[1:] == torch.cumsum(counts.long(), 0)).all()
    nz = torch.nonzero(masks)                       # row-major (env, m, i, j) == reference order
    assert (nz[:, 1:].to(torch.uint8) == pos[:, :3]).all()
    total = persp.shape[0]
    assert total == nz.shape[0]
    chunk = max(1024, (1 << 26) // (d * d))
    for lo in range(0, total, chunk):
        hi = min(total, lo + chunk)
        ref = _torch_reference_perspectives(state, pos[lo:hi], nz[lo:hi, 0], d)
        assert torch.equal(ref, persp[lo:hi])
    # roll/rotate preserve the number of defects in each plane
    sums = persp.sum(dim=(2, 3))
    assert torch.equal(sums, state.sum(dim=(2, 3))[nz[:, 0]].float())
    # incremental syndrome == full recompute after the steps
    before = eng.syn.clone()
    eng.syndrome()
    assert torch.equal(before, eng.syn)


@pytest.mark.parametrize("d,states", ((7, "u8"), (7, "rows"), (13, "rows"), (5, None), (7, "lazy"), (25, "lazy")))
def test_host_pipeline_matches_oracle(pkg, d, states):
    """HostActorPipeline (pinned host actions in, host results out, two steps in flight) == the numpy oracle."""
    n, steps = 300, 12
    eng = pkg.ToricEngine(d, n, seed=5, p_error=0.1, max_steps=1000)
    eng.reset()
    q = eng.get_qubits().cpu().numpy().astype(np.int64)
    s = eng.get_state().cpu().numpy().astype(np.int64)
    rng = np.random.default_rng(0)
    acts = [np.stack([rng.integers(0, 2, n), rng.integers(0, d, n), rng.integers(0, d, n), rng.integers(1, 4, n)],
                     axis=1).astype(np.int32) for _ in range(steps)]
    pinned = [torch.from_numpy(a).pin_memory() for a in acts]
    pipe = pkg.HostActorPipeline(eng, depth=2, states=states)
    got, probes = [], []

    def take(r):
        if states == "lazy":                       # the live view, indexed the way a host policy would
            assert isinstance(r["states"], pkg.hostpipe.LazyStates) and r["states"].shape == (n, 2, d, d)
            probes.append((r["states"][7].copy(), r["states"][[3, 299], 1].copy(), int(r["states"][5, 0, 1, 2])))
        got.append({key: np.array(val) for key, val in r.items()})
    for k in range(steps):
        pipe.submit(pinned[k], auto_reset=False)
        if k > 0:
            take(pipe.collect())
    take(pipe.collect())
    for k in range(steps):
        for e in range(n):
            q[e], s[e], rew, term = O.step(q[e], s[e], acts[k][e])
            assert got[k]["reward"][e] == rew and bool(got[k]["terminal"][e]) == term
        if states in ("u8", "lazy"):
            assert got[k]["states"].shape == s.shape and (got[k]["states"] == s).all()
            if states == "lazy":
                assert (probes[k][0] == s[7]).all() and (probes[k][1] == s[[3, 299], 1]).all()
                assert probes[k][2] == s[5, 0, 1, 2]
        elif states == "rows":
            assert (got[k]["state_rows"] == np.packbits(s.astype(np.uint8), axis=-1, bitorder="little")).all()
            assert (pipe.unpack_rows(got[k]["state_rows"]) == s).all()
        else:
            assert "states" not in got[k] and "state_rows" not in got[k]
        _, _, counts, _ = O.perspectives_batch(s)
        assert (got[k]["counts"] == counts).all()


def test_host_pipeline_lazy_states_equal_envset_golden(pkg):
    """The lazy `states` view of the closed-loop host pipeline (one step in flight: the next action is submitted
    only after the previous step was collected) equals what EnvSet.step returned in the reference's own recorded
    run (src/EnvSet.py:38-47; tests/golden/golden_batch_d7.npz), rewards and terminals included."""
    g = golden.batch_d7()
    n = len(g["qubits0"])
    eng = pkg.ToricEngine(7, n, seed=0, p_error=0.1, max_steps=1000)
    eng.set_qubits(g["qubits0"])
    pipe = pkg.HostActorPipeline(eng, depth=1)           # states="lazy" is the default
    for t in range(g["actions"].shape[0]):
        res = pipe.step(torch.from_numpy(g["actions"][t].astype(np.int32)).pin_memory(), auto_reset=False)
        st = res["states"]
        assert st.shape == g["states"][t].shape
        assert (np.asarray(st) == g["states"][t]).all()
        for e in (0, 17, n - 1):
            assert (st[e] == g["states"][t][e]).all()
        assert (res["reward"].astype(np.float64) == g["rewards"][t]).all()
        assert (res["terminal"] == g["terminals"][t]).all()
        assert res["state_rows"].nbytes * 7 == np.asarray(st).nbytes      # 1 byte per lattice row crosses PCIe at d=7


@pytest.mark.parametrize("d,n,steps", ((7, 65536, 40), (5, 100003, 12), (9, 20000, 8), (13, 3000, 6), (7, 1, 30), (7, 33, 30)))
def test_single_pass_step_equals_three_kernel_step(pkg, d, n, steps):
    """The single-pass step at BASELINE configs[1] size (and ragged / tiny / row-path sizes): every buffer the step
    touches -- states, results, counters, counts, offsets, positions, perspectives -- equals the three-kernel
    step's after every step; with host-supplied actions, kept previous syndromes and frozen envs too."""
    a = pkg.ToricEngine(d, n, seed=3, p_error=0.1, max_steps=9)
    b = pkg.ToricEngine(d, n, seed=3, p_error=0.1, max_steps=9, single_pass=True)
    for e in (a, b):
        e.reset()
    rng = np.random.default_rng(d)
    for t in range(steps):
        kw = {}
        if t % 3 == 1:                       # host policy + the transition's pre-reset syndrome
            acts = np.stack([rng.integers(0, 2, n), rng.integers(0, d, n), rng.integers(0, d, n),
                             rng.integers(0, 5, n)], 1)      # op 4 = a bad action, masked and counted
            kw = dict(actions=torch.from_numpy(acts), keep_prev=True)
        if t % 4 == 3:
            kw["freeze_done"] = True
            kw["auto_reset"] = False
        a.actor_step(**kw)
        b.actor_step(**kw)
        for name in ("qubits", "syn", "reward", "terminal", "step_gs", "status", "steps", "episode", "counts",
                     "actions", "counters", "overflow"):
            assert torch.equal(getattr(a, name), getattr(b, name)), (t, name)
        if "keep_prev" in kw:
            assert torch.equal(a.prev_syn, b.prev_syn)
        if t % 5 == 0 or t == steps - 1:
            for x, y in zip(a.perspective_views(), b.perspective_views()):
                assert torch.equal(x, y), t
    assert a.read_counters()["episodes"] > 0 or steps <= 10 or n < 10


@pytest.mark.parametrize("d,n,steps,single_pass", ((7, 1500, 400, False), (7, 40000, 120, False), (5, 300, 400, False),
                                                   (7, 1500, 400, True), (7, 40000, 120, True), (9, 700, 200, True)))
def test_pdl_chain_matches_serialized_launches(pkg, d, n, steps, single_pass):
    """Programmatic dependent launch (csrc/common.cuh: no exit path before pdl_wait) must not change results: long
    chains of back-to-back fused steps -- small batches, where a kernel's CTAs really are resident while the previous
    kernel drains -- against the TC_PDL=0 build of the same sources, where every launch takes a full dependency."""
    import os
    from toric_rl_decoder_b200 import _lib, buildlib
    if not os.path.exists(buildlib.NOPDL_LIB_PATH):
        pytest.skip("libtoric_b200_nopdl.so not built (__graft_entry__.build() makes it)")
    a = pkg.ToricEngine(d, n, seed=21, p_error=0.1, max_steps=12, single_pass=single_pass)
    b = pkg.ToricEngine(d, n, seed=21, p_error=0.1, max_steps=12, single_pass=single_pass)
    b.lib = _lib.bind(buildlib.NOPDL_LIB_PATH)
    for e in (a, b):
        e.reset()
    sums_a, sums_b = [], []
    for t in range(steps):
        a.actor_step()
        sums_a.append(a.out[:1024].sum() + a.offsets[n])        # reads behind the writer, enqueued in the chain
    for t in range(steps):
        b.actor_step()
        sums_b.append(b.out[:1024].sum() + b.offsets[n])
    torch.cuda.synchronize()
    assert torch.equal(torch.stack(sums_a), torch.stack(sums_b))
    for name in ("qubits", "syn", "steps", "episode", "counters", "counts", "offsets", "reward", "terminal"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    for x, y in zip(a.perspective_views(), b.perspective_views()):
        assert torch.equal(x, y)


@pytest.mark.parametrize("single_pass", (False, True))
def test_state_dict_exact_resume(pkg, tmp_path, single_pass):
    """run k steps, save, run k more == load into a fresh engine (through torch.save / torch.load), run k more."""
    d, n, k = 7, 5000, 40
    a = pkg.ToricEngine(d, n, seed=77, env_base=123456, p_error=0.07, max_steps=9, single_pass=single_pass)
    a.reset()
    for _ in range(k):
        a.actor_step(keep_prev=True)
    torch.save(a.state_dict(), tmp_path / "engine.pt")
    for _ in range(k):
        a.actor_step(keep_prev=True)
    b = pkg.ToricEngine(d, n, seed=0, env_base=0, p_error=0.5, max_steps=3, single_pass=single_pass)   # all overwritten
    b.load_state_dict(torch.load(tmp_path / "engine.pt"))
    for _ in range(k):
        b.actor_step(keep_prev=True)
    for name in pkg.ToricEngine._STATE_TENSORS + ("prev_syn", "counts", "offsets"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    for x, y in zip(a.perspective_views(), b.perspective_views()):
        assert torch.equal(x, y)
    assert a.read_counters() == b.read_counters() and a.read_counters()["episodes"] > 0
    with pytest.raises(ValueError):
        pkg.ToricEngine(5, n).load_state_dict(a.state_dict())


def test_single_pass_step_capacity_overflow_flag(pkg):
    eng = pkg.ToricEngine(7, 500, seed=1, p_error=0.1, single_pass=True)
    eng.reset()
    eng.ensure_capacity(1000)                       # far below sum P
    eng.actor_step()
    torch.cuda.synchronize()
    assert int(eng.overflow.item()) == 1
    persp, pos, counts, offsets = eng.perspective_views()       # regenerates into a large enough buffer
    ref = pkg.ToricEngine(7, 500, seed=1, p_error=0.1)
    ref.reset()
    ref.actor_step()
    for x, y in zip((persp, pos, counts, offsets), ref.perspective_views()):
        assert torch.equal(x, y)


@pytest.mark.parametrize("d,n", ((7, 3000), (15, 200)))
def test_cuda_graph_replay_equals_eager(pkg, d, n):
    """k fused steps captured into a CUDA graph (PDL edges included) and replayed == the same steps launched
    eagerly: states, counters, perspectives."""
    a = pkg.ToricEngine(d, n, seed=11, p_error=0.1, max_steps=6)
    b = pkg.ToricEngine(d, n, seed=11, p_error=0.1, max_steps=6, single_pass=(d == 7))   # graph replays reuse the
    for e in (a, b):                                       # kernel arguments: the look-back epoch lives on the device
        e.reset()
    for _ in range(20):
        a.actor_step()
    g = b.capture_actor_steps(5)
    for _ in range(4):
        g.replay()
    torch.cuda.synchronize()
    assert torch.equal(a.qubits, b.qubits) and torch.equal(a.syn, b.syn)
    assert torch.equal(a.steps, b.steps) and torch.equal(a.episode, b.episode)
    assert torch.equal(a.counters, b.counters)
    pa, pb = a.perspective_views(), b.perspective_views()
    for x, y in zip(pa, pb):
        assert torch.equal(x, y)


def test_multi_gpu_nccl(pkg):
    """scripts/gpu_multi_check.py under torchrun on 2 GPUs: weight broadcast, sharded evaluate == single-GPU
    evaluate, actor-step slices == one big batch, counters all-reduce to the big batch's.  Needs >= 2 GPUs."""
    import os
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(root, "scripts", "gpu_multi_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    assert '"multi_gpu_check": "ok"' in res.stdout
