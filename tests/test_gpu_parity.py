"""GPU parity: libtoric_b200.so (through the C ABI / the host mirror) against the golden fixtures
generated from the reference and against the numpy oracle on seeded inputs.  Bit-exact everywhere."""
import numpy as np
import pytest
import torch

import golden
from oracle import toric_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg(cuda_device):
    import toric_rl_decoder_b200 as pkg
    from toric_rl_decoder_b200 import _lib
    _lib.load()
    return pkg


def _views(eng, n=None):
    persp, pos, counts, offsets = eng.perspective_views(n)
    return (persp.cpu().numpy(), pos.cpu().numpy()[:, :3].astype(np.int64), counts.cpu().numpy().astype(np.int64),
            offsets.cpu().numpy())


@pytest.mark.parametrize("d", golden.ALL_D)
def test_inject_syndrome_logical(pkg, d):
    g = golden.per_d(d)
    k = len(g["qubits"])
    eng = pkg.ToricEngine(d, k)
    eng.set_qubits(g["qubits"])
    assert (eng.get_qubits().cpu().numpy() == g["qubits"]).all()          # pack/unpack round trip
    assert (eng.get_state().cpu().numpy() == g["state"]).all()            # createSyndromOpt
    assert (eng.ground_state().cpu().numpy().astype(bool) == g["gs"]).all()  # evalGroundState


@pytest.mark.parametrize("d", golden.ALL_D)
def test_perspectives_golden(pkg, d):
    g = golden.per_d(d)
    k = len(g["state"])
    eng = pkg.ToricEngine(d, k)
    for states, counts, pos, views in ((g["state"], g["counts"], g["positions"], g["views"]),
                                       (g["step_state"][:, -1], g["counts_last"], g["positions_last"],
                                        g["views_last"])):
        eng.set_state(states)
        eng.perspectives()
        p, ps, c, off = _views(eng)
        assert (c == counts).all()
        assert off[0] == 0 and (off[1:] == np.cumsum(counts)).all()
        assert (ps == pos).all()
        assert p.dtype == np.float32 and p.shape == views.shape
        assert (p == views.astype(np.float32)).all()


@pytest.mark.parametrize("d", (3, 5))
def test_exhaustive_masks(pkg, d):
    g = golden.exhaustive(d)
    k = len(g["qubits"])
    eng = pkg.ToricEngine(d, k)
    eng.set_qubits(g["qubits"])
    assert (eng.get_state().cpu().numpy() == g["state"]).all()
    assert (eng.ground_state().cpu().numpy().astype(bool) == g["gs"]).all()
    eng.perspectives()
    p, ps, c, off = _views(eng)
    assert (c == g["counts"]).all() and (ps == g["positions"]).all()
    assert (p == g["views"].astype(np.float32)).all()


@pytest.mark.parametrize("d", golden.ALL_D)
def test_step_sequences_golden(pkg, d):
    g = golden.per_d(d)
    k, T = g["actions"].shape[:2]
    eng = pkg.ToricEngine(d, k)
    eng.set_qubits(g["qubits"])
    for t in range(T):
        reward, terminal = eng.step(torch.from_numpy(g["actions"][:, t].astype(np.int64)))
        assert (eng.get_qubits().cpu().numpy() == g["step_qubits"][:, t]).all()
        assert (eng.get_state().cpu().numpy() == g["step_state"][:, t]).all()     # incremental == full recompute
        assert (reward.cpu().numpy().astype(np.float64) == g["step_reward"][:, t]).all()
        assert (terminal.cpu().numpy().astype(bool) == g["step_terminal"][:, t]).all()
        assert (eng.ground_state().cpu().numpy().astype(bool) == g["step_gs"][:, t]).all()   # sticky
    assert int(eng.status.max().item()) == 0


def test_envset_mirror_golden(pkg):
    """EnvSetB200 behaves like src/EnvSet.py on the reference's own recorded run (dtypes included)."""
    g = golden.batch_d7()
    n = len(g["qubits0"])
    es = pkg.EnvSetB200({"size": 7, "p_error": 0.1}, n)
    assert es.size == 7 and es.no_envs == n
    es.engine.set_qubits(g["qubits0"])
    for t in range(g["actions"].shape[0]):
        st, rw, tm, info = es.step(g["actions"][t])
        assert st.dtype == np.int64 and rw.dtype == np.float64 and tm.dtype == np.bool_ and info == {}
        assert (st == g["states"][t]).all() and (rw == g["rewards"][t]).all() and (tm == g["terminals"][t]).all()
    persp, pos, counts, offsets = pkg.generatePerspectiveBatch(3, 7, st)
    assert persp.dtype == torch.float32 and persp.is_cuda
    assert (offsets.cpu().numpy()[1:] == g["splice_idx"]).all()
    assert (pos.cpu().numpy() == g["positions"]).all()
    assert (persp.cpu().numpy() == g["views"].astype(np.float32)).all()


def test_perspective_batch_returns_fresh_tensors(pkg):
    """generatePerspectiveBatch returns fresh tensors like the reference (src/numba/util_actor.py:56-67): a second
    call at the same d (perspectives of next_state after state) leaves the first result untouched."""
    g = golden.batch_d7()
    s0, s1 = g["states"][0], g["states"][-1]
    first = pkg.generatePerspectiveBatch(3, 7, s0)
    snap = [t.clone() for t in first]
    second = pkg.generatePerspectiveBatch(3, 7, s1)
    for a, b in zip(first, snap):
        assert torch.equal(a, b)
    for a, b in zip(first, second):
        assert a.data_ptr() != b.data_ptr() or a.numel() == 0
    want, wpos, wcounts, _ = O.perspectives_batch(s1)
    assert (second[0].cpu().numpy() == want.astype(np.float32)).all()
    assert (second[1].cpu().numpy() == wpos).all() and (second[2].cpu().numpy() == wcounts).all()
    assert first[0].shape[0] == int(first[3][-1].item())           # sized to sum P, not to a worst-case buffer
    empty = pkg.generatePerspectiveBatch(3, 7, np.zeros((4, 2, 7, 7), np.int64))
    assert empty[0].shape == (0, 2, 7, 7) and empty[3].tolist() == [0, 0, 0, 0, 0]


def test_envset_envs_views(pkg):
    """EnvSet.envs[i] (src/EnvSet.py:14-16): per-env ToricCode surface on one row of the batch; the other envs of
    the batch are not touched by anything done through a view."""
    d, n = 5, 7
    es = pkg.EnvSetB200({"size": d, "p_error": 0.1}, n, seed=5)
    s0 = es.resetAll()
    assert len(es.envs) == n and len(es.envs[1:3]) == 2 and es.envs[-1]._i == n - 1
    with pytest.raises(IndexError):
        es.envs[n]
    q_all = es.engine.get_qubits().cpu().numpy().astype(np.int64)
    for i, env in enumerate(es.envs):
        assert (env.qubit_matrix == q_all[i]).all() and (env.state == s0[i]).all() and env.ground_state
        assert abs(env.p_error - 0.1) < 1e-9 and env.system_size == d
    # inject a mask into env 2, step it, evaluate its logical flag: oracle on that env, bit-identical rest
    rng = np.random.default_rng(3)
    q = (rng.random((2, d, d)) < 0.2) * rng.integers(1, 4, (2, d, d))
    v = es.envs[2]
    v.qubit_matrix = q
    s = O.syndrome(q)
    assert (v.qubit_matrix == q).all() and (v.state == s).all()
    qo, so, gs = q.astype(np.int64), s, True
    for t in range(6):
        a = [int(rng.integers(2)), int(rng.integers(d)), int(rng.integers(d)), int(rng.integers(1, 4))]
        st, r, term, info = v.step(a)
        qo, so2, ro, to = O.step(qo, so, a)
        assert (st == so2).all() and r == ro and term == to and info == {}
        so = so2
        gs = O.ground_state(qo, gs)
        assert v.evalGroundState() == gs
    now = es.engine.get_qubits().cpu().numpy().astype(np.int64)
    keep = [i for i in range(n) if i != 2]
    assert (now[keep] == q_all[keep]).all() and (now[2] == qo).all()
    assert (es.engine.steps.cpu().numpy() == np.array([0, 0, 6, 0, 0, 0, 0])).all()
    # state setter + reset of one env
    v.state = np.zeros((2, d, d), dtype=int)
    assert v.isTerminalState(v.state)
    st, r, term, _ = v.step([0, 1, 1, 3])                    # inconsistent pair: reward vs the assigned (empty) state
    q_v, ns_v, ro_v, to_v = O.step(qo, np.zeros((2, d, d), dtype=np.int64), [0, 1, 1, 3])
    assert (st == ns_v).all() and r == ro_v and term == to_v
    qo = q_v
    s3 = es.envs[3].reset(p_error=0.3)
    assert s3.any() and (O.syndrome(es.envs[3].qubit_matrix) == s3).all() and abs(es.envs[3].p_error - 0.3) < 1e-9
    now = es.engine.get_qubits().cpu().numpy().astype(np.int64)
    keep = [i for i in range(n) if i not in (2, 3)]
    assert (now[keep] == q_all[keep]).all()


def test_envset_step_after_state_assignment(pkg):
    """envs[i].state = X followed by the BATCHED EnvSet.step: env i takes its reward against X and gets its new state
    from the qubits (ToricCode.py:83-86), exactly as a per-env step would; the other envs take the incremental path;
    the mark is gone afterwards (the second batched step is incremental for everybody)."""
    d, n = 5, 6
    es = pkg.EnvSetB200({"size": d, "p_error": 0.15}, n, seed=9)
    s0 = es.resetAll()
    q0 = es.engine.get_qubits().cpu().numpy().astype(np.int64)
    fake = np.zeros((2, d, d), dtype=np.int64)
    fake[0, 1, 1] = fake[1, 3, 2] = fake[1, 0, 4] = fake[0, 4, 4] = 1
    es.envs[4].state = fake
    es.envs[1].state = np.zeros((2, d, d), dtype=int)
    rng = np.random.default_rng(21)
    acts = np.stack([rng.integers(0, 2, n), rng.integers(0, d, n), rng.integers(0, d, n), rng.integers(1, 4, n)], 1)
    st, r, term, _ = es.step(acts)
    start = {4: fake, 1: np.zeros((2, d, d), dtype=np.int64)}
    q1, s1 = [], []
    for i in range(n):
        qi, si, ri, ti = O.step(q0[i], start.get(i, s0[i]), list(acts[i]))
        assert (st[i] == si).all() and r[i] == ri and term[i] == ti, i
        q1.append(qi)
        s1.append(si)
    assert not es._state_assigned
    acts2 = np.stack([rng.integers(0, 2, n), rng.integers(0, d, n), rng.integers(0, d, n), rng.integers(1, 4, n)], 1)
    st2, r2, term2, _ = es.step(acts2)
    for i in range(n):
        _, si, ri, ti = O.step(q1[i], s1[i], list(acts2[i]))
        assert (st2[i] == si).all() and r2[i] == ri and term2[i] == ti, i


@pytest.mark.parametrize("kind", ("single", "view"))
def test_start_from_state_call_sequence(pkg, kind):
    """The call sequence of results/start_from_state.py:34-58 (reset, createSyndromOpt of a chosen error,
    assignment of qubit_matrix and state, plotToricCode, step loop) against a ToricCodeB200 and against
    EnvSet.envs[i]; the network is replaced by "first perspective, operator X" so that the test needs no weights."""
    d = 5
    if kind == "single":
        env = pkg.ToricCodeB200({"size": d, "p_error": 0.1})
    else:
        es = pkg.EnvSetB200({"size": d, "p_error": 0.1}, 4, seed=3)
        es.resetAll()
        env = es.envs[2]
    assert env.system_size == d and int(env.action_space.high[-1]) == 3 and env.min_qbit_errors == 0
    assert env.rule_table[1, 2] == 3 and env.observation_space.shape == (2, d, d)
    start_error = np.zeros((2, d, d), dtype=int)
    start_error[0, 1, 2] = 1
    start_error[1, 3, 3] = 2
    start_error[0, 4, 0] = 3
    env.reset()
    state = env.createSyndromOpt(start_error)
    assert (state == O.syndrome(start_error)).all()
    env.qubit_matrix = start_error
    env.state = state
    assert env.plotToricCode(state, "step_0") is None
    q, terminal_state, no_steps = start_error.astype(np.int64), False, 0
    while not terminal_state and no_steps < 12:
        _, pos = pkg.generatePerspectiveOptimized(d // 2, d, state)
        action = [int(pos[0][0]), int(pos[0][1]), int(pos[0][2]), 1]
        next_state, reward, terminal_state, _ = env.step(action)
        q, ns, ro, to = O.step(q, state, action)
        assert (next_state == ns).all() and reward == ro and terminal_state == to
        assert (env.next_state == ns).all() and (env.qubit_matrix == q).all()
        state = next_state
        no_steps += 1
    # the generators called on their own (ToricCode.py:130-166)
    m = env.generateRandomError(np.zeros((2, d, d), dtype=int), 0.3)
    assert m.shape == (2, d, d) and m.min() >= 0 and m.max() <= 3
    m = env.generateNRandomErrors(np.zeros((2, d, d), dtype=int), 4)
    assert int((m != 0).sum()) == 4
    assert (env.qubit_matrix == q).all()                       # neither generator touched the env


def test_single_env_api(pkg):
    """ToricCodeB200 mirrors gym_ToricCode's ToricCode, injected with a golden mask."""
    g = golden.per_d(5)
    env = pkg.ToricCodeB200({"size": 5, "min_qbit_errors": 0, "p_error": 0.1})
    assert env.system_size == 5 and int(env.action_space.high[-1]) == 3
    e = 1
    env.qubit_matrix = g["qubits"][e]
    assert (env.state == g["state"][e]).all() and env.state.dtype == np.int64
    assert (env.createSyndromOpt(g["qubits"][e]) == g["state"][e]).all()
    for t in range(g["actions"].shape[1]):
        st, r, term, info = env.step(list(g["actions"][e, t]))
        assert (st == g["step_state"][e, t]).all() and r == g["step_reward"][e, t]
        assert term == bool(g["step_terminal"][e, t]) and env.isTerminalState(st) == term
        assert env.evalGroundState() == bool(g["step_gs"][e, t])
    views, pos = pkg.generatePerspectiveOptimized(2, 5, g["state"][e].astype(np.int64))
    ov, op = O.perspectives(g["state"][e].astype(np.int64))
    assert len(views) == len(ov) and all((a == b).all() for a, b in zip(views, ov))
    assert pos == [tuple(int(x) for x in p) for p in op]
    pl = pkg.generatePerspective(2, 5, g["state"][e].astype(np.int64))
    assert all((a.perspective == b).all() and a.position == tuple(c) for a, b, c in zip(pl, ov, op))
    s = env.reset(p_error=0.2)
    assert s.shape == (2, 5, 5) and s.any() and env.ground_state is True
    assert (O.syndrome(env.qubit_matrix) == s).all()
    # an INCONSISTENT pair, as plain attribute assignment allows in the reference (results/small_p_error_test.py:119-120):
    # the step's reward is taken against the assigned state, the new state is recomputed from the qubits (ToricCode.py:83-86)
    q = g["qubits"][2].astype(np.int64)
    env.qubit_matrix = q
    fake = np.zeros((2, 5, 5), dtype=int)
    fake[0, 1, 1] = fake[1, 3, 2] = fake[1, 0, 4] = 1
    env.state = fake
    assert (env.state == fake).all()
    gs_before = env.ground_state
    a = [1, 2, 3, 2]
    st, r, term, _ = env.step(a)
    q1, ns, ro, to = O.step(q, fake, a)
    assert (st == ns).all() and r == ro and term == to and (env.qubit_matrix == q1).all()
    assert env.ground_state == gs_before                     # assignment and step leave the sticky flag alone
    st, r, term, _ = env.step([0, 0, 0, 1])                  # consistent again: the incremental path
    q2, ns2, ro2, to2 = O.step(q1, ns, [0, 0, 0, 1])
    assert (st == ns2).all() and r == ro2 and term == to2


@pytest.mark.parametrize("d,n", ((7, 1024), (9, 256), (4, 64), (16, 64), (17, 32), (32, 8)))
def test_seeded_random_vs_oracle(pkg, d, n):
    """Seeded masks and action sequences at sizes the oracle finishes in seconds."""
    rng = np.random.default_rng(100 + d)
    q = (rng.random((n, 2, d, d)) < 0.08) * rng.integers(1, 4, (n, 2, d, d))
    q = q.astype(np.uint8)
    q[0] = 0                                   # empty env: no perspectives, terminal syndrome
    q[1] = rng.integers(0, 4, (2, d, d))       # dense env
    eng = pkg.ToricEngine(d, n)
    eng.set_qubits(q)
    s = O.syndrome(q)
    assert (eng.get_state().cpu().numpy() == s).all()
    qq = q.astype(np.int64)
    T = 4
    for t in range(T):
        acts = np.stack([rng.integers(0, 2, n), rng.integers(0, d, n), rng.integers(0, d, n), rng.integers(0, 4, n)], 1)
        reward, terminal = eng.step(torch.from_numpy(acts))
        rw, tm = reward.cpu().numpy(), terminal.cpu().numpy()
        for e in range(n):
            qq[e], ns, r, term = O.step(qq[e], s[e], acts[e])
            s[e] = ns
            assert rw[e] == r and bool(tm[e]) == term
        assert (eng.get_state().cpu().numpy() == s).all()
        assert (eng.get_qubits().cpu().numpy() == qq).all()
    eng.perspectives()
    p, ps, c, off = _views(eng)
    op, opos, oc, ooff = O.perspectives_batch(s)
    assert (c == oc).all() and (off == ooff).all() and (ps == opos).all()
    assert (p == op).all()


def test_bad_actions_are_masked(pkg):
    d, n = 5, 8
    eng = pkg.ToricEngine(d, n)
    q = np.zeros((n, 2, d, d), np.uint8)
    q[:, 0, 1, 1] = 1
    eng.set_qubits(q)
    acts = np.array([[2, 0, 0, 1], [0, 5, 0, 1], [0, 0, -1, 1], [0, 0, 0, 4], [0, 0, 0, 0], [1, 4, 4, 3],
                     [0, 1, 1, 1], [-1, 0, 0, 1]])
    reward, terminal = eng.step(torch.from_numpy(acts))
    st = eng.status.cpu().numpy()
    assert list(st) == [1, 1, 1, 1, 0, 0, 0, 1]
    after = eng.get_qubits().cpu().numpy()
    for e in (0, 1, 2, 3, 4, 7):
        assert (after[e] == q[e]).all() and reward[e].item() == 0.0
    assert terminal[6].item() == 1 and reward[6].item() == 100.0


def test_capacity_overflow_regenerates(pkg):
    d, n = 7, 64
    rng = np.random.default_rng(5)
    q = ((rng.random((n, 2, d, d)) < 0.1) * rng.integers(1, 4, (n, 2, d, d))).astype(np.uint8)
    eng = pkg.ToricEngine(d, n)
    eng.set_qubits(q)
    eng.ensure_capacity(10)      # far too small on purpose
    eng.perspectives()
    p, ps, c, off = _views(eng)
    op, opos, oc, ooff = O.perspectives_batch(O.syndrome(q))
    assert (c == oc).all() and (p == op).all() and (ps == opos).all()


@pytest.mark.parametrize("d,n", ((3, 1000), (4, 777), (5, 1234), (7, 500), (9, 300), (15, 60)))
def test_partial_capacity_writes_the_envs_that_fit_and_nothing_else(pkg, d, n):
    """A perspective buffer smaller than sum P (every writer form: group writer d <= 5, plane path, row path): the envs
    whose perspectives fit are written exactly, nothing lands past the capacity, offsets are the true ones, the flag is
    set."""
    rng = np.random.default_rng(d * 1000 + n)
    q = ((rng.random((n, 2, d, d)) < 0.08) * rng.integers(1, 4, (n, 2, d, d))).astype(np.uint8)
    ref = pkg.ToricEngine(d, n)
    ref.set_qubits(q)
    ref.perspectives()
    p_ref, pos_ref, c_ref, off_ref = ref.perspective_views()
    total = int(off_ref[-1].item())
    for cap in (total // 2 + 1, total - 1, 1):
        eng = pkg.ToricEngine(d, n)
        eng.set_qubits(q)
        eng.ensure_capacity(cap + 64)
        eng.cap = cap                                   # the library sees `cap`; 64 guard perspectives behind it
        eng.out.view(torch.uint8).fill_(0x5a)
        eng.pos.fill_(0x5a)
        eng.perspectives()
        torch.cuda.synchronize()
        assert int(eng.overflow.item()) == 1
        assert torch.equal(eng.offsets[:n + 1], off_ref) and torch.equal(eng.counts[:n], c_ref)
        ends = off_ref[1:]
        fit = int((ends <= cap).sum().item())           # offsets grow with the env index: the fitting envs are a prefix
        kept = int(off_ref[fit].item())
        assert torch.equal(eng.out[:kept], p_ref[:kept]) and torch.equal(eng.pos[:kept, :3], pos_ref[:kept, :3])
        assert (eng.out[cap:].view(torch.uint8) == 0x5a).all() and (eng.pos[cap:] == 0x5a).all()


def test_abi_argument_errors(pkg):
    from toric_rl_decoder_b200 import _lib
    lib = _lib.load()
    assert lib.tc_syndrome(None, None, 7, 4, None) == -1
    assert lib.tc_syndrome(None, None, 2, 4, None) == -1
    assert lib.tc_syndrome(None, None, 33, 4, None) == -1
    eng = pkg.ToricEngine(7, 4)
    small = torch.zeros(8, dtype=torch.uint8, device="cuda")
    eng.ensure_capacity(16)
    rc = lib.tc_perspectives(eng.syn.data_ptr(), eng.out.data_ptr(), eng.pos.data_ptr(), eng.counts.data_ptr(),
                             eng.offsets.data_ptr(), eng.overflow.data_ptr(), eng.cap, small.data_ptr(), 8, 7, 4, None)
    assert rc == -2
    with pytest.raises(ValueError):
        pkg.ToricEngine(2, 4)


@pytest.mark.parametrize("dtype", ("bfloat16", "float16", "uint8"))
@pytest.mark.parametrize("d,n", ((3, 700), (4, 300), (5, 2000), (7, 4096), (8, 512), (9, 1024), (15, 256), (25, 64),
                                 (32, 16)))
def test_perspective_element_types(pkg, d, n, dtype):
    """tc_persp_write_as: the bf16 / fp16 / u8 tensors carry exactly the float32 tensor's 0/1 values (all
    head/tail alignments of the env runs occur: offsets are arbitrary), and nothing is written past the
    last perspective."""
    tdt = getattr(torch, dtype)
    rng = np.random.default_rng(7 * d + n)
    q = ((rng.random((n, 2, d, d)) < 0.06) * rng.integers(1, 4, (n, 2, d, d))).astype(np.uint8)
    q[0] = 0
    q[n // 2] = rng.integers(0, 4, (2, d, d))
    ref = pkg.ToricEngine(d, n)
    ref.set_qubits(q)
    ref.perspectives()
    p32, pos32, c32, off32 = ref.perspective_views()
    eng = pkg.ToricEngine(d, n, out_dtype=tdt)
    eng.set_qubits(q)
    total = int(off32[-1].item())
    eng.ensure_capacity(total + 3)
    guard = 0x5a
    eng.out.view(torch.uint8).fill_(guard)
    eng.perspectives()
    p, pos, c, off = eng.perspective_views()
    assert p.dtype == tdt and p.shape == p32.shape
    assert torch.equal(c, c32) and torch.equal(off, off32) and torch.equal(pos[:, :3], pos32[:, :3])
    assert torch.equal(p.float(), p32)
    assert (eng.out[total:].view(torch.uint8) == guard).all()
    if n <= 1024:
        op, opos, oc, ooff = O.perspectives_batch(O.syndrome(q))
        assert (p.float().cpu().numpy() == op).all()
    # fused actor step path with the same element type
    e32 = pkg.ToricEngine(d, n, seed=3, p_error=0.1)
    e16 = pkg.ToricEngine(d, n, seed=3, p_error=0.1, out_dtype=tdt)
    for e in (e32, e16):
        e.reset()
        e.actor_step()
        e.actor_step()
    a, b = e32.perspective_views(), e16.perspective_views()
    assert torch.equal(a[0], b[0].float()) and torch.equal(a[1], b[1]) and torch.equal(a[3], b[3])


def test_empty_batch(pkg):
    """Zero environments is a legal batch everywhere (EnvSet(env, 0) in the reference gives empty arrays)."""
    d = 5
    envs = pkg.EnvSetB200({"size": d, "p_error": 0.1}, 0)
    assert envs.resetAll().shape == (0, 2, d, d)
    st, rw, tm, _ = envs.step(np.zeros((0, 4), dtype=np.int64))
    assert st.shape == (0, 2, d, d) and rw.shape == (0,) and tm.shape == (0,)
    assert envs.resetTerminalEnvs(np.zeros(0, dtype=np.int64)).shape == (0, 2, d, d)
    eng = envs.engine
    eng.actor_step()
    eng.ground_state()
    persp, pos, counts, offsets = eng.perspective_views()
    assert persp.shape == (0, 2, d, d) and pos.shape[0] == 0 and counts.shape == (0,)
    assert offsets.cpu().tolist() == [0]
    out = pkg.generatePerspectiveBatch(d // 2, d, np.zeros((0, 2, d, d), dtype=np.int64))
    assert out[0].shape[0] == 0
    # a batch in which no env has a defect: all counts zero, nothing written
    eng2 = pkg.ToricEngine(d, 9)
    eng2.set_qubits(np.zeros((9, 2, d, d), np.uint8))
    eng2.perspectives()
    p2, _, c2, o2 = eng2.perspective_views()
    assert p2.shape[0] == 0 and (c2 == 0).all() and (o2 == 0).all()


def test_engine_on_another_device_fails_loudly(pkg):
    """The library launches on the current device: an engine living elsewhere raises instead of launching on the
    wrong GPU, and works inside torch.cuda.device(...)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    with torch.cuda.device(1):
        eng = pkg.ToricEngine(5, 16, device="cuda:1", seed=2)
        eng.reset()
        s1 = eng.get_state().cpu()
    assert torch.cuda.current_device() == 0
    with pytest.raises(RuntimeError):
        eng.reset()
    with torch.cuda.device(1):
        eng.actor_step()
        assert eng.get_state().device.index == 1 and s1.shape == (16, 2, 5, 5)
