"""GPU: policy-side kernels (tc_select_action, tc_transitions, tc_td_priorities, tc_segmented_max) through
the C ABI against the reference-made fixtures (tests/golden/golden_policy.npz, transitions_d9.npz) and
the numpy oracle."""
import numpy as np
import pytest
import torch

import golden as G
from oracle import c_oracle as C
from oracle import toric_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg(cuda_device):
    import toric_rl_decoder_b200 as pkg
    return pkg


@pytest.fixture(scope="module")
def pol():
    return G.load("golden_policy.npz")


class TableModel(torch.nn.Module):
    """Stand-in network that returns a fixed Q table (rows in perspective order)."""

    def __init__(self, q):
        super().__init__()
        self.q = q

    def forward(self, x):
        assert x.shape[0] == self.q.shape[0] and x.dtype == torch.float32
        return self.q


@pytest.mark.parametrize("d", (5, 7))
def test_select_action_greedy_matches_reference(pkg, pol, d):
    states = pol["sel_states%d" % d]
    q = torch.from_numpy(pol["sel_q%d" % d]).cuda()
    actions, q_v = pkg.selectActionBatch(3, 0.0, d // 2, d, states, TableModel(q))
    assert actions.dtype == np.int64
    assert (actions == pol["sel_actions%d" % d]).all()
    assert (q_v == pol["sel_qv%d" % d]).all()
    # single-env form
    a0, q0 = pkg.selectAction(3, 0.0, d // 2, d, states[0], TableModel(q[: int(pol["sel_counts%d" % d][0])]))
    assert a0 == [int(x) for x in pol["sel_actions%d" % d][0]] and (q0 == pol["sel_qv%d" % d][0]).all()


@pytest.mark.parametrize("d,n,eps", ((7, 4096, 0.3), (5, 1000, 1.0), (9, 512, 0.05)))
def test_select_action_epsilon_draws(pkg, d, n, eps):
    """Device epsilon-greedy draws == the documented Philox numbering (checked with the C oracle's Philox),
    and the random fraction is binomially consistent with epsilon."""
    seed, base = 99, 12345
    eng = pkg.ToricEngine(d, n, seed=seed, env_base=base, p_error=0.1)
    eng.reset()
    eng.steps.copy_(torch.arange(n, dtype=torch.int32, device="cuda") % 7)
    eng.perspectives()
    persp, pos, counts, offsets = eng.perspective_views()
    total = persp.shape[0]
    gen = torch.Generator(device="cuda").manual_seed(5)
    q = (torch.randint(-8, 9, (total, 3), generator=gen, device="cuda").float() * 0.125).contiguous()
    was = torch.empty(n, dtype=torch.uint8, device="cuda")
    sel = torch.empty(n, dtype=torch.int64, device="cuda")
    actions, q_sel = pkg.select_action_device(q, pos, offsets, n, epsilon=eps, episode=eng.episode, steps=eng.steps,
                                              seed=seed, env_base=base, sel_idx=sel, was_greedy=was)
    torch.cuda.synchronize()
    th = int(pkg.p_to_thresh(eps))
    ep, st = eng.episode.cpu().numpy(), eng.steps.cpu().numpy()
    cn, off = counts.cpu().numpy(), offsets.cpu().numpy()
    greedy = np.zeros(n, bool)
    rp, ra = np.zeros(n, np.int64), np.zeros(n, np.int64)
    for e in range(n):
        g = base + e
        u = C.philox((g & 0xffffffff, g >> 32, int(ep[e]) & 0xffffffff, 0x60000000 | int(st[e])), (seed & 0xffffffff, seed >> 32))
        greedy[e] = not (int(u[0]) < th)
        rp[e] = (int(u[1]) * int(cn[e])) >> 32
        ra[e] = (int(u[2]) * 3) >> 32
    oa, oq, orow = O.select_action_prime(q.cpu().numpy(), off, pos.cpu().numpy()[:, :3], greedy, rp, ra)
    assert (was.cpu().numpy().astype(bool) == greedy).all()
    assert (actions.cpu().numpy() == oa).all() and (q_sel.cpu().numpy() == oq).all()
    assert (sel.cpu().numpy() == orow).all()
    k = int((~greedy).sum())
    sigma = max(np.sqrt(n * eps * (1 - eps)), 1e-9)
    assert abs(k - n * eps) <= 5 * sigma + 1e-6


def test_select_and_first_error_draws_are_independent(pkg):
    """The epsilon-greedy draw of (episode E, step 0) and the first-error draw (attempt 0) of the conditioned reset
    that runs with episode counter E live in different Philox counter domains (csrc/common.cuh).  When they shared
    0x40000000, "random action" meant U < eps for the reset as well, i.e. a first error at a low linear index: the
    mean first-error position of the exploring envs then sat far below that of the greedy ones."""
    d, n, eps, seed, base = 5, 1 << 16, 0.5, 4242, 77
    eng = pkg.ToricEngine(d, n, seed=seed, env_base=base, p_error=0.05)      # p * 2d^2 = 2.5 < 8: conditioned sampler
    st = np.zeros((n, 2, d, d), np.uint8)
    st[:, 0, 1, 1] = 1
    eng.set_state(st)                                                         # one defect: every env has perspectives
    eng.perspectives()
    persp, pos, counts, offsets = eng.perspective_views()
    q = torch.zeros((persp.shape[0], 3), dtype=torch.float32, device="cuda")
    was = torch.empty(n, dtype=torch.uint8, device="cuda")
    pkg.select_action_device(q, pos, offsets, n, epsilon=eps, episode=eng.episode, steps=eng.steps, seed=seed,
                             env_base=base, was_greedy=was)
    eng.reset()                                                               # same episode counter as the draw above
    torch.cuda.synchronize()
    codes = eng.get_qubits().cpu().numpy().reshape(n, -1)
    first = (codes != 0).argmax(axis=1).astype(np.float64)                    # q = m*d*d + i*d + j of the first error
    explore = was.cpu().numpy() == 0
    k = int(explore.sum())
    assert abs(k - n * eps) <= 5 * np.sqrt(n * eps * (1 - eps))
    diff = first[explore].mean() - first[~explore].mean()
    sigma = first.std() * np.sqrt(1.0 / k + 1.0 / (n - k))
    assert abs(diff) <= 5 * sigma, (diff, sigma)


def test_select_action_empty_segments(pkg):
    d, n = 5, 6
    eng = pkg.ToricEngine(d, n)
    st = np.zeros((n, 2, d, d), np.uint8)
    st[1, 0, 2, 2] = 1
    st[4, 1, 0, 4] = 1
    eng.set_state(st)
    eng.perspectives()
    persp, pos, counts, offsets = eng.perspective_views()
    q = torch.arange(persp.shape[0] * 3, dtype=torch.float32, device="cuda").reshape(-1, 3)
    actions, q_sel = pkg.select_action_device(q, pos, offsets, n)
    a = actions.cpu().numpy()
    assert (a[[0, 2, 3, 5]] == 0).all() and (q_sel.cpu().numpy()[[0, 2, 3, 5]] == 0).all()
    assert a[1, 3] == 3 and a[4, 3] == 3      # last row, last column holds the maximum


@pytest.mark.parametrize("d", (4, 5, 7, 9))
def test_transitions_match_reference(pkg, pol, d):
    st, ns, act = pol["tr_state%d" % d], pol["tr_next%d" % d], pol["tr_action%d" % d]
    op, ao, _, on, _ = pkg.generateTransitionParallel(act, pol["tr_reward%d" % d], st, ns, pol["tr_terminal%d" % d],
                                                      d // 2)
    assert (op.cpu().numpy() == pol["tr_out_prev%d" % d]).all()
    assert (on.cpu().numpy() == pol["tr_out_next%d" % d]).all()
    aon = ao.cpu().numpy()
    assert (aon[:, :3] == pol["tr_out_pos%d" % d]).all() and (aon[:, 3] == pol["tr_out_op%d" % d]).all()
    # record-array form with the reference's structured dtype (src/Actor_mp.py:52-56)
    action_type = np.dtype([('position', (int, 3)), ('op', int)])
    ttype = np.dtype([('perspective', (int, (2, d, d))), ('action', action_type), ('reward', float),
                      ('next_perspective', (int, (2, d, d))), ('terminal', bool)])
    rec = pkg.generateTransitionParallel(act, pol["tr_reward%d" % d], st, ns, pol["tr_terminal%d" % d], d // 2, ttype)
    assert (rec['perspective'] == pol["tr_out_prev%d" % d]).all()
    assert (rec['reward'] == pol["tr_reward%d" % d]).all() and (rec['terminal'] == pol["tr_terminal%d" % d]).all()


def test_transitions_known_answers_d9(pkg):
    """The reference's stored transitions (output_speed_test/transitions_0.npy): stepping the centred state
    with the stored action at (m,g,g) reproduces the stored next_state and reward."""
    g = G.load("transitions_d9.npz")
    d = g["state"].shape[-1]
    n = len(g["state"])
    act = np.concatenate([g["position"], g["op"][:, None]], axis=1).astype(np.int64)
    assert (act[:, 1:3] == d // 2).all()
    # a view centred at (g,g) is a fixed point of the m=0 transform
    m0 = act[:, 0] == 0
    op, ao, _, on, _ = pkg.generateTransitionParallel(act[m0], g["reward"][m0], g["state"][m0], g["next_state"][m0],
                                                      g["terminal"][m0], d // 2)
    assert (op.cpu().numpy() == g["state"][m0]).all() and (on.cpu().numpy() == g["next_state"][m0]).all()
    assert n >= 1


def test_td_priorities(pkg, pol):
    out = pkg.computePrioritiesParallel(pol["pr_A"], pol["pr_R"], pol["pr_Q"], pol["pr_Qns"], float(pol["pr_discount"]))
    assert out.shape == pol["pr_out"].shape
    assert np.allclose(out.cpu().numpy(), pol["pr_out"], rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("d", (5, 7))
def test_predict_max_matches_reference(pkg, pol, d):
    W = torch.from_numpy(pol["pm_W%d" % d]).cuda()
    b = torch.from_numpy(pol["pm_bias%d" % d]).cuda()

    class Lin(torch.nn.Module):
        def forward(self, x):
            return x.reshape(x.shape[0], -1) @ W + b

    torch.backends.cuda.matmul.allow_tf32 = False
    out = pkg.predictMaxOptimized(Lin(), torch.from_numpy(pol["pm_states%d" % d]), d // 2, d)
    assert (out.numpy() == pol["pm_out%d" % d]).all()


def test_freeze_done_leaves_solved_envs_untouched(pkg):
    d, n = 5, 8
    eng = pkg.ToricEngine(d, n, max_steps=1000)
    q = np.zeros((n, 2, d, d), np.uint8)
    q[1, 0, 1, 1] = 1
    q[3, 1, 2, 2] = 3
    eng.set_qubits(q)
    acts = np.tile(np.array([[0, 0, 0, 1]], np.int32), (n, 1))
    acts[1] = (0, 1, 1, 1)          # corrects env 1
    eng.reward.fill_(-7.0)
    flags = pkg.engine.ACTOR_FREEZE_DONE
    eng.actor_step(actions=torch.from_numpy(acts), auto_reset=False, freeze_done=True)
    torch.cuda.synchronize()
    st = eng.get_state().cpu().numpy()
    steps = eng.steps.cpu().numpy()
    assert flags == 4
    assert (st[[0, 2, 4, 5, 6, 7]] == 0).all() and (steps[[0, 2, 4, 5, 6, 7]] == 0).all()
    assert (eng.reward.cpu().numpy()[[0, 2]] == -7.0).all()
    assert steps[1] == 1 and steps[3] == 1 and eng.terminal.cpu().numpy()[1] == 1 and st[1].sum() == 0
    assert st[3].sum() > 0
    assert eng.read_counters()["env_steps"] == 2


@pytest.mark.parametrize("d", (3, 4, 5, 7, 8, 9, 16, 25))
def test_rotate_and_shift_state(pkg, d):
    """rotate_state / shift_state host mirrors (src/util.py:87-102) == numpy restatement of the reference's
    np.rot90 / np.roll lines, every (row, col) of the lattice for the shift."""
    from oracle import toric_oracle as O
    rng = np.random.default_rng(d)
    for _ in range(3):
        st = (rng.random((2, d, d)) < 0.3).astype(np.int64)
        prev = (rng.random((2, d, d)) < 0.3).astype(np.int64)
        want = np.stack((np.roll(np.rot90(st[0]), 1, axis=0), np.rot90(st[1])))
        got = pkg.rotate_state(st)
        assert got.dtype == np.int64 and (got == want).all() and (got == O.rotate_state(st)).all()
        g = d // 2
        for row, col in ((0, 0), (d - 1, d - 1), (g, g), (int(rng.integers(d)), int(rng.integers(d)))):
            a, b = pkg.shift_state(row, col, prev, st, g)
            wa = np.roll(np.roll(prev, g - row, axis=1), g - col, axis=2)
            wb = np.roll(np.roll(st, g - row, axis=1), g - col, axis=2)
            assert (a == wa).all() and (b == wb).all()


def test_make_mirrors_gym_make(pkg):
    env = pkg.make("toric-code-v0", config={"size": 5, "p_error": 0.1})
    assert env.system_size == 5 and env.action_space.high[-1] == 3
    s = env.reset()
    assert s.shape == (2, 5, 5) and s.any()
    with pytest.raises(ValueError):
        pkg.make("cartpole")


@pytest.mark.parametrize("d", (5, 7))
def test_single_env_and_parallel_spellings(pkg, pol, d):
    """src/util_actor.py's other spellings of the same operations -- selectActionParallel (:132-175),
    generatePerspectiveParallel (:178-186), generateTransition (:67-103), computePriorities (:105-124) -- against the
    reference-made fixtures and the numpy oracle."""
    states = pol["sel_states%d" % d]
    q = torch.from_numpy(pol["sel_q%d" % d]).cuda()
    actions, q_v = pkg.selectActionParallel(3, 0.0, d // 2, d, states, TableModel(q), "cuda")
    assert (actions == pol["sel_actions%d" % d]).all() and (q_v == pol["sel_qv%d" % d]).all()
    per, pos = pkg.generatePerspectiveParallel(d // 2, d, states[:6])
    for e in range(6):
        ov, op_ = O.perspectives(states[e])
        assert len(per[e]) == len(ov) == len(pos[e])
        assert all((a == b).all() for a, b in zip(per[e], ov)) and [tuple(x) for x in op_] == pos[e]
    st, ns, act = pol["tr_state%d" % d], pol["tr_next%d" % d], pol["tr_action%d" % d]
    trans, qs, qns = [], [], []
    rng = np.random.default_rng(d)
    for k in range(8):
        t = pkg.generateTransition(act[k], float(pol["tr_reward%d" % d][k]), d // 2, st[k], ns[k],
                                   bool(pol["tr_terminal%d" % d][k]))
        assert isinstance(t, pkg.Transition) and isinstance(t.action, pkg.Action)
        assert (t.state == pol["tr_out_prev%d" % d][k]).all() and (t.next_state == pol["tr_out_next%d" % d][k]).all()
        assert tuple(t.action.position) == tuple(int(x) for x in pol["tr_out_pos%d" % d][k])
        assert t.action.action == int(pol["tr_out_op%d" % d][k]) and t.reward == float(pol["tr_reward%d" % d][k])
        trans.append(t)
        qs.append(rng.normal(size=3).astype(np.float32) * 50)
        qns.append(rng.normal(size=3).astype(np.float32) * 50)
    pr = pkg.computePriorities(trans, qs, qns, 0.95)
    A = np.array([list(t.action.position) + [t.action.action] for t in trans])
    want = O.td_priorities(A, [t.reward for t in trans], np.array(qs), np.array(qns), 0.95)
    assert pr.shape == (8,) and np.allclose(pr, want, rtol=1e-6, atol=1e-5)
