"""GPU (BASELINE.json configs[0]): greedy rollouts through the CUDA engine + PyTorch Q-network on cuda:0 replay
the reference's CPU rollouts (golden_rollout_d*.npz: NN_11 + selectAction + ToricCode.step): same greedy
actions, Q-values within 1e-5 relative (fp32, TF32 off), same step counts, terminal and ground-state flags.
Also the batched evaluate() / prediction_smart() mirrors."""
import os

import numpy as np
import pytest
import torch

import golden as G
from oracle import toric_oracle as O

pytestmark = pytest.mark.gpu
WEIGHTS = {5: "Size_5_NN_11_state.pt", 7: "Size_7_NN_11_random_state.pt"}
RTOL = 1e-5


@pytest.fixture(scope="module")
def pkg(cuda_device):
    import toric_rl_decoder_b200 as pkg
    from toric_rl_decoder_b200 import networks
    networks.strict_fp32()
    return pkg


def load_model(d):
    from toric_rl_decoder_b200 import networks
    return networks.load_state(networks.NN_11(d, 3), os.path.join(G.GOLDEN_DIR, WEIGHTS[d]), device="cuda")


@pytest.mark.parametrize("d", (5, 7))
def test_batched_greedy_rollout_matches_reference(pkg, d):
    g = G.load("golden_rollout_d%d.npz" % d)
    model = load_model(d)
    n = len(g["qubits0"])
    eng = pkg.ToricEngine(d, n, max_steps=75)
    eng.set_qubits(g["qubits0"])
    q_sel = torch.empty((n, 3), dtype=torch.float32, device="cuda")
    eng.perspectives()
    T = int(g["n_steps"].max())
    near_ties = 0
    for t in range(T):
        total = eng.total_perspectives()
        with torch.no_grad():
            q = model(eng.out[:total]).contiguous()
        pkg.select_action_device(q, eng.pos, eng.offsets, n, actions=eng.actions, q_sel=q_sel)
        acts = eng.actions.cpu().numpy()
        qv = q_sel.cpu().numpy()
        qh, off = q.cpu().numpy(), eng.offsets.cpu().numpy()
        fix = False
        for e in range(n):
            if t < g["n_steps"][e]:
                ref = g["q_values"][e, t]
                tol = RTOL * np.abs(ref).max()
                if not (acts[e] == g["actions"][e, t]).all():
                    # allowed only when the top-2 gap is inside the tolerance (north_star: "within 1e-5 relative");
                    # the reference's action is then forced so that the trajectories stay comparable
                    seg = np.sort(qh[off[e]:off[e + 1]].reshape(-1))
                    assert seg[-1] - seg[-2] <= 2 * tol, (e, t, acts[e], g["actions"][e, t], seg[-3:])
                    near_ties += 1
                    acts[e] = g["actions"][e, t]
                    fix = True
                    continue
                assert np.allclose(qv[e], ref, rtol=RTOL, atol=tol), (e, t, qv[e], ref)
            else:
                assert (acts[e] == 0).all()            # solved env: identity action, frozen
        if fix:
            eng.actions.copy_(torch.from_numpy(acts))
        eng.actor_step(actions=eng.actions, auto_reset=False, freeze_done=True)
    assert near_ties <= 2, near_ties
    assert (eng.steps.cpu().numpy() == g["n_steps"]).all()
    assert (eng.get_qubits().cpu().numpy() == g["qubits_final"]).all()
    assert (eng.get_state().cpu().numpy().reshape(n, -1).sum(1) == 0).tolist() == g["terminal"].tolist()
    eng.gs.fill_(1)
    assert (eng.ground_state().cpu().numpy().astype(bool) == g["ground_state"]).all()


def test_evaluate_mirror_small(pkg):
    """evaluate(): same return structure as src/evaluation.py:10-124; rates consistent with a per-episode
    replay of the same device-sampled episodes through the numpy oracle + the same network."""
    from toric_rl_decoder_b200 import evaluation
    d, n = 5, 192
    model = load_model(d)
    cfg = {"size": d, "min_qubit_errors": 0, "p_error": 0.1}
    out = evaluation.evaluate(model, 'toric-code-v0', cfg, d // 2, 'cuda', [0.1, 0.05], num_of_episodes=n,
                              epsilon=0.0, num_of_steps=75, seed=3)
    ec, gs, steps, mq, failed = out
    assert ec.shape == gs.shape == steps.shape == mq.shape == (2,)
    assert (ec > 0.95).all() and (gs > 0.8).all() and (gs <= 1).all()
    assert (steps > 1).all() and (steps < 20).all() and (mq > 50).all()
    assert len(failed) % 2 == 0 and all(f.shape == (2, d, d) for f in failed)
    # replay p=0.1 through the oracle: the engine's episode e starts from the sampler state (seed, env e, episode 0)
    eng = pkg.ToricEngine(d, n, seed=3, env_base=0, p_error=0.1)
    eng.reset()
    Q0 = eng.get_qubits().cpu().numpy().astype(np.int64)
    kept = solved = nsteps = 0
    model_cpu = load_model(d).cpu()
    for e in range(n):
        q = Q0[e]
        s = O.syndrome(q)
        term, t = False, 0
        while not term and t < 75:
            views, pos = O.perspectives(s)
            with torch.no_grad():
                qt = model_cpu(torch.from_numpy(np.asarray(views, np.float32))).numpy()
            acts, _, _ = O.select_action_prime(qt, np.array([0, len(pos)]), pos, [True])
            q, s, _, term = O.step(q, s, acts[0])
            t += 1
        solved += term
        kept += O.ground_state(q)
        nsteps += t
    assert abs(ec[0] - solved / n) < 1e-12 and abs(gs[0] - kept / n) <= 1.0 / n + 1e-12
    assert abs(steps[0] - round(nsteps / n, 1)) <= 0.1 + 1e-9


@pytest.mark.parametrize("d,n", ((5, 1), (7, 1), (5, 24)))
def test_graphed_decoder_equals_eager(pkg, d, n):
    """evaluate() on launch-bound small batches replays one CUDA graph per decoding step
    (evaluation.GraphedGreedyDecoder); results must equal the eager loop's on the same seeded episodes.  The
    decoder (engine + graph) is built once per model and shape and reused across p_error values and seeds."""
    from toric_rl_decoder_b200 import evaluation
    model = load_model(d)
    cfg = {"size": d, "min_qubit_errors": 0, "p_error": 0.1}
    ps = [0.1, 0.05]
    flipped = 0
    seeds = 6 if n == 1 else 2
    for seed in range(seeds):
        a = evaluation.evaluate(model, 'toric-code-v0', cfg, d // 2, 'cuda', ps, num_of_episodes=n, num_of_steps=75,
                                seed=seed, use_graph=False)
        b = evaluation.evaluate(model, 'toric-code-v0', cfg, d // 2, 'cuda', ps, num_of_episodes=n, num_of_steps=75,
                                seed=seed, use_graph=True)
        same = all((x == y).all() for x, y in zip(a[:3], b[:3])) and len(a[4]) == len(b[4]) and \
            all((u == v).all() for u, v in zip(a[4], b[4]))
        if d == 5 and n == 1:
            # both paths run the network on the same 64-row batch (50 rows padded) -> bit-identical Q-values
            assert same, (seed, a[:4], b[:4])
        elif not same:
            # other shapes: the eager loop pads sumP to a multiple of 64, the graph runs all `cap` rows; cuDNN may
            # pick different kernels per batch size (Q differs in the last bits), so a near-tie may flip
            flipped += 1
            continue
        assert np.allclose(a[3], b[3], rtol=1e-4), (a[3], b[3])
    assert flipped <= 1, flipped
    assert len(evaluation.decoders_of(model)) == 1
    import copy
    copy.deepcopy(model)                                # the cache lives outside the module: still copyable
    # a batch past the launch-bound range keeps the eager loop
    assert evaluation.graphed_decoder(model, d, evaluation.GRAPH_MAX_ROWS, 'cuda', 75) is None


def test_prediction_smart_mirror_small(pkg, tmp_path):
    from toric_rl_decoder_b200 import evaluation
    d, n = 5, 256
    model = load_model(d)
    cfg = {"size": d, "min_qubit_errors": 0, "p_error": 0.1}
    q = d // 2 + 1
    res = evaluation.prediction_smart(model, 'toric-code-v0', cfg, d // 2, 'cuda', [0.01], num_of_episodes=n,
                                      num_of_steps=75, nbr_of_qubit_errors=q, seed=1)
    ec, gs, steps, mq, hist, n_fail, p_l, failed = res
    assert hist.shape == (3, d * d) and (hist[0] == np.arange(d * d)).all()
    assert hist[1:].sum() == n and hist[1:, :q].sum() == 0          # every episode has >= q errors
    assert hist[2].sum() == round((1 - gs[0]) * n) == len(failed)
    assert 0 <= p_l[0] < 1e-3
    path = tmp_path / "eval.txt"
    evaluation.write_evaluation_file(str(path), d, [0.01], gs, ec, mq, p_l, steps)
    lines = open(path).read().strip().split("\n")
    assert lines[0] == "# " + evaluation.RESULT_HEADER and len(lines[1].split(",")) == 9
