"""GPU: bounded versions of the decoder-quality sweeps, so that the numbers that pin the evaluation path end to end
(sampler -> perspectives -> Q-network -> select-action epilogue -> step -> logical-failure check -> estimator) are
re-checked by every `-m gpu` run, not only by scripts/c4_big.py / scripts/c_depolarizing_sweep.py.

References (2*10^6 and 10^4 episodes per point, converged NN_11 weights of the checkout):
  results/evaluation_size_5.txt:3    d=5, p=0.005, generateNPlusQRandomErrors(q=3): ground-state rate 0.996427,
                                     mean steps 3.2 (results/small_p_error_test.py)
  data/evaluation_size_7_p_0.005.txt:2 / results/evaluation_size_7.txt   d=7, p=0.005, q=4: 0.9998975 / 0.9998785
  runs/test_size_5_steps_10000       d=5, plain depolarizing p=0.10, greedy, 75-step cap: ground-state rate 0.9043
                                     (tests/evaluate_network.py protocol; results/results_mats/RL_5.txt's 0.9286 at the
                                     same p is the paper's figure data from another checkpoint and is NOT the target)
Two-sample z-scores, |z| <= 4."""
import math
import os

import pytest
import torch

import golden as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tools(cuda_device):
    from toric_rl_decoder_b200 import evaluation, networks
    torch.backends.cudnn.allow_tf32 = True           # bulk statistics: TF32 convolutions (labelled; parity tests use fp32)
    torch.backends.cuda.matmul.allow_tf32 = True
    yield evaluation, networks
    networks.strict_fp32()


def _model(networks, d, name):
    m = networks.load_state(networks.NN_11(d, 3), os.path.join(G.GOLDEN_DIR, name), device="cuda")
    return networks.pad_channels(m, 8).to(memory_format=torch.channels_last)


def _z(f, n, f_ref, n_ref):
    pooled = (f * n + f_ref * n_ref) / (n + n_ref)
    return (f - f_ref) / math.sqrt(max(pooled * (1 - pooled), 1e-30) * (1.0 / n + 1.0 / n_ref))


def test_importance_sampled_sweep_point_d5(tools):
    evaluation, networks = tools
    d, p, n = 5, 0.005, 1 << 17
    model = _model(networks, d, "Size_5_NN_11_converged_state.pt")
    ec, gs, st, mq, hist, nfail, pl, _ = evaluation.prediction_smart(
        model, "toric-code-v0", {"size": d, "p_error": 0.1}, d // 2, "cuda", [p], num_of_episodes=n, num_of_steps=75,
        nbr_of_qubit_errors=d // 2 + 1, seed=5, keep_failed=False)
    f = 1.0 - float(gs[0])
    assert abs(_z(f, n, 1.0 - 0.996427, 2000000)) <= 4.0, f
    assert abs(float(st[0]) - 3.2) < 0.1
    assert float(ec[0]) > 0.999
    assert 1e-6 < float(pl[0]) < 1e-5                 # reference: 3.159e-6


def test_importance_sampled_sweep_point_d7(tools):
    evaluation, networks = tools
    d, p, n = 7, 0.005, 1 << 17
    model = _model(networks, d, "Size_7_NN_11_random_state.pt")
    ec, gs, st, mq, hist, nfail, pl, _ = evaluation.prediction_smart(
        model, "toric-code-v0", {"size": d, "p_error": 0.1}, d // 2, "cuda", [p], num_of_episodes=n, num_of_steps=75,
        nbr_of_qubit_errors=d // 2 + 1, seed=6, keep_failed=False)
    f = 1.0 - float(gs[0])
    # the reference's own two runs: 1.025e-4 and 1.215e-4 (they differ by 1.8 sigma); pooled 1.12e-4 over 4e6 episodes
    assert abs(_z(f, n, 1.12e-4, 4000000)) <= 4.0, f
    assert abs(float(st[0]) - 4.5) < 0.1


def test_depolarizing_point_d5(tools):
    evaluation, networks = tools
    d, p, n = 5, 0.10, 1 << 15
    model = _model(networks, d, "Size_5_NN_11_converged_state.pt")
    ec, gs, st, mq, _ = evaluation.evaluate(model, "toric-code-v0", {"size": d, "p_error": p}, d // 2, "cuda", [p],
                                            num_of_episodes=n, num_of_steps=75, seed=7, keep_failed=False)
    assert abs(_z(float(gs[0]), n, 0.9043, 10000)) <= 4.0, float(gs[0])
    assert float(ec[0]) > 0.998
