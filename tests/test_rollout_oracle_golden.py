"""CPU: the PyTorch Q-network restatement (toric_rl_decoder_b200/networks.py) and the numpy oracle replay the
reference's greedy rollouts (tests/golden/golden_rollout_d*.npz, made by oracle/gen_golden_rollout.py with the
reference's NN_11 + selectAction + ToricCode on CPU): same actions, Q-values within 1e-5 relative."""
import os

import numpy as np
import pytest
import torch

import golden as G
from oracle import toric_oracle as O

WEIGHTS = {5: "Size_5_NN_11_state.pt", 7: "Size_7_NN_11_random_state.pt"}
RTOL = 1e-5


def load_model(d, device="cpu"):
    from toric_rl_decoder_b200 import networks
    m = networks.NN_11(d, 3)
    return networks.load_state(m, os.path.join(G.GOLDEN_DIR, WEIGHTS[d]), device=device)


@pytest.mark.parametrize("d", (5, 7))
def test_network_matches_reference_forward(d):
    g = G.load("golden_rollout_d%d.npz" % d)
    model = load_model(d)
    s0 = O.syndrome(g["qubits0"][0].astype(np.int64))
    views, _ = O.perspectives(s0)
    with torch.no_grad():
        q = model(torch.from_numpy(np.asarray(views, np.float32))).numpy()
    ref = g["first_q"]
    assert q.shape == ref.shape
    assert np.allclose(q, ref, rtol=RTOL, atol=RTOL * np.abs(ref).max())


@pytest.mark.parametrize("d,episodes", ((5, 32), (7, 4)))
def test_oracle_replays_reference_rollouts(d, episodes):
    g = G.load("golden_rollout_d%d.npz" % d)
    model = load_model(d)
    torch.set_num_threads(4)
    for j in range(episodes):
        q = g["qubits0"][j].astype(np.int64)
        s = O.syndrome(q)
        for t in range(int(g["n_steps"][j])):
            views, pos = O.perspectives(s)
            with torch.no_grad():
                qt = model(torch.from_numpy(np.asarray(views, np.float32))).numpy()
            acts, qv, _ = O.select_action_prime(qt, np.array([0, len(pos)]), pos, [True])
            assert (acts[0] == g["actions"][j, t]).all(), (j, t)
            ref = g["q_values"][j, t]
            assert np.allclose(qv[0], ref, rtol=RTOL, atol=RTOL * np.abs(ref).max())
            q, s, _, term = O.step(q, s, acts[0])
        assert term == bool(g["terminal"][j]) and (q == g["qubits_final"][j]).all()
        assert O.ground_state(q) == bool(g["ground_state"][j])
