"""Golden greedy rollouts (BASELINE.json configs[0]) produced by the UNMODIFIED reference on CPU.

Run in the build container only (needs /root/reference):

    python oracle/gen_golden_rollout.py

For each (d, checkpoint) the reference's own classes are used end to end: the network is
src/nn/torch/NN.py:NN_11 with the reference checkpoint, actions come from src/util_actor.py:selectAction
(epsilon = 0) on gym_ToricCode's ToricCode, episodes start from env.reset(p_error) under a fixed numpy
seed and run until terminal or 75 steps (the reference scripts' num_of_steps).

Writes
  tests/golden/Size_5_NN_11_state.pt        state_dict extracted from network/Size_5_NN_11.pt (the legacy
                                            full-module pickle BASELINE.json names), float32, unchanged values
  tests/golden/Size_7_NN_11_random_state.pt copy of the state_dict network/converged/Size_7_NN_11_random_*.pt
  tests/golden/golden_rollout_d{5,7}.npz    initial qubit matrices, per-step actions and Q-value rows,
                                            step counts, terminal / ground-state flags, and the reference
                                            network's Q-values on the first state's perspectives
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
MAX_STEPS = 75


def load_legacy_pickle(path):
    """network/Size_5_NN_11.pt pickles the whole module as NN.NN_11 with the removed thnn backend."""
    import torch
    import src.nn.torch.NN as NN
    sys.modules.setdefault("NN", NN)
    if not hasattr(torch.nn, "backends"):
        torch.nn.backends = types.ModuleType("torch.nn.backends")
        sys.modules["torch.nn.backends"] = torch.nn.backends
    thnn = types.ModuleType("torch.nn.backends.thnn")
    thnn._get_thnn_function_backend = lambda: None
    torch.nn.backends.thnn = thnn
    sys.modules["torch.nn.backends.thnn"] = thnn
    return torch.load(path, map_location="cpu", weights_only=False)


def rollouts(ref, model, d, p_error, episodes, seed):
    import torch
    torch.set_num_threads(4)
    np.random.seed(seed)
    import random
    random.seed(seed)
    env = ref.gym.make("toric-code-v0", config={"size": d, "min_qbit_errors": 0, "p_error": p_error})
    g = int(d / 2)
    Q0 = np.zeros((episodes, 2, d, d), np.uint8)
    A = np.full((episodes, MAX_STEPS, 4), -1, np.int64)
    QV = np.zeros((episodes, MAX_STEPS, 3), np.float32)
    NS = np.zeros(episodes, np.int64)
    TERM = np.zeros(episodes, bool)
    GS = np.zeros(episodes, bool)
    QF = np.zeros((episodes, 2, d, d), np.uint8)
    first_q = None
    for j in range(episodes):
        state = env.reset(p_error=p_error)
        Q0[j] = env.qubit_matrix
        if j == 0:
            per, _ = ref.util.generatePerspectiveOptimized(g, d, state)
            with torch.no_grad():
                first_q = model(torch.from_numpy(np.array(per)).type('torch.Tensor')).numpy()
        terminal, t = False, 0
        while not terminal and t < MAX_STEPS:
            action, q_values = ref.util_actor.selectAction(number_of_actions=3, epsilon=0.0, grid_shift=g,
                                                           toric_size=d, state=state, model=model, device='cpu')
            A[j, t] = action
            QV[j, t] = q_values
            state, _, terminal, _ = env.step(action)
            t += 1
        NS[j], TERM[j], GS[j], QF[j] = t, terminal, env.evalGroundState(), env.qubit_matrix
        print("d=%d episode %d: %d steps terminal=%s gs=%s" % (d, j, t, terminal, GS[j]))
    return {"qubits0": Q0, "actions": A, "q_values": QV, "n_steps": NS, "terminal": TERM, "ground_state": GS,
            "qubits_final": QF, "first_q": first_q, "p_error": p_error, "d": d}


def main():
    import torch
    os.makedirs(OUT, exist_ok=True)
    ref = ref_loader.load()
    import src.nn.torch.NN as NN
    net = ref_loader.REF_ROOT + "/network/"

    m5 = load_legacy_pickle(net + "Size_5_NN_11.pt").eval()
    sd5 = {k: v.detach().float().clone() for k, v in m5.state_dict().items()}
    torch.save(sd5, os.path.join(OUT, "Size_5_NN_11_state.pt"))
    np.savez_compressed(os.path.join(OUT, "golden_rollout_d5.npz"), **rollouts(ref, m5, 5, 0.1, 32, seed=5))

    sd7 = torch.load(net + "converged/Size_7_NN_11_random_18_Mar_2020_18_17_52.pt", map_location="cpu",
                     weights_only=True)
    m7 = NN.NN_11(7, 3, 'cpu')
    m7.load_state_dict(sd7)
    m7.eval()
    torch.save({k: v.detach().float().clone() for k, v in sd7.items()},
               os.path.join(OUT, "Size_7_NN_11_random_state.pt"))
    np.savez_compressed(os.path.join(OUT, "golden_rollout_d7.npz"), **rollouts(ref, m7, 7, 0.1, 12, seed=7))


def export_converged_d5():
    """tests/golden/Size_5_NN_11_converged_state.pt: copy of the state_dict
    network/converged/Size_5_NN_11_17_Mar_2020_22_33_59.pt, the network behind results/evaluation_size_5.txt
    (results/small_p_error_test.py:12) -- used by scripts/c4_big.py --d 5 for the statistical check on a second
    lattice size.  Does not touch the rollout fixtures."""
    import torch
    sd = torch.load(ref_loader.REF_ROOT + "/network/converged/Size_5_NN_11_17_Mar_2020_22_33_59.pt",
                    map_location="cpu", weights_only=True)
    torch.save({k: v.detach().float().clone() for k, v in sd.items()},
               os.path.join(OUT, "Size_5_NN_11_converged_state.pt"))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "converged5":
        export_converged_d5()
    else:
        main()
