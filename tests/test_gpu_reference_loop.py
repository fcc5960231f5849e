"""The reference actor's loop body (src/Actor_mp.py:104-183) run through the mirrors that carry the reference's names
-- generatePerspectiveBatch, EnvSet.step, generateTransitionParallel, resetTerminalEnvs, `next_state[idx] =
reset_states` -- in both host contracts (numpy_compat=True: the reference's numpy dtypes; False: device tensors).
Every step of every env is checked against the oracle's restatement of ToricCode.step, every reset against the
oracle's syndrome of the qubits the engine drew, and the two contracts against each other."""
import numpy as np
import pytest
import torch

from oracle import toric_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg(cuda_device):
    import toric_rl_decoder_b200 as pkg
    return pkg


def _loop(pkg, numpy_compat, d=5, n=48, steps=30, p=0.12, seed=17, max_actions=9):
    g = d // 2
    es = pkg.EnvSetB200({"size": d, "p_error": p}, n, seed=seed, numpy_compat=numpy_compat)
    trans_type = pkg.transition_type(d) if numpy_compat else None
    state = es.resetAll()
    steps_per_episode = np.zeros(n, dtype=np.int64) if numpy_compat else torch.zeros(n, dtype=torch.int64, device="cuda")
    qubits = es.engine.get_qubits().cpu().numpy().astype(np.int64)
    log = []
    for k in range(steps):
        steps_per_episode += 1
        persp, pos, counts, offsets = pkg.generatePerspectiveBatch(g, d, state, raw_positions=True)
        q = torch.zeros((persp.shape[0], 3), dtype=torch.float32, device="cuda")
        step_no = torch.full((n,), k, dtype=torch.int32, device="cuda")
        act, _ = pkg.select_action_device(q, pos, offsets, n, epsilon=1.0, steps=step_no, seed=seed)
        action = act.cpu().numpy().astype(np.int64) if numpy_compat else act
        next_state, reward, terminal_state, _ = es.step(action)
        transition = pkg.generateTransitionParallel(action, reward, state, next_state, terminal_state, g, trans_type)

        # against the oracle, env by env
        s_np = np.asarray(state if numpy_compat else state.cpu().numpy()).astype(np.int64)
        ns_np = np.asarray(next_state if numpy_compat else next_state.cpu().numpy()).astype(np.int64)
        a_np = np.asarray(action if numpy_compat else action.cpu().numpy()).astype(np.int64)
        r_np = np.asarray(reward if numpy_compat else reward.cpu().numpy()).astype(np.float64)
        t_np = np.asarray(terminal_state if numpy_compat else terminal_state.cpu().numpy()).astype(bool)
        for i in range(n):
            if s_np[i].any():
                assert a_np[i, 3] in (1, 2, 3)                       # an operator on a qubit next to a defect
            qubits[i], ns, r, t = O.step(qubits[i], s_np[i], list(a_np[i]))
            assert (ns_np[i] == ns).all() and r_np[i] == r and t_np[i] == t, (k, i)
        if numpy_compat:                                            # the record array the reference's actor stores
            assert transition.dtype == trans_type and transition.shape == (n,)
            assert (transition["reward"] == r_np).all() and (transition["terminal"] == t_np).all()
            assert (transition["action"]["op"] == a_np[:, 3]).all()
            assert (transition["action"]["position"][:, 1:] == g).all()
        log.append((a_np.copy(), ns_np.copy(), r_np.copy(), t_np.copy()))

        too_many_steps = steps_per_episode > max_actions
        if numpy_compat:
            idx = np.argwhere(np.logical_or(terminal_state, too_many_steps)).flatten()
        else:
            idx = torch.nonzero(terminal_state | too_many_steps).flatten()
        if len(idx):
            p_errors = np.full(len(idx), p) if numpy_compat else None
            reset_states = es.resetTerminalEnvs(idx, p_errors=p_errors)
            next_state[idx] = reset_states                          # src/Actor_mp.py:182
            steps_per_episode[idx] = 0
            ii = np.asarray(idx if numpy_compat else idx.cpu().numpy())
            fresh = es.engine.get_qubits().cpu().numpy().astype(np.int64)
            rs = np.asarray(reset_states if numpy_compat else reset_states.cpu().numpy())
            for j, i in enumerate(ii):
                qubits[i] = fresh[i]
                assert (O.syndrome(fresh[i]) == rs[j]).all() and rs[j].any()
            if numpy_compat:
                assert reset_states.dtype == np.float64 and next_state.dtype == np.int64
            log.append(ii.copy())
        state = next_state
    assert sum(1 for x in log if not isinstance(x, tuple)) >= 3     # the run did see resets
    return log


def test_reference_loop_both_contracts(pkg):
    a = _loop(pkg, True)
    b = _loop(pkg, False)
    assert len(a) == len(b)
    for x, y in zip(a, b):
        if isinstance(x, tuple):
            assert all((u == v).all() for u, v in zip(x, y))
        else:
            assert (x == y).all()
