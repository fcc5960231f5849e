"""Batched greedy evaluation on the GPU: src/evaluation.py:10-124 `evaluate` and
results/small_p_error_test.py:56-191 `prediction_smart` (importance-sampled low-p statistics), with the
reference's argument names, return tuples and result-file columns
(data/evaluation_size_7_p_*.txt:1).

The reference runs one episode at a time; here all episodes of one p_error run as one engine batch
(chunked to fit HBM): perspectives -> Q-network (PyTorch) -> tc_select_action -> tc_actor_step with
TC_ACTOR_FREEZE_DONE, until every episode is solved or num_of_steps is reached.  Episodes are independent,
so with torch.distributed initialised the episode range is sharded over the ranks and only the sums are
all-reduced.
"""
import math
import weakref

import numpy as np
import torch

from . import _lib
from .engine import (SAMPLE_DEPOLARIZING, SAMPLE_FIXED_PLUS_P, SAMPLE_FIXED_WEIGHT, ToricEngine, _ptr, _stream)
from .policy import _forward, select_action_device
from .sharding import shard_range


def default_chunk(d, budget_bytes=12 << 30):
    """Episodes per engine batch so that the worst-case perspective buffer (n * 2d^2 views of 8d^2 bytes)
    stays within budget_bytes."""
    per_env = (2 * d * d) * (8 * d * d + 4)
    return max(1, min(1 << 18, budget_bytes // per_env))


def _unpack_qubits(eng, packed):
    out = torch.empty((packed.shape[0], 2, eng.d, eng.d), dtype=torch.uint8, device=eng.device)
    if packed.shape[0]:
        _lib.check(eng.lib.tc_unpack_qubits(_ptr(packed.contiguous()), _ptr(out), eng.d, packed.shape[0], _stream()),
                   "tc_unpack_qubits")
    return out


def run_episodes(model, eng, num_of_steps, epsilon=0.0, fwd_chunk=1 << 15):
    """Drive every env of `eng` (already reset or injected) greedily until solved or num_of_steps.
    Returns (q_sum float64, q_count int) of the chosen actions' Q-values; per-env results stay in the
    engine (steps, terminal via the syndrome, qubits)."""
    n = eng.n
    q_sum = torch.zeros((), dtype=torch.float64, device=eng.device)
    q_cnt = 0
    q_sel = torch.empty((n, 3), dtype=torch.float32, device=eng.device)
    eng.perspectives()
    for _ in range(int(num_of_steps)):
        total = eng.total_perspectives()          # host read: the network input is sliced to sumP
        if int(eng.overflow.item()) != 0:
            raise RuntimeError("perspective buffer overflow")
        if total == 0:
            break
        q = _forward(model, eng.out[:total], chunk=fwd_chunk)
        select_action_device(q, eng.pos, eng.offsets, n, epsilon=epsilon, episode=eng.episode, steps=eng.steps,
                             seed=eng.seed, env_base=eng.env_base, actions=eng.actions, q_sel=q_sel)
        active = eng.counts > 0
        chosen = q_sel.gather(1, (eng.actions[:, 3].long() - 1).clamp_(0, 2).unsqueeze(1)).squeeze(1)
        q_sum += (chosen.double() * active).sum()
        q_cnt += int(active.sum().item())
        eng.actor_step(actions=eng.actions, auto_reset=False, freeze_done=True)
    return q_sum, q_cnt


GRAPH_MAX_ROWS = 4096     # worst-case perspective rows (n * 2d^2) up to which a decoding step is launch-bound
GRAPH_CACHE_SIZE = 8      # graphed decoders kept per model (one per lattice size / batch size / step cap)


class GraphedGreedyDecoder:
    """One whole greedy decoding step -- Q-network forward over the full fixed-capacity perspective buffer,
    tc_select_action, Q bookkeeping, tc_actor_step(FREEZE_DONE) -- captured as ONE CUDA graph.

    For the reference's own use of the path (src/evaluation.py:62-117: one episode at a time, BASELINE
    configs[0]) a step is ~45 kernel launches for a few microseconds of work each (12 convolutions + padding +
    activations of NN_11, plus the engine's kernels), and run_episodes() synchronises three times per step to
    slice the network input to sumP.  Here the network always runs over all `cap` rows (rows past sumP are
    never read: tc_select_action walks `offsets`), so every shape is static, the step is a single graph launch
    and the only host read is a 4-byte "unsolved envs" counter.  Needs a network whose rows are independent
    (`pad_small_batches`, true of the reference's conv decoders) and epsilon = 0 (the epsilon draw's Philox key
    would be baked into the graph).  The engine is reused across run() calls; the caller reseeds it through
    `reseed()` so that its sampler state equals a freshly built engine's."""

    def __init__(self, model, d, n, device, num_of_steps):
        self.n = int(n)
        self._model = model                        # only until the capture below: the graph holds the kernels
        self.eng = eng = ToricEngine(d, n, device=device, max_steps=num_of_steps)
        cap = (eng.worst_case_capacity() + 63) // 64 * 64
        eng.ensure_capacity(cap)
        eng.out.zero_()
        dev = eng.device
        self.q_sel = torch.zeros((self.n, 3), dtype=torch.float32, device=dev)
        self.q_sum = torch.zeros((), dtype=torch.float64, device=dev)
        self.q_cnt = torch.zeros((), dtype=torch.int64, device=dev)
        self.unsolved = torch.zeros((), dtype=torch.int64, device=dev)
        self.param_ptr = next(model.parameters()).data_ptr()
        eng.perspectives()                         # trivial syndromes: every env is "solved", the body is a no-op
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):              # cuDNN picks its kernels outside the capture
            for _ in range(3):
                self._body()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._body()
        self._model = None                         # no strong reference back to the cache key (see _DECODERS)

    def _body(self):
        eng, n = self.eng, self.n
        with torch.no_grad():
            q = self._model(eng.out).float().contiguous()
        select_action_device(q, eng.pos, eng.offsets, n, actions=eng.actions, q_sel=self.q_sel)
        active = eng.counts > 0
        chosen = self.q_sel.gather(1, (eng.actions[:, 3].long() - 1).clamp_(0, 2).unsqueeze(1)).squeeze(1)
        self.q_sum += (chosen.double() * active).sum()
        self.q_cnt += active.sum()
        eng.actor_step(actions=eng.actions, auto_reset=False, freeze_done=True)
        self.unsolved.copy_((eng.counts > 0).sum())

    def reseed(self, seed, env_base, p_error):
        """Put the engine in the state of ToricEngine(d, n, seed=seed, env_base=env_base, p_error=p_error)."""
        eng = self.eng
        eng.seed, eng.env_base = int(seed), int(env_base)
        eng.set_p_error(p_error)
        eng.episode.zero_()
        eng.status.zero_()
        eng.counters.zero_()
        return eng

    def run(self, num_of_steps):
        """Same contract as run_episodes(model, eng, num_of_steps): the engine has been reset or injected."""
        eng = self.eng
        self.q_sum.zero_()
        self.q_cnt.zero_()
        eng.perspectives()
        for _ in range(int(num_of_steps)):
            self.graph.replay()
            if int(self.unsolved.item()) == 0:
                break
        if int(eng.overflow.item()) != 0:
            raise RuntimeError("perspective buffer overflow")
        return self.q_sum.clone(), int(self.q_cnt.item())


# model -> {shape key: GraphedGreedyDecoder}.  Kept outside the module object (copy.deepcopy(model) and state_dict
# stay untouched) and weakly keyed: the decoders go when the network does.
_DECODERS = weakref.WeakKeyDictionary()


def decoders_of(model):
    """The graphed decoders currently cached for `model` ({shape key: GraphedGreedyDecoder})."""
    return dict(_DECODERS.get(model, {}))


def graphed_decoder(model, d, n, device, num_of_steps, epsilon=0.0):
    """The cached GraphedGreedyDecoder of `model` for this shape, or None when the step is not launch-bound
    (more than GRAPH_MAX_ROWS rows), epsilon > 0, or the network does not declare independent rows."""
    if epsilon != 0.0 or n * 2 * d * d > GRAPH_MAX_ROWS or not getattr(model, "pad_small_batches", False):
        return None
    dev = torch.device(device)
    if dev.type == "cuda" and dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    key = (int(d), int(n), str(dev), int(num_of_steps), torch.backends.cudnn.allow_tf32,
           torch.backends.cuda.matmul.allow_tf32, torch.is_autocast_enabled())
    cache = _DECODERS.setdefault(model, {})
    dec = cache.get(key)
    if dec is None or dec.param_ptr != next(model.parameters()).data_ptr():
        cache.pop(key, None)
        while len(cache) >= GRAPH_CACHE_SIZE:      # each entry owns a graph memory pool: keep the most recent shapes
            cache.pop(next(iter(cache)))
        dec = cache[key] = GraphedGreedyDecoder(model, d, n, dev, num_of_steps)
    return dec


def _all_reduce_sums(vec):
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(vec, op=dist.ReduceOp.SUM)
    return vec


def _rank_world():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def _evaluate_one_p(model, d, device, p_error, num_of_episodes, epsilon, num_of_steps, mode, n_fixed, seed,
                    chunk, keep_failed, flips_hist=None, episode_base=0, use_graph=True):
    """All episodes of one p_error owned by this rank.  Returns the float64 sums
    [episodes, solved, ground_state_kept, steps, q_sum, q_count] (device) and the failed-syndrome list."""
    rank, world = _rank_world()
    base, count = shard_range(num_of_episodes, rank, world)
    sums = torch.zeros(6, dtype=torch.float64, device=device)
    failed = []
    chunk = int(chunk or default_chunk(d))
    done = 0
    while done < count:
        n = min(chunk, count - done)
        dec = graphed_decoder(model, d, n, device, num_of_steps, epsilon) if use_graph else None
        if dec is not None:       # launch-bound small batch (the reference's one-episode-at-a-time use)
            eng = dec.reseed(seed, episode_base + base + done, p_error)
        else:
            eng = ToricEngine(d, n, device=device, seed=seed, env_base=episode_base + base + done, p_error=p_error,
                              max_steps=num_of_steps)
        eng.reset(mode=mode, n_fixed=n_fixed)
        init = eng.qubits.clone()
        q_sum, q_cnt = dec.run(num_of_steps) if dec is not None else run_episodes(model, eng, num_of_steps, epsilon)
        solved = (eng.syn.reshape(n, -1) != 0).sum(dim=1) == 0
        eng.gs.fill_(1)
        gs = eng.ground_state().bool()
        sums += torch.stack([torch.tensor(float(n), dtype=torch.float64, device=device), solved.sum().double(),
                             gs.sum().double(), eng.steps.sum().double(), q_sum,
                             torch.tensor(float(q_cnt), dtype=torch.float64, device=device)])
        bad = (~solved) | (~gs)
        if flips_hist is not None:                     # prediction_smart's per-weight success / failure counts
            q0 = _unpack_qubits(eng, init)
            flips = (q0 != 0).reshape(n, -1).sum(dim=1)
            size = flips_hist.shape[1]
            ok_w = flips[gs & (flips < size)]
            bad_w = flips[(~gs) & (flips < size)]
            flips_hist[1] += torch.bincount(ok_w, minlength=size)[:size].double()
            flips_hist[2] += torch.bincount(bad_w, minlength=size)[:size].double()
            bad = ~gs
        if keep_failed:
            idx = torch.nonzero(bad).reshape(-1)
            if idx.numel():
                qi = _unpack_qubits(eng, init[idx]).cpu().numpy().astype(np.int64)
                qf = _unpack_qubits(eng, eng.qubits[idx]).cpu().numpy().astype(np.int64)
                if flips_hist is None:
                    for a, b in zip(qi, qf):           # evaluation.py:117-119: initial then final qubit matrix
                        failed.append(a)
                        failed.append(b)
                else:
                    failed.extend(list(qi))            # small_p_error_test.py:162
        done += n
        eng = None
    return sums, failed


def _config_fields(env_config):
    d = int(env_config["size"])
    return d


def evaluate(model, env, env_config, grid_shift, device, prediction_list_p_error, num_of_episodes=1,
             num_actions=3, epsilon=0.0, num_of_steps=50, plot_one_episode=False, minimum_nbr_of_qubit_errors=0,
             seed=0, chunk=None, keep_failed=True, use_graph=True):
    """evaluate (src/evaluation.py:10-124).  Returns (error_corrected_list, ground_state_list,
    average_number_of_steps_list, mean_q_list, failed_syndroms) like the reference.
    `env` (the gym id) and plot_one_episode are accepted for signature compatibility."""
    d = _config_fields(env_config)
    if int(grid_shift) != d // 2:
        raise ValueError("grid_shift must be size // 2")
    dev = torch.device(device)
    model.to(dev)
    model.eval()
    k = len(prediction_list_p_error)
    error_corrected_list, ground_state_list = np.zeros(k), np.zeros(k)
    average_number_of_steps_list, mean_q_list = np.zeros(k), np.zeros(k)
    failed_syndroms = []
    # the reference passes 'min_qubit_errors', which the env never reads (ToricCode.py:39 reads
    # 'min_qbit_errors'), so its evaluate always samples depolarizing errors; honour the explicit argument only
    mode = SAMPLE_FIXED_WEIGHT if int(env_config.get("min_qbit_errors", 0)) > 0 else SAMPLE_DEPOLARIZING
    n_fixed = int(env_config.get("min_qbit_errors", 0))
    for i, p_error in enumerate(prediction_list_p_error):
        sums, failed = _evaluate_one_p(model, d, dev, float(p_error), int(num_of_episodes), epsilon, num_of_steps,
                                       mode, n_fixed, seed, chunk, keep_failed, episode_base=i * int(num_of_episodes),
                                       use_graph=use_graph)
        s = _all_reduce_sums(sums).cpu().numpy()
        n_ep = max(s[0], 1.0)
        error_corrected_list[i] = s[1] / n_ep
        ground_state_list[i] = s[2] / n_ep
        average_number_of_steps_list[i] = np.round(s[3] / n_ep, 1)
        mean_q_list[i] = np.round(s[4] / max(s[5], 1.0), 3)
        failed_syndroms.extend(failed)
    return error_corrected_list, ground_state_list, average_number_of_steps_list, mean_q_list, failed_syndroms


def prediction_smart(model, env, env_config, grid_shift, device, prediction_list_p_error, num_of_episodes=1,
                     epsilon=0.0, num_of_steps=50, plot_one_episode=False, show_network=False, show_plot=False,
                     nbr_of_qubit_errors=0, print_Q_values=False, checkpoint=10000, seed=0, chunk=None,
                     keep_failed=True, use_graph=True):
    """prediction_smart (results/small_p_error_test.py:56-191): episodes start from
    generateNPlusQRandomErrors(q = nbr_of_qubit_errors, p) (fixed q errors plus depolarizing errors on the
    other qubits, resampled while the syndrome is trivial), failures are binned by initial error weight and
    turned into the logical failure estimate P_l (:169-181).  Returns the reference's 8-tuple."""
    d = _config_fields(env_config)
    dev = torch.device(device)
    model.to(dev)
    model.eval()
    k = len(prediction_list_p_error)
    size = d * d                                                     # max_number_of_errors (:76)
    error_corrected_list, ground_state_list = np.zeros(k), np.zeros(k)
    average_number_of_steps_list, mean_q_list, P_l_list = np.zeros(k), np.zeros(k), np.zeros(k)
    failed_syndromes = []
    hist = torch.zeros((3, size), dtype=torch.float64, device=dev)   # accumulates over p like the reference
    hist[0] = torch.arange(size, dtype=torch.float64, device=dev)
    N_fail = 0.0
    q = int(nbr_of_qubit_errors)
    mode = SAMPLE_FIXED_PLUS_P if q > 0 else SAMPLE_DEPOLARIZING
    for i, p_error in enumerate(prediction_list_p_error):
        local = torch.zeros_like(hist)
        sums, failed = _evaluate_one_p(model, d, dev, float(p_error), int(num_of_episodes), epsilon, num_of_steps,
                                       mode, q, seed, chunk, keep_failed, flips_hist=local,
                                       episode_base=i * int(num_of_episodes), use_graph=use_graph)
        s = _all_reduce_sums(sums).cpu().numpy()
        hist[1:] += _all_reduce_sums(local[1:].contiguous())
        n_ep = max(s[0], 1.0)
        fails = hist[2].cpu().numpy()
        n_fail = np.zeros(size)
        for kk in range(size):
            n_fail[kk] = 0 if kk < q else fails[kk] / math.comb(kk, q)
        N_fail = float(np.sum(n_fail))
        nq = 2 * d * d
        p_conf = float(p_error) ** q * (1 - float(p_error)) ** (nq - q)
        P_l_list[i] = math.comb(nq, q) * p_conf * N_fail / n_ep
        error_corrected_list[i] = s[1] / n_ep
        ground_state_list[i] = s[2] / n_ep
        average_number_of_steps_list[i] = np.round(s[3] / n_ep, 1)
        mean_q_list[i] = np.round(s[4] / max(s[5], 1.0), 3)
        failed_syndromes.extend(failed)
    return (error_corrected_list, ground_state_list, average_number_of_steps_list, mean_q_list, hist.cpu().numpy(),
            N_fail, P_l_list, failed_syndromes)


GROUND_STATE_FAILED_THEORY = {5: 0.000472411186696901, 7: 6.35851871947911e-06, 9: 7.268875712609422e-08}
GROUND_STATE_CONSERVED_THEORY = {5: 0.9995275888133031, 7: 0.9999936414812806, 9: 0.9999999273112429}
RESULT_HEADER = ("p_error, ground_state_list, error_corrected_list, mean_q_list, failure_rate, asymptotic_fail, "
                 "asymptotic_success, P_l, average_number_of_steps_list")


def write_evaluation_file(path, size, p_error, ground_state_list, error_corrected_list, mean_q_list, P_l,
                          average_number_of_steps_list):
    """The reference's result file (results/small_p_error_test.py:232-245; data/evaluation_size_7_p_*.txt):
    one comma-separated row per p_error under the same header."""
    gs = np.asarray(ground_state_list, dtype=np.float64)
    failure_rate = 1 - gs
    ft, ct = GROUND_STATE_FAILED_THEORY.get(int(size)), GROUND_STATE_CONSERVED_THEORY.get(int(size))
    asymptotic_fail = (failure_rate - ft) / ft * 100 if ft else np.full_like(gs, np.nan)
    asymptotic_success = (gs - ct) / ct * 100 if ct else np.full_like(gs, np.nan)
    data = np.array([np.asarray(p_error, dtype=np.float64), gs, np.asarray(error_corrected_list, dtype=np.float64),
                     np.asarray(mean_q_list, dtype=np.float64), failure_rate, asymptotic_fail, asymptotic_success,
                     np.asarray(P_l, dtype=np.float64), np.asarray(average_number_of_steps_list, dtype=np.float64)])
    with open(path, "a") as f:
        np.savetxt(f, np.transpose(data), header=RESULT_HEADER, delimiter=",", fmt="%s")
    return data
