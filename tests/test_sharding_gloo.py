"""Host-side multi-rank logic on CPU: world_size-2 gloo process group (the N>1 path of bench.py uses the
same functions over NCCL)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from toric_rl_decoder_b200 import sharding


def test_shard_range_partitions():
    for n in (0, 1, 7, 65536, 1000003):
        for w in (1, 2, 3, 8):
            seen = 0
            for r in range(w):
                base, count = sharding.shard_range(n, r, w)
                assert base == seen and count >= 0
                seen += count
            assert seen == n
            sizes = [sharding.shard_range(n, r, w)[1] for r in range(w)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    r, w, _ = sharding.init_from_env(backend="gloo")
    assert (r, w) == (rank, world)
    # counters: each rank contributes its shard's totals
    base, count = sharding.shard_range(1001, rank, world)
    counters = torch.zeros(16, dtype=torch.int64)
    counters[0] = count
    counters[1] = base
    sharding.allreduce_counters(counters)
    # weights: rank 0 holds the "learner" weights
    torch.manual_seed(rank)
    model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Linear(5, 3))
    n_params = sharding.broadcast_weights(model, src=0)
    flat = torch.nn.utils.parameters_to_vector(model.parameters()).detach()
    t = sharding.max_over_ranks(10.0 + rank)
    out.put((rank, counters.tolist(), flat.tolist(), n_params, t))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo():
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, c0, w0, n0, t0), (r1, c1, w1, n1, t1) = res
    assert c0 == c1 and c0[0] == 1001 and c0[1] == sharding.shard_range(1001, 1, 2)[0]
    assert w0 == w1 and n0 == n1 == 6 * 5 + 5 + 5 * 3 + 3
    torch.manual_seed(0)
    ref = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Linear(5, 3))
    assert w0 == torch.nn.utils.parameters_to_vector(ref.parameters()).tolist()
    assert t0 == t1 == 11.0
