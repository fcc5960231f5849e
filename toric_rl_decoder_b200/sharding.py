"""Multi-GPU plumbing: one process per GPU, envs sharded by global index, no data-path collective.

Environments never interact (src/EnvSet.py:14-16: independent deep copies), so rank r of W owns a
contiguous slice of the global env range and runs the same kernels on it.  The Philox counter carries
the GLOBAL env id, so trajectories do not depend on W.  The only collectives are the ones the
reference's actors/learner exchange through shared memory or MPI:
  - all_reduce(SUM) of the int64 counter vector (success / failure statistics; evaluation.py:119-122);
  - broadcast of the flat fp32 policy weights from the learner rank (replaces the shared Array('d') of
    Distributed_mp.py:107-113 / the MPI bcast of src/mpi/IO_mpi.py:49-57).
Backends: "nccl" on GPUs (NVLink 5 / NVSwitch), "gloo" for the CPU tests of this host logic.
"""
import os

import torch
import torch.distributed as dist


def shard_range(n_global, rank, world):
    """(base, count) of the contiguous slice of [0, n_global) owned by `rank`; remainders go to the
    lowest ranks, every env is owned exactly once."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    q, r = divmod(int(n_global), int(world))
    count = q + (1 if rank < r else 0)
    base = rank * q + min(rank, r)
    return base, count


def init_from_env(backend=None):
    """Initialise torch.distributed from RANK / WORLD_SIZE / MASTER_* (torchrun).  Returns (rank, world, local)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend, **kw)
    return rank, world, local


def bind_to_gpu_numa(local_gpu):
    """Pin the calling process to the CPU cores NVML reports as closest to GPU `local_gpu` (same NUMA node /
    PCIe root), so that the pinned host buffers it allocates afterwards are NUMA-local to that GPU.  Matters
    when 8 ranks stream results to host memory at once; a no-op when NVML or the affinity call is unavailable.
    Returns the number of CPUs in the mask (0 if nothing was changed)."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        idx = int(local_gpu)
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                idx = int(vis.split(",")[idx])
            except Exception:
                pass
        h = nv.nvmlDeviceGetHandleByIndex(idx)
        ncpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if not cpus or cpus == allowed:
            return 0
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return 0


def allreduce_counters(counters, group=None):
    """Sum the engine's int64 counter vector over all ranks (in place) and return it."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(counters, op=dist.ReduceOp.SUM, group=group)
    return counters


def max_over_ranks(value, device="cpu", group=None):
    """Device-timed durations are reported as the max over ranks."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def broadcast_weights(model, src=0, group=None):
    """Push the policy network's parameters from `src` to every rank as ONE flat fp32 vector."""
    params = [p for p in model.parameters()]
    if not params:
        return 0
    flat = torch.nn.utils.parameters_to_vector([p.detach().float() for p in params])
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.broadcast(flat, src=src, group=group)
    torch.nn.utils.vector_to_parameters(flat.to(params[0].dtype), params)
    return flat.numel()


def make_sharded_engine(d, n_global, device=None, seed=0, p_error=0.1, max_steps=75, rank=None, world=None):
    """The rank's ToricEngine over its slice of the global env range."""
    from .engine import ToricEngine
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
    base, count = shard_range(n_global, rank, world)
    if device is None:
        device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    return ToricEngine(d, count, device=device, seed=seed, env_base=base, p_error=p_error, max_steps=max_steps)
