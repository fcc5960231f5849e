"""tc_perspectives_static: the perspective generator for callers that cannot allocate device memory.

Some users of libtoric_b200.so are handed their device buffers by somebody else and have no allocator of their own
at the moment of the call: a plug-in inside a framework whose memory pool it must not bypass, a process that gave
all device memory to one arena at start-up, a latency-bound decoding loop that must not wait on the allocator's
lock.  They can pass inputs and outputs but not the scratch block tc_perspectives wants, so
tc_perspectives_static takes its scratch from a 1 MiB array built into the library (one per GPU, shared by every
call, never cleared).

What is checked here, through the C ABI only: the new entry point gives exactly what tc_perspectives gives with a
caller-owned workspace -- counts, offsets, positions and the 0/1 perspective tensor, bit for bit (the tolerance of
all the integer/bit parity tests of this suite is zero) -- on several calls in a row whose lattice size and batch
size change, so that whatever the previous call left in the shared array would show up if anything depended on it;
and that a batch too large for the array is refused with -2 before anything is launched.

The file sorts last on purpose: under `pytest -x` a failure here cannot hide another test.  It needs nothing from
the tests before it (first use on the device does the one-time address look-up itself) and leaves nothing behind.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_BYTES = 1 << 20      # TC_STATIC_SCRATCH_BYTES
STATIC_MAX_ENVS = 65536             # TC_STATIC_MAX_ENVS
POISON = 0x5a


@pytest.fixture(scope="module")
def lib(cuda_device):
    from toric_rl_decoder_b200 import _lib
    return _lib.load()


def _syndromes(lib, d, n, seed, density=0.08):
    """n random defect patterns of a d x d lattice as the library's row-packed uint32 [n][2][d] planes."""
    rng = np.random.default_rng(seed)
    state01 = torch.from_numpy((rng.random((n, 2, d, d)) < density).astype(np.uint8)).cuda()
    syn = torch.zeros((n, 2, d), dtype=torch.int32, device="cuda")
    assert lib.tc_pack_state(state01.data_ptr(), syn.data_ptr(), d, n, None) == 0
    return syn


def _outputs(d, n, cap):
    """Caller-owned result buffers, pre-filled so that nothing stale can pass for a result."""
    dev = "cuda"
    out = torch.full((cap, 2, d, d), float("nan"), dtype=torch.float32, device=dev)
    pos = torch.full((cap, 4), POISON, dtype=torch.uint8, device=dev)
    counts = torch.full((n,), -7, dtype=torch.int32, device=dev)
    offsets = torch.full((n + 1,), -7, dtype=torch.int64, device=dev)
    overflow = torch.zeros(1, dtype=torch.int32, device=dev)
    return out, pos, counts, offsets, overflow


def _with_workspace(lib, syn, d, n, cap):
    nbytes = ctypes.c_size_t(0)
    assert lib.tc_workspace_bytes(d, n, ctypes.byref(nbytes)) == 0
    ws = torch.full((max(nbytes.value, 256),), POISON, dtype=torch.uint8, device="cuda")
    o = _outputs(d, n, cap)
    rc = lib.tc_perspectives(syn.data_ptr(), o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(), o[3].data_ptr(),
                             o[4].data_ptr(), cap, ws.data_ptr(), ws.numel(), d, n, None)
    assert rc == 0
    torch.cuda.synchronize()
    return o


def _without_allocation(lib, syn, d, n, cap, check_free_memory=False):
    o = _outputs(d, n, cap)                 # the caller's buffers exist before the call
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = lib.tc_perspectives_static(syn.data_ptr(), o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(),
                                    o[3].data_ptr(), o[4].data_ptr(), cap, d, n, None)
    assert rc == 0
    torch.cuda.synchronize()
    if check_free_memory:                   # the driver's own count of free device memory did not move
        assert torch.cuda.mem_get_info()[0] == free_before
    return o


def _same(ref, got, n):
    r_out, r_pos, r_counts, r_offsets, r_over = ref
    g_out, g_pos, g_counts, g_offsets, g_over = got
    assert torch.equal(r_counts, g_counts) and torch.equal(r_offsets, g_offsets)
    assert int(r_over.item()) == int(g_over.item())
    total = int(r_offsets[n].item())
    assert total == int(r_counts.sum().item()) and total > 0
    assert torch.equal(r_out[:total], g_out[:total])
    assert torch.equal(r_pos[:total, :3], g_pos[:total, :3])
    # nothing behind the last perspective was touched by either path
    assert torch.isnan(g_out[total:]).all() and (g_pos[total:] == POISON).all()


# back-to-back calls on the one shared array: lattice size, group width of the count kernels (d <= 8 / <= 16 / <= 32)
# and batch size all change from one call to the next, and the first shape comes back at the end
SEQUENCE = ((7, 4096, 1), (7, 4096, 2), (12, 777, 3), (3, 30000, 4), (25, 96, 5), (5, 1, 6), (7, 4096, 1))


def test_static_scratch_matches_caller_workspace_on_consecutive_calls(lib):
    for d, n, seed in SEQUENCE:
        syn = _syndromes(lib, d, n, seed)
        cap = n * 2 * d * d                 # every qubit next to a defect: cannot overflow
        ref = _with_workspace(lib, syn, d, n, cap)
        # twice in a row on the same input, then on to the next shape; by the second call every kernel of this
        # shape has been loaded, so a change in free device memory could only be an allocation
        _same(ref, _without_allocation(lib, syn, d, n, cap), n)
        _same(ref, _without_allocation(lib, syn, d, n, cap, check_free_memory=True), n)


def test_static_scratch_overflow_flag_and_empty_batch(lib):
    d, n = 7, 2048
    syn = _syndromes(lib, d, n, 11)
    full = _with_workspace(lib, syn, d, n, n * 2 * d * d)
    cap = int(full[3][n].item()) // 2       # room for half of the perspectives: the tail envs are skipped, flag set
    ref = _with_workspace(lib, syn, d, n, cap)
    got = _without_allocation(lib, syn, d, n, cap)
    assert int(ref[4].item()) == 1 and int(got[4].item()) == 1
    assert torch.equal(ref[2], got[2]) and torch.equal(ref[3], got[3])
    fit = int((ref[3][1:] <= cap).sum().item())
    kept = int(ref[3][fit].item())
    assert torch.equal(ref[0][:kept], got[0][:kept]) and torch.equal(ref[1][:kept, :3], got[1][:kept, :3])
    # n = 0: offsets = [0] and nothing else
    offsets = torch.full((1,), -7, dtype=torch.int64, device="cuda")
    assert lib.tc_perspectives_static(None, None, None, None, offsets.data_ptr(), None, 0, d, 0, None) == 0
    torch.cuda.synchronize()
    assert int(offsets.item()) == 0


def test_static_scratch_size_limit(lib, capfd):
    nbytes = ctypes.c_size_t(0)
    for d in range(3, 33):                  # the promised batch size fits for every lattice size
        assert lib.tc_workspace_bytes(d, STATIC_MAX_ENVS, ctypes.byref(nbytes)) == 0
        assert nbytes.value <= STATIC_SCRATCH_BYTES
    d, n = 7, 1 << 18                       # needs more scratch than the array has: refused, nothing launched
    assert lib.tc_workspace_bytes(d, n, ctypes.byref(nbytes)) == 0 and nbytes.value > STATIC_SCRATCH_BYTES
    canary = torch.full((64,), -7, dtype=torch.int64, device="cuda")
    rc = lib.tc_perspectives_static(canary.data_ptr(), canary.data_ptr(), canary.data_ptr(), canary.data_ptr(),
                                    canary.data_ptr(), canary.data_ptr(), 1, d, n, None)
    assert rc == -2
    assert "TC_STATIC_SCRATCH_BYTES" in capfd.readouterr().err
    torch.cuda.synchronize()
    assert (canary == -7).all()
    assert lib.tc_perspectives_static(None, None, None, None, None, None, 0, 2, 4, None) == -1     # d out of range

    # the largest promised batch runs (d = 3 keeps the output tensor small)
    n = STATIC_MAX_ENVS
    syn = _syndromes(lib, 3, n, 12)
    cap = n * 18
    _same(_with_workspace(lib, syn, 3, n, cap), _without_allocation(lib, syn, 3, n, cap), n)
