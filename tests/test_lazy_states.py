"""LazyStates (toric_rl_decoder_b200/hostpipe.py): the row-packed host form of EnvSet.step's `states` array
(src/EnvSet.py:39,47) must index like the (n,2,d,d) uint8 array it stands for.  Host-side logic only -- no GPU."""
import numpy as np
import pytest

from toric_rl_decoder_b200.hostpipe import LazyStates


def _pair(n, d, seed):
    rng = np.random.default_rng(seed)
    full = (rng.random((n, 2, d, d)) < 0.2).astype(np.uint8)
    return full, LazyStates(np.packbits(full, axis=-1, bitorder="little"), d)


@pytest.mark.parametrize("d", (3, 5, 7, 8, 9, 16, 17, 25, 32))
def test_lazy_states_indexing(d):
    full, lazy = _pair(37, d, d)
    assert lazy.shape == full.shape and len(lazy) == len(full) and lazy.dtype == np.uint8 and lazy.ndim == 4
    assert (np.asarray(lazy) == full).all()
    assert (lazy.astype(np.int64) == full).all() and lazy.astype(np.int64).dtype == np.int64
    for idx in (0, 36, -1, slice(3, 20, 4), [5, 1, 1, 30], np.array([2, 7]), np.arange(37) % 3 == 0,
                (4, 1), (4, 1, 2), (4, 1, 2, d - 1), (slice(0, 9), 0), (slice(None), 1, 0, 0), ([3, 4], slice(None), 1),
                Ellipsis, (Ellipsis, 0)):
        got, want = lazy[idx], full[idx]
        assert np.shape(got) == np.shape(want), idx
        assert (got == want).all(), idx
    with pytest.raises(IndexError):
        lazy[37]


def test_lazy_states_copy_detaches():
    full, lazy = _pair(8, 7, 1)
    snap = lazy.copy()
    lazy.packed_rows[:] = 0                       # the pipeline reuses its pinned slot
    assert (np.asarray(snap) == full).all() and not np.asarray(lazy).any()
    assert snap.packed_rows.shape == (8, 2, 7, 1)


def test_lazy_states_empty_batch():
    _, lazy = _pair(0, 5, 0)
    assert np.asarray(lazy).shape == (0, 2, 5, 5) and lazy[0:0].shape == (0, 2, 5, 5)


@pytest.mark.parametrize("d", (3, 7, 9, 17))
def test_lazy_states_assignment_follows_numpy(d):
    """Item assignment on the view (the reference's `next_state[idx] = reset_states`, src/Actor_mp.py:182) gives
    what the same assignment gives on the expanded array, and never writes through to the download slot."""
    from toric_rl_decoder_b200.hostpipe import LazyStates
    rng = np.random.default_rng(d)
    n = 23
    full = (rng.random((n, 2, d, d)) < 0.3).astype(np.uint8)
    slot = np.packbits(full, axis=-1, bitorder="little")
    slot_before = slot.copy()
    view = LazyStates(slot, d)
    ref = full.astype(np.int64)
    idx = np.array([1, 5, 6, 20])
    reset_states = (rng.random((4, 2, d, d)) < 0.4).astype(np.float64)          # resetTerminalEnvs returns float64
    view[idx] = reset_states
    ref[idx] = reset_states
    assert (np.asarray(view) == ref).all() and (slot == slot_before).all()
    mask = rng.random(n) < 0.5
    view[mask] = 0
    ref[mask] = 0
    view[3] = np.ones((2, d, d))
    ref[3] = 1
    view[2:9:3, 1] = 1
    ref[2:9:3, 1] = 1
    view[7, 0, 1, 2] = 1
    ref[7, 0, 1, 2] = 1
    view[idx, 1, :, 0] = np.array([0, 1, 0, 1])[:, None]
    ref[idx, 1, :, 0] = np.array([0, 1, 0, 1])[:, None]
    assert (np.asarray(view) == ref).all() and (view[idx] == ref[idx]).all()
    assert (slot == slot_before).all()
    c = view.copy()
    c[0] = 1 - ref[0]
    assert (np.asarray(view) == ref).all()
    view[...] = 0
    assert not np.asarray(view).any()
