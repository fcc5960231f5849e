"""ctypes binding of libtoric_b200.so (the C ABI declared in include/toric_b200.h).

There is no CPU fallback: if the library is missing this raises, and every entry point needs
device pointers.
"""
import ctypes
import os

from .buildlib import LIB_PATH

_c = ctypes
_vp, _i, _i64, _u64, _sz = _c.c_void_p, _c.c_int, _c.c_int64, _c.c_uint64, _c.c_size_t

class HostStep(_c.Structure):
    """struct tc_host_step of include/toric_b200.h (argument block of tc_host_actor_step), field for field."""
    _fields_ = [(n, _vp) for n in ("qubits", "syn", "prev_syn", "gs", "status", "steps", "episode", "p_thresh",
                                   "counters", "offsets", "overflow", "out", "pos")] + \
               [("cap", _i64), ("out_dtype", _i), ("workspace", _vp), ("workspace_bytes", _sz),
                ("flags", _i), ("max_steps", _i), ("mode", _i), ("n_fixed", _i), ("seed", _u64), ("env_base", _i64),
                ("d", _i), ("n", _i64)] + \
               [(n, _vp) for n in ("h_actions", "d_actions", "d_stage", "h_stage")] + \
               [("stage_bytes", _sz), ("o_reward", _sz), ("o_counts", _sz), ("o_term", _sz), ("o_state", _sz),
                ("state_mode", _i)] + \
               [(n, _vp) for n in ("h2d_stream", "compute_stream", "d2h_stream", "ev_actions", "ev_compute",
                                   "ev_unpacked_prev", "ev_unpacked", "ev_result")]


# name -> argtypes, exactly the prototypes of include/toric_b200.h
SIGNATURES = {
    "tc_abi_version": [],
    "tc_workspace_bytes": [_i, _i64, _c.POINTER(_sz)],
    "tc_pack_qubits": [_vp, _vp, _i, _i64, _vp],
    "tc_unpack_qubits": [_vp, _vp, _i, _i64, _vp],
    "tc_pack_state": [_vp, _vp, _i, _i64, _vp],
    "tc_unpack_state": [_vp, _vp, _i, _i64, _vp],
    "tc_state_rows": [_vp, _vp, _i, _i64, _vp],
    "tc_syndrome": [_vp, _vp, _i, _i64, _vp],
    "tc_step": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i64, _vp],
    "tc_reset": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _u64, _i64, _i, _vp],
    "tc_ground_state": [_vp, _vp, _i, _i64, _vp],
    "tc_perspectives": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _sz, _i, _i64, _vp],
    "tc_perspectives_static": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _i64, _vp],
    "tc_persp_count": [_vp, _vp, _vp, _vp, _sz, _i, _i64, _vp],
    "tc_persp_write": [_vp, _vp, _vp, _vp, _vp, _i64, _vp, _sz, _i, _i64, _vp],
    "tc_persp_write_as": [_vp, _vp, _i, _vp, _vp, _vp, _i64, _vp, _sz, _i, _i64, _vp],
    "tc_perspectives_as": [_vp, _vp, _i, _vp, _vp, _vp, _vp, _i64, _vp, _sz, _i, _i64, _vp],
    "tc_actor_pre": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _u64, _i64,
                     _vp, _vp, _vp, _sz, _i, _i64, _vp],
    "tc_actor_step": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _u64, _i64,
                      _vp, _vp, _vp, _vp, _vp, _i64, _vp, _sz, _i, _i64, _vp],
    "tc_host_actor_step": [_c.POINTER(HostStep)],
    "tc_select_action": [_vp, _vp, _vp, _vp, _vp, _vp, _c.c_uint32, _vp, _u64, _i64, _vp, _vp, _vp, _vp, _i64, _vp],
    "tc_p_error_curriculum": [_vp, _vp, _vp, _vp, _c.c_uint32, _c.c_uint32, _c.c_uint32, _i, _i, _u64, _i64, _i64, _vp],
    "tc_transitions": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i64, _vp],
    "tc_td_priorities": [_vp, _vp, _vp, _vp, _c.c_float, _vp, _i64, _vp],
    "tc_segmented_max": [_vp, _vp, _i, _vp, _vp, _i64, _vp],
}

ABI_VERSION = 6
_lib = None


def bind(path):
    """ctypes handle of one build of the library with every prototype set (a second build of the same sources,
    e.g. the TC_PDL=0 variant the dependent-launch stress test compares against, can be bound next to the default)."""
    lib = ctypes.CDLL(path)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError if the symbol is missing
        fn.argtypes = argtypes
        fn.restype = _i
    if lib.tc_abi_version() != ABI_VERSION:
        raise RuntimeError("%s: ABI version mismatch" % path)
    return lib


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libtoric_b200.so is not built (%s). Run `python -m toric_rl_decoder_b200.buildlib` or "
            "`__graft_entry__.build()`. There is no CPU fallback." % LIB_PATH)
    _lib = bind(LIB_PATH)
    return _lib


def check(rc, what):
    if rc == 0:
        return
    if rc == -1:
        raise ValueError("%s: bad argument" % what)
    if rc == -2:
        raise ValueError("%s: workspace or capacity too small" % what)
    raise RuntimeError("%s: CUDA error %d" % (what, rc))
