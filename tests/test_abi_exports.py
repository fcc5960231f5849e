"""CPU-side checks of the drop-in boundary: the shared library builds for sm_100a, loads, and exports
every symbol include/toric_b200.h declares (no compute without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    from toric_rl_decoder_b200 import buildlib
    return buildlib.build()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "toric_b200.h")).read()
    return sorted(set(re.findall(r"^int\s+(tc_\w+)\s*\(", text, flags=re.M)))


def test_header_symbols_exported(lib_path):
    lib = ctypes.CDLL(lib_path)
    names = declared_symbols()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), "missing export %s" % n
    assert lib.tc_abi_version() == 6


def test_binding_covers_header(lib_path):
    from toric_rl_decoder_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    _lib.load()


def test_host_step_struct_layout_matches_header(tmp_path):
    """_lib.HostStep (ctypes) against struct tc_host_step as the C compiler lays it out: size and every field offset."""
    import subprocess
    from toric_rl_decoder_b200 import _lib
    names = [f[0] for f in _lib.HostStep._fields_]
    src = ['#include <stdio.h>', '#include <stddef.h>', '#include "toric_b200.h"', 'int main(void) {',
           '  printf("%zu\\n", sizeof(tc_host_step));']
    src += ['  printf("%%zu\\n", offsetof(tc_host_step, %s));' % n for n in names]
    src += ['  return 0;', '}']
    c = tmp_path / "layout.c"
    c.write_text("\n".join(src))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(c)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert got[0] == ctypes.sizeof(_lib.HostStep)
    assert got[1:] == [getattr(_lib.HostStep, n).offset for n in names]


def test_argument_validation_without_gpu(lib_path):
    from toric_rl_decoder_b200 import _lib
    lib = _lib.load()
    nbytes = ctypes.c_size_t(0)
    assert lib.tc_workspace_bytes(7, 65536, ctypes.byref(nbytes)) == 0 and nbytes.value > 65536 * 4
    assert lib.tc_workspace_bytes(2, 10, ctypes.byref(nbytes)) == -1
    assert lib.tc_workspace_bytes(33, 10, ctypes.byref(nbytes)) == -1
    assert lib.tc_syndrome(None, None, 7, 10, None) == -1
    assert lib.tc_step(None, None, None, None, None, None, 7, 10, None) == -1
    assert lib.tc_persp_write_as(None, None, 0, None, None, None, 0, None, 0, 7, 10, None) == -1
    assert lib.tc_host_actor_step(None) == -1
    hs = _lib.HostStep()
    hs.d, hs.n = 7, 16
    assert lib.tc_host_actor_step(ctypes.byref(hs)) == -1          # no streams / events / buffers


def test_no_cpu_fallback():
    import torch
    import toric_rl_decoder_b200 as pkg
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        pkg.ToricEngine(7, 4)


def test_product_does_not_import_oracle():
    pkg_dir = os.path.join(ROOT, "toric_rl_decoder_b200")
    for fn in os.listdir(pkg_dir):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg_dir, fn)).read()
            assert "oracle" not in src, fn


def test_graphed_decoder_gating_without_gpu():
    """evaluation.graphed_decoder only takes launch-bound, greedy, row-independent cases; everything else keeps
    the eager loop (decided before any CUDA work, so checkable here)."""
    from toric_rl_decoder_b200 import evaluation, networks
    model = networks.NN_11(5, 3)
    rows = evaluation.GRAPH_MAX_ROWS
    assert evaluation.graphed_decoder(model, 5, rows // 50 + 1, "cuda", 75) is None          # not launch-bound
    assert evaluation.graphed_decoder(model, 5, 1, "cuda", 75, epsilon=0.1) is None          # epsilon draw needs the seed
    model.pad_small_batches = False
    assert evaluation.graphed_decoder(model, 5, 1, "cuda", 75) is None                       # rows not declared independent
    assert evaluation.decoders_of(model) == {}


def test_empty_batches_and_bad_sizes_without_gpu(lib_path):
    """n = 0 is legal through the state entry points and returns before any pointer is looked at or any kernel is
    launched (so it is checkable without a device); negative n and out-of-range d are -1 whatever the pointers."""
    from toric_rl_decoder_b200 import _lib
    lib = _lib.load()
    assert lib.tc_syndrome(None, None, 7, 0, None) == 0
    assert lib.tc_step(None, None, None, None, None, None, 7, 0, None) == 0
    assert lib.tc_ground_state(None, None, 7, 0, None) == 0
    assert lib.tc_pack_qubits(None, None, 7, 0, None) == 0 and lib.tc_unpack_qubits(None, None, 7, 0, None) == 0
    assert lib.tc_pack_state(None, None, 7, 0, None) == 0 and lib.tc_unpack_state(None, None, 7, 0, None) == 0
    assert lib.tc_state_rows(None, None, 7, 0, None) == 0
    for d, n in ((7, -1), (2, 0), (33, 0), (0, 5)):
        assert lib.tc_syndrome(None, None, d, n, None) == -1
        assert lib.tc_ground_state(None, None, d, n, None) == -1
        assert lib.tc_state_rows(None, None, d, n, None) == -1


def test_reference_type_surface():
    """The reference's record types on the path (src/util.py:9-15, src/Actor_mp.py:52-56) exist under the same names
    with the same fields."""
    import numpy as np
    import toric_rl_decoder_b200 as pkg
    assert pkg.Action._fields == ("position", "action")
    assert pkg.Perspective._fields == ("perspective", "position")
    assert pkg.Transition._fields == ("state", "action", "reward", "next_state", "terminal")
    assert pkg.action_type.names == ("position", "op") and pkg.action_type["position"].shape == (3,)
    t = pkg.transition_type(7)
    assert t.names == ("perspective", "action", "reward", "next_perspective", "terminal")
    assert t["perspective"].shape == (2, 7, 7) and t["terminal"] == np.dtype(bool) and t["reward"] == np.dtype(float)
