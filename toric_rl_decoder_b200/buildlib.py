"""In-tree build of libtoric_b200.so (hand-written sm_100a CUDA + C ABI).

nvcc cross-compiles without a GPU, so this runs in the build container; the .so is git-ignored but
travels to the GPU box with the snapshot.
"""
import os
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
# TORIC_B200_LIB selects another build of the same sources (kernel-variant experiments); default: in-tree
LIB_PATH = os.environ.get("TORIC_B200_LIB") or os.path.join(PKG_DIR, "libtoric_b200.so")
SOURCES = [os.path.join(CSRC, "toric_b200.cu")]
HEADERS = [os.path.join(CSRC, h) for h in ("common.cuh", "state_kernels.cuh", "sampler.cuh", "reset_env.cuh", "persp_write.cuh", "fused_step.cuh",
                                            "policy_kernels.cuh")] + \
          [os.path.join(os.path.dirname(PKG_DIR), "include", "toric_b200.h")]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC", "-diag-suppress", "63"]


# the same sources with programmatic dependent launch compiled out (every launch takes a full stream dependency):
# the reference build tests/test_gpu_actor.py::test_pdl_chain_matches_serialized_launches compares long chains with
NOPDL_LIB_PATH = os.path.join(PKG_DIR, "libtoric_b200_nopdl.so")


def build_nopdl():
    return build(defines=("TC_PDL=0",), out=NOPDL_LIB_PATH)


def is_stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(p) > t for p in SOURCES + HEADERS)


def build(force=False, verbose=False, defines=(), out=None):
    """Compile the CUDA library for sm_100a.  Returns the path of the .so."""
    if out is not None:
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        cmd = [nvcc] + NVCC_FLAGS + ["-D%s" % d for d in defines] + ["-o", out] + SOURCES
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed:\n%s\n%s" % (" ".join(cmd), res.stderr))
        return out
    if not force and not is_stale():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not os.path.exists(nvcc):
        nvcc = "nvcc"
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + SOURCES
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n%s\n%s" % (" ".join(cmd), res.stderr))
    if verbose:
        print(res.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force=True, verbose=True))
