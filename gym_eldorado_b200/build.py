"""In-tree build of the native pieces (nvcc, sm_100a). Used by __graft_entry__.build()."""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libcog_b200.so")

NVCC_FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared"]


def _nvcc():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the CUDA library cannot be built (there is no CPU fallback)")


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def cuda_sources():
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh", ".h"))]
    return srcs + [os.path.join(ROOT, "include", "cog_b200.h")]


def build_cuda(force=False, verbose=False):
    """nvcc -> gym_eldorado_b200/libcog_b200.so (C ABI of include/cog_b200.h)."""
    if force or _stale(LIB, cuda_sources()):
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [
            "-I" + os.path.join(ROOT, "include"), "-I" + CSRC, os.path.join(CSRC, "cog_capi.cu"), "-o", LIB]
        subprocess.check_call(cmd)
    return LIB


def build_all(force=False):
    build_cuda(force)
    try:
        from . import pybind_build
        pybind_build.build(force)
    except ImportError:
        pass


if __name__ == "__main__":
    build_all("--force" in sys.argv)
