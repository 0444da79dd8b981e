"""Builds libssps_b200.so (CUDA kernels + C ABI) in-tree with nvcc for sm_100a.

-fmad=false: every fused multiply-add in the kernels is an explicit fma(), so the operation order
is the one the CPU oracle's kernel-faithful path restates (bit-exact replay parity).
"""
from __future__ import annotations

import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.environ.get("SSPS_LIB") or os.path.join(_HERE, "libssps_b200.so")   # SSPS_LIB: experiment builds
SOURCES = [os.path.join(CSRC, "ssps_capi.cu")]
DEPS = SOURCES + [os.path.join(CSRC, "ssps_kernels.cuh"), os.path.join(CSRC, "ssps_mh.cuh"), os.path.join(CSRC, "ssps_poly.cuh"), os.path.join(CSRC, "ssps_neff.cuh"),
                  os.path.join(os.path.dirname(_HERE), "include", "ssps.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-fmad=false", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libssps_b200.so cannot be built (there is no CPU fallback)")


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(p) > t for p in DEPS if os.path.exists(p))


def build(force: bool = False, verbose: bool = False) -> str:
    if force or is_stale():
        extra = os.environ.get("SSPS_NVCC_EXTRA", "").split()       # experiments: -DSSPS_KS=5 -DSSPS_MIN_CTAS=4 ...
        cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + SOURCES
        subprocess.check_call(cmd)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force=True, verbose=True))
