"""Builds libsmsgpu.so (the product: sm_100a kernels + C ABI) in-tree with nvcc.

nvcc cross-compiles without a GPU, so this runs in the CPU container; the built .so is git-ignored and
travels to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "libsmsgpu.so")
LIB_DEBUG = os.path.join(CSRC, "libsmsgpu_dbg.so")  # -DSMSGPU_DEBUG_CHECKS: device asserts (SMSGPU_LIB selects it)
SOURCES = ["smsgpu.cu"]
HEADERS = ["ocw_types.cuh", "ocw_search.cuh", "ocw_search_fast.cuh", "ocw_search_batch.cuh", "ocw_walks.cuh", "ocw_kernels.cuh", "dedup_kernels.cuh",
           "small_cycles.cuh", os.path.join("..", "..", "include", "smsgpu.h")]

NVCC_FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-shared",
              "-Xcompiler", "-fPIC", "-cudart", "static"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False, verbose: bool = False, debug: bool = False) -> str:
    """Compile sms_b200/csrc/*.cu for sm_100a into sms_b200/csrc/libsmsgpu.so; returns the path."""
    if debug:
        cmd = [_nvcc(), *NVCC_FLAGS, "-DSMSGPU_DEBUG_CHECKS", "-o", LIB_DEBUG, *[os.path.join(CSRC, s) for s in SOURCES], "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
        return LIB_DEBUG
    if not force and not needs_build():
        return LIB
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", LIB, *[os.path.join(CSRC, s) for s in SOURCES], "-ldl"]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


if __name__ == "__main__":
    import sys
    if "--debug" in sys.argv:
        print(build(debug=True))
    else:
        print(build(force=True, verbose=True))
