"""Builds libsmsgpu.so (the product: sm_100a kernels + C ABI) in-tree with nvcc.

nvcc cross-compiles without a GPU, so this runs in the CPU container; the built .so is git-ignored and
travels to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import glob
import hashlib
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "libsmsgpu.so")
LIB_DEBUG = os.path.join(CSRC, "libsmsgpu_dbg.so")  # -DSMSGPU_DEBUG_CHECKS: device asserts (SMSGPU_LIB selects it)
SOURCES = ["smsgpu.cu"]
STAMP = LIB + ".srchash"  # sha256 of every source the library was built from (bench.py prints it)

NVCC_FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-shared",
              "-Xcompiler", "-fPIC", "-cudart", "static"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def dependencies() -> list[str]:
    """Every file the library is compiled from: all of csrc/*.cu, csrc/*.cuh and the C-ABI header."""
    deps = sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cuh")))
    deps.append(os.path.normpath(os.path.join(HERE, "..", "include", "smsgpu.h")))
    return deps


def source_hash() -> str:
    h = hashlib.sha256()
    for d in dependencies():
        h.update(os.path.basename(d).encode())
        with open(d, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()[:16]


def built_hash() -> str | None:
    """Source hash recorded when libsmsgpu.so was built (None: no library / no stamp)."""
    try:
        with open(STAMP) as f:
            return f.read().strip()
    except OSError:
        return None


def needs_build() -> bool:
    """The library is stale when its recorded source hash differs from the sources' (content, not mtime: the
    snapshot that travels to the GPU box does not keep mtimes in order)."""
    return not os.path.exists(LIB) or built_hash() != source_hash()


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
    with open(STAMP, "w") as f:
        f.write(source_hash() + "\n")
    return LIB


def build_variant(name: str, flags: list[str]) -> str:
    """Experiment builds (profiles/README.md): csrc/libsmsgpu_<name>.so with extra nvcc flags, selected at run time with
    SMSGPU_LIB=<path>.  Never the default library."""
    out = os.path.join(CSRC, f"libsmsgpu_{name}.so")
    cmd = [_nvcc(), *NVCC_FLAGS, *flags, "-o", out, *[os.path.join(CSRC, s) for s in SOURCES], "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    return out


if __name__ == "__main__":
    import sys
    if "--variant" in sys.argv:
        i = sys.argv.index("--variant")
        print(build_variant(sys.argv[i + 1], sys.argv[i + 2:]))
    elif "--debug" in sys.argv:
        print(build(debug=True))
    else:
        print(build(force=True, verbose=True))
