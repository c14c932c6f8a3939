"""CPU checks of the drop-in boundary: libsmsgpu.so loads without a GPU, exports exactly the symbols
include/smsgpu.h declares, the ctypes structs match the header, and the product never touches oracle/."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "smsgpu.h")


@pytest.fixture(scope="module")
def lib():
    from sms_b200 import build, capi
    build.build()
    return capi.load_library()


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(smsgpu_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_match_binding_list():
    from sms_b200 import capi
    assert declared_symbols() == sorted(capi.SYMBOLS)


def test_library_exports_every_declared_symbol(lib):
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/smsgpu.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", lib._name], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\bT (smsgpu_[a-z0-9_]+)", out))
    assert exported == set(declared_symbols())


def test_abi_version_and_default_params(lib):
    from sms_b200 import capi
    assert lib.smsgpu_abi_version() == 3
    p = capi.Params()
    lib.smsgpu_default_params(C.byref(p))
    assert p.max_intermediates == 10 and p.best_dist_init == 1.0 - 1e-6 and p.dedup == 1 and p.mode == 0 and p.min_efficacy == 0.0


def test_struct_sizes_match_header():
    """compile a tiny C program against the header and compare sizeof with the ctypes mirrors"""
    from sms_b200 import capi
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "s.c")
        open(src, "w").write('#include <stdio.h>\n#include "smsgpu.h"\nint main(void){printf("%zu %zu %zu\\n",'
                             'sizeof(smsgpu_params),sizeof(smsgpu_cuts),sizeof(smsgpu_stats));return 0;}')
        exe = os.path.join(d, "s")
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), src, "-o", exe], check=True)
        sizes = [int(v) for v in subprocess.run([exe], capture_output=True, text=True).stdout.split()]
    assert sizes == [C.sizeof(capi.Params), C.sizeof(capi.CutsStruct), C.sizeof(capi.StatsStruct)]


def test_create_rejects_bad_graphs_and_fails_loudly_without_gpu(lib):
    """Argument validation happens before any CUDA call; on a machine without a GPU a valid graph must fail with
    SMSGPU_ERR_CUDA (-2) -- there is no CPU fallback."""
    import numpy as np
    from sms_b200 import capi, instances as I
    inst = I.torus2d(4)
    h = C.c_void_p()
    rp = inst.rowptr.copy()
    rp[-1] += 1
    assert lib.smsgpu_create(C.byref(h), inst.n, inst.m, rp.ctypes.data, inst.col.ctypes.data, inst.eid.ctypes.data, 0) == -1
    rp = inst.rowptr.copy()
    rp[1] = 2 * inst.m + 5  # an inner row pointer beyond the arc arrays: rejected before any arc is read
    assert lib.smsgpu_create(C.byref(h), inst.n, inst.m, rp.ctypes.data, inst.col.ctypes.data, inst.eid.ctypes.data, 0) == -1
    col = inst.col.copy()
    col[0] = 0  # self loop at vertex 0
    assert lib.smsgpu_create(C.byref(h), inst.n, inst.m, inst.rowptr.ctypes.data, col.ctypes.data, inst.eid.ctypes.data, 0) == -1
    assert b"self loop" in lib.smsgpu_last_error(None)
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(capi.SmsGpuError) as ei:
            capi.Context(inst.n, inst.rowptr, inst.col, inst.eid)
        assert ei.value.code == -2


def test_product_never_imports_oracle():
    """the oracle is test infrastructure: nothing in sms_b200/ may import, include or link it"""
    bad = []
    for dp, _, files in os.walk(os.path.join(ROOT, "sms_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"^\s*(from|import)\s+oracle|#include\s+[\"<].*oracle|liborc|libsmsref", txt, flags=re.M):
                    bad.append(f)
    assert not bad, bad


def test_mc_files_round_trip(tmp_path):
    """'.mc' writer / reader (harness I/O; format of the reference's src/io/io.cpp): the written file
    reads back to the same numbering -- adjacency order and edge ids are functions of the edge order in the file."""
    import numpy as np
    from sms_b200 import instances as I
    for inst in (I.torus2d(7), I.torus3d(4), I.random_graph(300, 1200, seed=4)):
        path = tmp_path / (inst.name + ".mc")
        I.write_mc(inst, str(path), seed=3)
        lines = path.read_text().splitlines()
        assert lines[0].split() == [str(inst.n), str(inst.m)] and len(lines) == inst.m + 1
        assert all(ln.split()[2] in ("1", "-1") for ln in lines[1:])
        back = I.read_mc(str(path))
        assert back.n == inst.n and back.m == inst.m
        for a in ("rowptr", "col", "eid", "eu", "ev"):
            assert np.array_equal(getattr(back, a), getattr(inst, a)), a
