"""The C++ host side above the C ABI: sms::GpuOcwSeparator (sms_b200/host/gpu_ocw_separator.hpp) and the drop-in
SCIP handler `cycleparity` (sms_b200/host/conshdlr_oscw.cpp), compiled against the recording SCIP fake and the
NetworKit stand-in (tests/cpp/shim) because neither SCIP nor a built NetworKit exists on the test boxes.

CPU: the sources compile, link against libsmsgpu.so, and construction fails loudly without a device.
GPU (-m gpu): one CONSSEPALP / CONSENFOLP / CONSSEPASOL round through the handler delivers to (fake) SCIP exactly
the rows the reference's handler would build from the oracle's walks (reference conshdlr_oscw.cpp:196-243):
name "osc", lhs -inf, rhs numCrosses-1, -1 on stay / +1 on cross edges, efficacious ones only, no duplicates."""
import math
import os
import subprocess

import numpy as np
import pytest

from helpers import canonical
from sms_b200 import build
from sms_b200 import instances as I

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "cpp", "_build")
DRIVER = os.path.join(OUT, "conshdlr_driver")


@pytest.fixture(scope="module")
def driver():
    lib = build.build()
    os.makedirs(OUT, exist_ok=True)
    srcs = [os.path.join(ROOT, "tests", "cpp", "conshdlr_driver.cpp"), os.path.join(ROOT, "sms_b200", "host", "conshdlr_oscw.cpp")]
    hdrs = [os.path.join(ROOT, "sms_b200", "host", "gpu_ocw_separator.hpp"), os.path.join(ROOT, "tests", "cpp", "shim", "scip", "scip.h")]
    if not os.path.exists(DRIVER) or any(os.path.getmtime(s) > os.path.getmtime(DRIVER) for s in srcs + hdrs + [lib]):
        cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", f"-I{ROOT}/sms_b200/host", f"-I{ROOT}/tests/cpp/shim",
               f"-I{ROOT}/include", *srcs, "-o", DRIVER, f"-L{os.path.dirname(lib)}", "-lsmsgpu", f"-Wl,-rpath,{os.path.dirname(lib)}"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    return DRIVER


def write_instance(path, inst, x, fixed=(), cutpool=1):
    with open(path, "w") as f:
        f.write(f"{inst.n} {inst.m} {inst.root()} {len(fixed)} {cutpool}\n")
        for u, v in zip(inst.eu.tolist(), inst.ev.tolist()):
            f.write(f"{u} {v}\n")
        f.write("\n".join(repr(float(v)) for v in x) + "\n")
        f.write(" ".join(str(int(u)) for u in fixed) + "\n")


def run(driver, path, mode):
    r = subprocess.run([driver, path, mode], capture_output=True, text=True, timeout=300)
    return r.returncode, r.stdout, r.stderr


def parse(stdout):
    rows, result = [], None
    for ln in stdout.splitlines():
        f = ln.split()
        if f[0] in ("pool", "lp"):
            n = int(f[3])
            terms = frozenset((int(t.split(":")[0]), float(t.split(":")[1])) for t in f[4:4 + n])
            tail = ln.split("|")[1].split()
            rows.append((f[0], f[1], float(f[2]), terms, float(tail[1]), tail[3]))
        elif f[0] == "result":
            result = dict(zip(f[0::2], (int(v) for v in f[1::2])))
    return rows, result


def expected_rows(inst, x, bl, orc, min_eff=1e-6):
    """rows the reference handler builds from the walks of one round, deduplicated, efficacious only"""
    exp = orc.sweep(inst, x, blacklist=bl, mode="I", canonical=True)
    xc = np.clip(x, 0.0, 1.0)
    out = set()
    for c in {canonical(n, s) for _, n, s in exp.walks()}:
        L = len(c)
        terms, act = [], 0.0
        for i in range(L):
            u, v, stay = c[i][0], c[(i + 1) % L][0], c[i][1]
            row = inst.col[inst.rowptr[u]:inst.rowptr[u + 1]]
            e = int(inst.eid[int(inst.rowptr[u]) + int(np.nonzero(row == v)[0][0])])
            coef = -1.0 if stay else 1.0
            terms.append((e, coef))
            act += coef * xc[e]
        rhs = sum(1 for _, s in c if not s) - 1
        if (act - rhs) / math.sqrt(L) > min_eff:
            out.add((float(rhs), frozenset(terms)))
    return out


def test_host_sources_compile_and_fail_loudly_without_a_device(driver, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present; covered by the gpu tests")
    inst = I.torus2d(4)
    p = str(tmp_path / "i.txt")
    write_instance(p, inst, I.xstar_uniform(inst.m, seed=1))
    rc, out, err = run(driver, p, "sepalp")
    assert rc == 3 and "no CUDA device" in err and "pool" not in out


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["sepalp", "sepasol", "enfolp"])
@pytest.mark.parametrize("mk,xf", [(lambda: I.torus2d(12), I.xstar_uniform), (lambda: I.torus3d(5), I.xstar_lp_like),
                                   (lambda: I.random_graph(300, 1200, seed=4), I.xstar_uniform)])
def test_handler_round_builds_the_reference_rows(driver, orc, tmp_path, mode, mk, xf):
    inst = mk()
    x = xf(inst.m, seed=11)
    rng = np.random.default_rng(3)
    fixed = [int(u) for u in rng.choice(inst.n, 5, replace=False) if int(u) != inst.root()]
    bl = inst.root_blacklist()
    bl[fixed] = 1
    p = str(tmp_path / "i.txt")
    write_instance(p, inst, x, fixed=fixed, cutpool=1)
    rc, out, err = run(driver, p, mode)
    assert rc == 0, err
    rows, res = parse(out)
    exp = expected_rows(inst, x, bl, orc)
    got = {(r[2], r[3]) for r in rows}
    assert len(rows) == len(got), "duplicate rows reached SCIP"
    assert got == exp
    assert all(r[0] == "pool" and r[1] == "osc" and r[4] <= -1e20 and r[5] == "001" for r in rows)
    assert res["rows_created"] == res["rows_released"] and res["addvar"] == res["cyclelen"]
    SEPARATED, DIDNOTFIND, FEASIBLE = 8, 3, 4
    assert res["result"] == (SEPARATED if exp else (FEASIBLE if mode == "enfolp" else DIDNOTFIND))


@pytest.mark.gpu
def test_handler_sepastore_path_cutoff_and_other_callbacks(driver, orc, tmp_path):
    inst = I.torus2d(8)
    x = I.xstar_uniform(inst.m, seed=2)
    p = str(tmp_path / "i.txt")
    write_instance(p, inst, x, cutpool=0)  # constraints/cycleparity/sepacutpool = FALSE -> SCIPaddRow
    rc, out, err = run(driver, p, "sepalp")
    assert rc == 0, err
    rows, res = parse(out)
    assert rows and all(r[0] == "lp" for r in rows) and res["result"] == 8
    assert {(r[2], r[3]) for r in rows} == expected_rows(inst, x, inst.root_blacklist(), orc)
    # integral points: a cut is feasible, a frustrated assignment is not (CONSCHECK)
    side = np.random.default_rng(1).integers(0, 2, inst.n)
    cut = (side[inst.eu] ^ side[inst.ev]).astype(np.float64)
    # x is indexed by edge id; build it through the CSR's ids
    xe = np.zeros(inst.m)
    for u in range(inst.n):
        for a in range(int(inst.rowptr[u]), int(inst.rowptr[u + 1])):
            xe[inst.eid[a]] = side[u] ^ side[inst.col[a]]
    write_instance(p, inst, xe)
    rc, out, _ = run(driver, p, "check")
    assert rc == 0 and parse(out)[1]["result"] == 4
    rc, out, _ = run(driver, p, "sepalp")  # nothing to separate at a cut
    assert rc == 0 and parse(out)[0] == [] and parse(out)[1]["result"] == 3
    xe[0] = 1 - xe[0]
    write_instance(p, inst, xe)
    rc, out, _ = run(driver, p, "check")
    assert rc == 0 and parse(out)[1]["result"] == 5
    rc, out, _ = run(driver, p, "lock")
    assert rc == 0 and f"locks {2 * inst.m}" in out
    rc, out, _ = run(driver, p, "enfops")
    assert "threw Not implemented: scip_enfops" in out
    rc, out, _ = run(driver, p, "clone")
    assert "clone ok valid 1" in out
    del cut
