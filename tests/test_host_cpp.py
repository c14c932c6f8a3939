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

from helpers import canonical, first_finder
from sms_b200 import build
from sms_b200 import instances as I

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "cpp", "_build")
DRIVER = os.path.join(OUT, "conshdlr_driver")
TRI_DRIVER = os.path.join(OUT, "triangles_driver")


def _build_driver(exe, main_cpp, handler_cpp):
    lib = build.build()
    os.makedirs(OUT, exist_ok=True)
    srcs = [os.path.join(ROOT, "tests", "cpp", main_cpp), os.path.join(ROOT, "sms_b200", "host", handler_cpp)]
    hdrs = [os.path.join(ROOT, "sms_b200", "host", "gpu_ocw_separator.hpp"), os.path.join(ROOT, "tests", "cpp", "shim", "scip", "scip.h"),
            os.path.join(ROOT, "sms_b200", "host", "sms", "conshdlr", "conshdlr_oscw.hpp"),
            os.path.join(ROOT, "sms_b200", "host", "sms", "conshdlr", "conshdlr_triangles.hpp"), os.path.join(ROOT, "include", "smsgpu.h")]
    if not os.path.exists(exe) or any(os.path.getmtime(s) > os.path.getmtime(exe) for s in srcs + hdrs + [lib]):
        cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", f"-I{ROOT}/sms_b200/host", f"-I{ROOT}/tests/cpp/shim",
               f"-I{ROOT}/include", *srcs, "-o", exe, f"-L{os.path.dirname(lib)}", "-lsmsgpu", f"-Wl,-rpath,{os.path.dirname(lib)}"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    return exe


@pytest.fixture(scope="module")
def driver():
    return _build_driver(DRIVER, "conshdlr_driver.cpp", "conshdlr_oscw.cpp")


@pytest.fixture(scope="module")
def tri_driver():
    return _build_driver(TRI_DRIVER, "triangles_driver.cpp", "conshdlr_triangles.cpp")


def write_instance(path, inst, x, fixed=(), cutpool=1):
    with open(path, "w") as f:
        f.write(f"{inst.n} {inst.m} {inst.root()} {len(fixed)} {cutpool}\n")
        for u, v in zip(inst.eu.tolist(), inst.ev.tolist()):
            f.write(f"{u} {v}\n")
        f.write("\n".join(repr(float(v)) for v in x) + "\n")
        f.write(" ".join(str(int(u)) for u in fixed) + "\n")


def run(driver, path, mode, *extra):
    r = subprocess.run([driver, path, mode, *[str(e) for e in extra]], capture_output=True, text=True, timeout=300)
    return r.returncode, r.stdout, r.stderr


def parse(stdout):
    rows, result = [], None
    for ln in stdout.splitlines():
        f = ln.split()
        if f[0] == "triangles":
            continue
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


def test_host_sources_compile_and_fail_loudly_without_a_device(driver, tri_driver, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present; covered by the gpu tests")
    inst = I.torus2d(4)
    p = str(tmp_path / "i.txt")
    write_instance(p, inst, I.xstar_uniform(inst.m, seed=1))
    for d in (driver, tri_driver):
        rc, out, err = run(d, p, "sepalp")
        assert rc == 3 and "no CUDA device" in err and "pool" not in out


def ordered_rows(inst, x, bl, orc, min_eff=1e-6):
    """[(first-finding source, rhs, terms)] of the efficacious cuts in the order the handler receives them"""
    exp = orc.sweep(inst, x, blacklist=bl, mode="I", canonical=True)
    ff = first_finder(exp.walks())
    xc = np.clip(x, 0.0, 1.0)
    out = []
    for c, (src, _) in sorted(ff.items(), key=lambda kv: kv[1]):
        L = len(c)
        terms, act = [], 0.0
        for i in range(L):
            u, v, stay = c[i][0], c[(i + 1) % L][0], c[i][1]
            row = inst.col[inst.rowptr[u]:inst.rowptr[u + 1]]
            e = int(inst.eid[int(inst.rowptr[u]) + int(np.nonzero(row == v)[0][0])])
            terms.append((e, -1.0 if stay else 1.0))
            act += terms[-1][1] * xc[e]
        rhs = sum(1 for _, s in c if not s) - 1
        if (act - rhs) / math.sqrt(L) > min_eff:
            out.append((src, float(rhs), frozenset(terms)))
    return out


@pytest.mark.gpu
def test_cutoff_protocol_of_the_reference(driver, orc, tmp_path):
    """conshdlr_oscw.cpp:188-243 with sepacutpool = FALSE: SCIPaddRow reports `infeasible` for one row.  The reference sets
    CUTOFF, still processes the remaining walks of the SAME source (a later feasible row of that source turns the result
    back into SEPARATED), and stops in front of the next source only while the result is CUTOFF."""
    for inst, x in ((I.random_graph(300, 1200, seed=4), None), (I.torus3d(5), None), (I.torus2d(8), None)):
        x = I.xstar_lp_like(inst.m, seed=2)
        rows = ordered_rows(inst, x, inst.root_blacklist(), orc)
        srcs = [r[0] for r in rows]
        last_of_source = [k for k in range(len(rows)) if k + 1 == len(rows) or srcs[k + 1] != srcs[k]]
        inner = [k for k in range(len(rows) - 1) if srcs[k + 1] == srcs[k]]
        if len(last_of_source) > 2 and inner:
            break
    assert len(last_of_source) > 2 and inner, "the instance must have sources with several cuts and several sources"
    p = str(tmp_path / "i.txt")
    write_instance(p, inst, x, cutpool=0)
    CUTOFF, SEPARATED = 7, 8
    for k in (last_of_source[0], last_of_source[len(last_of_source) // 2], inner[0], inner[-1]):
        rc, out, err = run(driver, p, "sepalp", k)
        assert rc == 0, err
        got, res = parse(out)
        # simulate the reference's loop
        result, n_added, cur = 3, 0, None
        for j, (src, _, _) in enumerate(rows):
            if src != cur:
                if result == CUTOFF:
                    break
                cur = src
            result = CUTOFF if j == k else SEPARATED
            n_added += 1
        assert [(r[2], r[3]) for r in got] == [(r[1], r[2]) for r in rows[:n_added]], f"infeasible at row {k}"
        assert res["result"] == result


def expected_triangle_rows(inst, x, orc, min_eff=1e-6):
    tri, _ = orc.small_ccs(inst)
    xc = np.clip(x, 0.0, 1.0)
    out = []
    for _, nodes, stay in (orc.small_cycle_scan(inst, x, tri) if len(tri) else []):
        terms, act = [], 0.0
        for i in range(3):
            u, v = nodes[i], nodes[(i + 1) % 3]
            row = inst.col[inst.rowptr[u]:inst.rowptr[u + 1]]
            e = int(inst.eid[int(inst.rowptr[u]) + int(np.nonzero(row == v)[0][0])])
            terms.append((e, -1.0 if stay[i] else 1.0))
            act += terms[-1][1] * xc[e]
        rhs = sum(1 for s in stay if not s) - 1
        if (act - rhs) / math.sqrt(3) > min_eff:
            out.append((float(rhs), frozenset(terms)))
    return len(tri), out


@pytest.mark.gpu
@pytest.mark.parametrize("cutpool", [0, 1])
@pytest.mark.parametrize("mk,xf", [(lambda: I.random_graph(80, 700, seed=4), I.xstar_uniform), (lambda: I.random_graph(300, 4000, seed=2), I.xstar_lp_like),
                                   (lambda: I.torus2d(8), I.xstar_uniform)])
def test_triangles_handler_round_builds_the_reference_rows(tri_driver, orc, tmp_path, mk, xf, cutpool):
    """ConshdlrTriangles (reference conshdlr_triangles.cpp:73-132, 193-203) over the fake SCIP: scip_initsol enumerates
    the triangles on the device, CONSSEPALP / CONSENFOLP deliver the rows the reference would build from
    TriangleSeparator::getViolatedTriangles, in list order: "oc", lhs -inf, rhs numCrosses-1, efficacious ones only."""
    inst = mk()
    x = xf(inst.m, seed=11)
    p = str(tmp_path / "i.txt")
    write_instance(p, inst, x, cutpool=cutpool)
    ntri, exp = expected_triangle_rows(inst, x, orc)
    for mode in ("sepalp", "enfolp"):
        rc, out, err = run(tri_driver, p, mode)
        assert rc == 0, err
        assert f"triangles {ntri}" in out
        rows, res = parse(out)
        assert [(r[2], r[3]) for r in rows] == exp
        assert all(r[0] == ("pool" if cutpool else "lp") and r[1] == "oc" and r[4] <= -1e20 and r[5] == "001" for r in rows)
        assert res["rows_created"] == res["rows_released"]
        SEPARATED, DIDNOTFIND, FEASIBLE = 8, 3, 4
        assert res["result"] == (SEPARATED if exp else (FEASIBLE if mode == "enfolp" else DIDNOTFIND))
    if not cutpool and len(exp) > 3:  # CUTOFF: the reference stops at the first infeasible row (:96-98, :121-123)
        rc, out, err = run(tri_driver, p, "sepalp", 2)
        rows, res = parse(out)
        assert [(r[2], r[3]) for r in rows] == exp[:3] and res["result"] == 7
    for mode, msg in (("sepasol", "threw Not implemented: scip_sepasol."), ("enfops", "threw Not implemented: scip_enfops"), ("clone", "clone ok valid 1")):
        rc, out, _ = run(tri_driver, p, mode)
        assert msg in out
    rc, out, _ = run(tri_driver, p, "check")
    assert parse(out)[1]["result"] == 4


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
