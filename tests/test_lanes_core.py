"""The sequential core of the lanes engine (sms_b200/csrc/ocw_lanes_core.cuh: one source per thread) compiled for the
HOST and run on the CPU against the oracle: labels (dist, hops) bit for bit, arcs scanned / settled / number of
intermediates per source, and predecessors that are consistent with the labels.  The same code runs per thread on the
GPU (tests/test_lanes_gpu.py compares its walks with the oracle and its whole round with the pairs engine's)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from sms_b200 import instances as I

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "cpp", "_build")
LIB = os.path.join(OUT, "liblaneshost.so")


@pytest.fixture(scope="module")
def lanes():
    os.makedirs(OUT, exist_ok=True)
    srcs = [os.path.join(ROOT, "tests", "cpp", "lanes_host.cpp"), os.path.join(ROOT, "sms_b200", "csrc", "ocw_lanes_core.cuh"),
            os.path.join(ROOT, "sms_b200", "csrc", "ocw_types.cuh")]
    if not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        cmd = ["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-x", "c++", f"-I{ROOT}/sms_b200/csrc", "-I/usr/local/cuda/include",
               srcs[0], "-o", LIB]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    return ctypes.CDLL(LIB)


def run_core(lib, inst, x, bl, src, tab_lg=13, nbuckets=2048, park=8192):  # park: queue entries
    nsrc = len(src)
    arcs = np.zeros(nsrc, np.int64)
    settled = np.zeros(nsrc, np.int64)
    histn = np.zeros(nsrc, np.int64)
    flags = np.zeros(nsrc, np.int32)
    best = np.zeros(nsrc, np.float64)
    lab_off = np.zeros(nsrc + 1, np.int64)
    cap = 64 * nsrc * 64 + 1024
    ln = np.zeros(cap, np.int32)
    ld = np.zeros(cap, np.float64)
    lh = np.zeros(cap, np.int32)
    lp = np.zeros(cap, np.int32)
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    rp = np.ascontiguousarray(inst.rowptr, np.int64)
    col = np.ascontiguousarray(inst.col, np.int32)
    eid = np.ascontiguousarray(inst.eid, np.int32)
    xx = np.ascontiguousarray(x, np.float64)
    bb = np.ascontiguousarray(bl, np.uint8)
    ss = np.ascontiguousarray(src, np.int32)
    lib.lanes_host_sweep.restype = ctypes.c_int
    rc = lib.lanes_host_sweep(ctypes.c_int64(inst.n), ctypes.c_int64(inst.m), P(rp), P(col), P(eid), P(xx), P(bb), P(ss),
                              ctypes.c_int64(nsrc), ctypes.c_double(1.0 - 1e-6), ctypes.c_int32(tab_lg), ctypes.c_int32(nbuckets),
                              ctypes.c_int32(park), P(arcs), P(settled), P(histn), P(flags), P(best), P(lab_off), ctypes.c_int64(cap),
                              P(ln), P(ld), P(lh), P(lp))
    assert rc == 0
    return arcs, settled, histn, flags, best, lab_off, ln, ld, lh, lp


CASES = [lambda: I.torus2d(10), lambda: I.torus3d(6), lambda: I.random_graph(400, 1600, seed=3), lambda: I.random_graph(150, 1500, seed=5),
         lambda: I.random_graph(3000, 12000, seed=11)]
XS = [I.xstar_uniform, I.xstar_lp_like, I.xstar_half_integral]


@pytest.mark.parametrize("mk", CASES)
@pytest.mark.parametrize("xf", XS)
def test_lanes_core_equals_oracle(lanes, orc, mk, xf):
    inst = mk()
    x = xf(inst.m, seed=7)
    bl = inst.root_blacklist()
    src = inst.handler_sources()
    if len(src) > 1200:
        src = src[::5]
    exp = orc.sweep(inst, x, blacklist=bl, sources=src, mode="I", want_labels=True, canonical=True)
    arcs, settled, histn, flags, best, lo, ln, ld, lh, lp = run_core(lanes, inst, x, bl, src)
    done = (flags >= 0) & ((flags & 1) == 0)
    # LP-like points overflow a 16-key bucket at once (exact ties): those sources belong to the runs engine
    if xf is I.xstar_uniform:
        assert done.sum() >= 0.9 * (flags >= 0).sum()
    assert np.array_equal(arcs[done], exp.stats[done, 0]), "arcs scanned differ"
    assert np.array_equal(settled[done], exp.stats[done, 1]), "settled differ"
    assert np.array_equal(histn[done], exp.stats[done, 4]), "number of intermediates differs"
    for k in np.nonzero(done)[0]:
        ea, eb = int(exp.label_off[k]), int(exp.label_off[k + 1])
        omap = {int(n): (d, int(h)) for n, d, h in zip(exp.label_node[ea:eb], exp.label_dist[ea:eb], exp.label_hops[ea:eb])}
        mine = {}
        for n, d, h, p in zip(ln[lo[k]:lo[k + 1]], ld[lo[k]:lo[k + 1]], lh[lo[k]:lo[k + 1]], lp[lo[k]:lo[k + 1]]):
            assert int(n) in omap, f"source {k}: label for node {n} unknown to the oracle"
            assert (d, int(h)) == omap[int(n)], f"source {k} node {n}: ({d},{h}) oracle {omap[int(n)]}"
            mine[int(n)] = (d, int(h), int(p))
        # predecessors: one hop less, and the label is the predecessor's plus the arc
        s = int(src[k])
        for n, (d, h, p) in mine.items():
            if n == s:
                continue
            assert p in mine and mine[p][1] == h - 1 and mine[p][0] <= d, f"source {k} node {n}: predecessor {p}"


def test_lanes_core_small_table_overflows_cleanly(lanes, orc):
    """a table / parking area too small for the ball ends the search with the overflow flag, never with a wrong result"""
    inst = I.random_graph(3000, 12000, seed=11)
    x = I.xstar_uniform(inst.m, seed=7)
    bl = inst.root_blacklist()
    src = inst.handler_sources()[::40]
    exp = orc.sweep(inst, x, blacklist=bl, sources=src, mode="I", want_labels=False, canonical=True)
    for kw in (dict(tab_lg=6), dict(park=64), dict(nbuckets=64)):
        arcs, settled, histn, flags, *_ = run_core(lanes, inst, x, bl, src, **kw)
        done = (flags >= 0) & ((flags & 1) == 0)
        assert (flags[flags >= 0] & 1).any(), kw
        assert np.array_equal(arcs[done], exp.stats[done, 0]) and np.array_equal(settled[done], exp.stats[done, 1])
