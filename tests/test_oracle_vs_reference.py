"""Pins the CPU restatement (oracle/ocw_oracle.hpp) against
  (a) every known-answer test the reference holds for this path and
  (b) the unmodified reference compiled from /root/reference (oracle/_ref/libsmsref.so), bit for bit.
Runs on CPU (`-m "not gpu"`)."""
import numpy as np
import pytest

from sms_b200 import instances as I

TRI = I.from_edges("tri", 3, [(0, 1), (1, 2), (2, 0)])


def x_for(inst, assign):
    """x by edge id from {(u,v): value}"""
    x = np.zeros(inst.m)
    for (u, v), w in assign.items():
        row = inst.col[inst.rowptr[u]:inst.rowptr[u + 1]]
        a = int(np.nonzero(row == v)[0][0]) + int(inst.rowptr[u])
        x[inst.eid[a]] = w
    return x


def engines(request):
    out = [request.getfixturevalue("orc")]
    from oracle import oracle_py
    import os
    if oracle_py.have_ref() or os.path.isdir("/root/reference/src"):
        out.append(request.getfixturevalue("ref"))
    return out


# ---- reference KATs: src/graph/test/ocw_separator_gtest.cpp ------------------------------------------------------
def test_kat_simple_no_path(request):  # :5-19
    x = x_for(TRI, {(0, 1): .3, (1, 2): .3, (2, 0): .3})
    for e in engines(request):
        s = e.sweep(TRI, x, blacklist=np.zeros(3, np.uint8), sources=[0])
        assert s.ncuts == 0


def test_kat_most_violated_simple_path(request):  # :21-36
    x = x_for(TRI, {(0, 1): .1, (1, 2): .4, (2, 0): .1})
    for e in engines(request):
        w = e.most_violated(TRI, x, 0)
        assert len(w) == 1 and len(w[0][1]) == 3


def test_kat_simple_path(request):  # :38-53
    x = x_for(TRI, {(0, 1): .9, (1, 2): .9, (2, 0): .9})
    for e in engines(request):
        s = e.sweep(TRI, x, blacklist=np.zeros(3, np.uint8), sources=[0])
        assert s.ncuts >= 1
        _, nodes, stay = s.walks()[0]
        assert (len(stay) - sum(stay)) % 2 == 1
        assert set(nodes) <= {0, 1, 2}


def test_kat_extract_simple1(request):  # :55-78
    g = I.from_edges("g", 4, [(0, 1), (0, 2), (1, 2), (0, 3)])
    x = x_for(g, {(0, 1): 1., (0, 2): 1., (1, 2): 1., (0, 3): .1})
    for e in engines(request):
        res = e.simple_odd_cycles(g, x, [3, 0, 1, 2, 0], [1, 0, 0, 0, 1])
        assert len(res) != 0


def test_kat_extract1(request):  # :80-108
    g = I.from_edges("g", 4, [(0, 1), (0, 2), (0, 3), (1, 2), (2, 3)])
    x = x_for(g, {(0, 1): .5, (0, 2): .3, (0, 3): .1, (1, 2): .1, (2, 3): .1})
    for e in engines(request):
        res = e.extract_cycles(g, x, [0, 1, 2, 3], [0, 1, 1, 1])
        assert len(res) == 2 and len(res[0][1]) == 3 and len(res[1][1]) == 3


def test_kat_extract2(request):  # :110-149
    g = I.from_edges("g", 7, [(0, 1), (0, 6), (1, 2), (1, 6), (2, 3), (2, 4), (2, 5), (3, 4), (4, 5), (5, 6)])
    x = x_for(g, {(0, 1): .1, (0, 6): .1, (1, 2): .9, (1, 6): .5, (2, 3): .9, (2, 4): .1, (2, 5): .1, (3, 4): .1,
                  (4, 5): .1, (5, 6): .9})
    for e in engines(request):
        res = e.extract_cycles(g, x, [0, 1, 2, 3, 4, 5, 6], [1, 0, 0, 1, 1, 0, 1])
        assert len(res) == 3


# ---- src/graph/test/odd_closed_walk_gtest.cpp:149-177 -------------------------------------------------------------
@pytest.mark.parametrize("w,violated", [([.95, .95, .95, .6], True), ([.95, .95, .95, .95], False),
                                        ([.95, .95, .1, .4], True), ([.1, .1, .4, .1], True)])
def test_kat_odd_selection(request, w, violated):
    for e in engines(request):
        st = e.odd_selection([1, 2, 3, 4], w)
        assert (st is not None) == violated
        if violated:
            assert len(st) == 4 and (4 - int(st.sum())) % 2 == 1


# ---- src/conshdlr/test/triangles_gtest.cpp, holes_gtest.cpp ------------------------------------------------------
def test_kat_triangles_and_holes(request):
    for e in engines(request):
        assert len(e.small_cycle_scan(TRI, np.array([1., 1., 1.]), np.zeros((0, 3), np.int32))) == 0  # none registered
        assert len(e.small_cycle_scan(TRI, np.array([1., 1., 1.]), [[0, 1, 2]])) == 1
        assert len(e.small_cycle_scan(TRI, x_for(TRI, {(0, 1): 0., (1, 2): 0., (2, 0): 1.}), [[0, 1, 2]])) == 1
        sq = I.from_edges("sq", 4, [(0, 1), (1, 2), (2, 3), (3, 0)])
        xs = x_for(sq, {(0, 1): .9, (1, 2): .3, (2, 3): .2, (3, 0): .1})
        assert len(e.small_cycle_scan(sq, xs, [[0, 1, 2, 3]])) == 1
        assert len(e.small_cycle_scan(sq, xs, np.zeros((0, 4), np.int32))) == 0


# ---- src/graph/test/small_ccs_gtest.cpp:128-146: K_c has c(c-1)(c-2)/6 triangles and no 4-holes -------------------
@pytest.mark.parametrize("c", [3, 4, 5, 8])
def test_kat_small_ccs_complete(request, c):
    g = I.from_edges("K", c, [(u, v) for u in range(c) for v in range(u + 1, c)])
    for e in engines(request):
        tri, holes = e.small_ccs(g)
        assert len(tri) == c * (c - 1) * (c - 2) // 6 and len(holes) == 0


# ---- restatement == unmodified reference, bit for bit ------------------------------------------------------------
INSTANCES = [lambda: I.torus2d(12), lambda: I.torus3d(5), lambda: I.random_graph(300, 1200, seed=3),
             lambda: I.random_graph(120, 900, seed=5)]
XS = [("uniform", I.xstar_uniform), ("lp_like", I.xstar_lp_like), ("half", I.xstar_half_integral)]


@pytest.mark.parametrize("mk", INSTANCES)
@pytest.mark.parametrize("xname,xf", XS)
@pytest.mark.parametrize("mode", ["I", "V"])
def test_restatement_equals_reference(orc, ref, mk, xname, xf, mode):
    inst = mk()
    x = xf(inst.m, seed=11)
    a = ref.sweep(inst, x, mode=mode, want_labels=True)
    b = orc.sweep(inst, x, mode=mode, want_labels=True)
    assert a.walks() == b.walks()
    assert np.array_equal(a.label_off, b.label_off) and np.array_equal(a.label_node, b.label_node)
    assert np.array_equal(a.label_dist, b.label_dist) and np.array_equal(a.label_hops, b.label_hops)  # bitwise


def test_csr_numbering_equals_networkit(orc, ref):
    inst = I.random_graph(500, 2500, seed=9)
    for got in (orc.csr(inst.n, inst.eu, inst.ev), ref.csr(inst.n, inst.eu, inst.ev)):
        assert np.array_equal(got[0], inst.rowptr) and np.array_equal(got[1], inst.col) and np.array_equal(got[2], inst.eid)


@pytest.mark.parametrize("mk", INSTANCES)
def test_canonical_mode_equals_reference_on_tie_free_input(orc, ref, mk):
    """The canonical tie-breaking (the CUDA path's contract) changes nothing when no ties exist."""
    inst = mk()
    x = I.xstar_uniform(inst.m, seed=5)
    a = ref.sweep(inst, x, mode="I", want_labels=True)
    c = orc.sweep(inst, x, mode="I", want_labels=True, canonical=True)
    assert a.walks() == c.walks()
    assert np.array_equal(a.label_dist, c.label_dist) and np.array_equal(a.label_hops, c.label_hops)


def test_canonical_mode_finds_valid_violated_cycles_on_ties(orc):
    """On tie-heavy x* canonical and reference tie-breaking may pick different equally short paths; every
    canonical-mode walk must still be an odd closed walk of length < 1."""
    from helpers import walk_length
    inst = I.torus2d(10)
    x = I.xstar_half_integral(inst.m, seed=2)
    s = orc.sweep(inst, x, mode="I", canonical=True)
    assert s.ncuts > 0
    for _, nodes, stay in s.walks():
        assert (len(stay) - sum(stay)) % 2 == 1
        assert walk_length(inst, x, nodes, stay) < 1.0


def test_small_cycle_scan_equals_reference(orc, ref):
    g = I.random_graph(60, 500, seed=4)
    tri, holes = ref.small_ccs(g)
    tri2, holes2 = orc.small_ccs(g)
    assert np.array_equal(tri, tri2) and np.array_equal(holes, holes2) and len(tri) > 20
    for xf in (I.xstar_uniform, I.xstar_lp_like, I.xstar_half_integral):
        x = xf(g.m, seed=3)
        assert ref.small_cycle_scan(g, x, tri) == orc.small_cycle_scan(g, x, tri)
        assert ref.small_cycle_scan(g, x, holes[:500]) == orc.small_cycle_scan(g, x, holes[:500])
