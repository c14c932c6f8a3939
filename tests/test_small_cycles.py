"""K6 end to end: the triangle / four-hole lists of SmallChordlessCycles::computeSmallCCs (src/graph/small_ccs.cpp:18-43)
and the scans of TriangleSeparator / HoleSeparator (triangles.hpp:34-53, holes.hpp:38-54) on the reference's own fixtures.
Goldens: tests/golden/smallcc_*.npz, produced by the UNMODIFIED reference (tests/golden/make_small_cycle_golden.py); the
counts are the ones the reference's small_ccs_gtest.cpp pins (ca-netscience 921 triangles, ego-facebook 91,
tripleSquareTriangle 2 + 2).
CPU: the restatement reproduces them.  GPU: smsgpu_small_cycles_enumerate produces the identical lists IN THE IDENTICAL
ORDER, and the scans return the identical violated walks."""
import glob
import hashlib
import os

import numpy as np
import pytest

from sms_b200 import instances as I

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "smallcc_*.npz")))
KAT = {"ca-netscience": 921, "ego-facebook": 91, "tripleSquareTriangle": 2}  # small_ccs_gtest.cpp:45-126


def _id(p):
    return os.path.basename(p)[8:-4]


def load(path):
    g = np.load(path)
    return I.from_edges(_id(path), int(g["n"]), np.stack([g["eu"], g["ev"]], axis=1)), g


def digest(a):
    return hashlib.blake2b(np.ascontiguousarray(a, np.int32).tobytes(), digest_size=8).hexdigest()


def check_lists(tri, holes, g, name):
    assert len(tri) == int(g["ntri"]) and len(holes) == int(g["nholes"])
    if name in KAT:
        assert len(tri) == KAT[name]
    assert digest(tri) == str(g["tri_hash"]) and digest(holes) == str(g["holes_hash"])  # same cycles, same order
    if "tri" in g:
        assert np.array_equal(tri, g["tri"])
    if "holes" in g:
        assert np.array_equal(holes, g["holes"])


def test_goldens_present():
    assert {_id(p) for p in GOLD} >= {"ca-netscience", "ego-facebook", "tripleSquareTriangle", "be120.3.5", "bqp250-3"}


@pytest.mark.parametrize("path", GOLD, ids=_id)
def test_oracle_reproduces_reference_lists_and_scans(orc, path):
    inst, g = load(path)
    tri, holes = orc.small_ccs(inst)
    check_lists(tri, holes, g, _id(path))
    vt = orc.small_cycle_scan(inst, g["x"], tri) if len(tri) else []
    assert [n for _, n, _ in vt] == [tuple(r) for r in g["viol_tri_nodes"].tolist()]
    assert [s for _, _, s in vt] == [tuple(r) for r in g["viol_tri_stay"].tolist()]
    vh = orc.small_cycle_scan(inst, g["x"], holes[:20000]) if len(holes) else []
    assert [n for _, n, _ in vh] == [tuple(r) for r in g["viol_hole_nodes"].tolist()]
    assert [s for _, _, s in vh] == [tuple(r) for r in g["viol_hole_stay"].tolist()]


# ---- GPU ---------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def gpu():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from sms_b200 import build
    build.build()
    return True


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLD, ids=_id)
def test_gpu_enumeration_and_scans_equal_the_reference(gpu, path):
    from sms_b200 import capi
    inst, g = load(path)
    ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid)
    try:
        nt, nh = ctx.small_cycles_enumerate(True)
        tri, holes = ctx.small_cycles_get(3), ctx.small_cycles_get(4)
        assert (nt, nh) == (len(tri), len(holes))
        check_lists(tri, holes, g, _id(path))
        ctx.set_round_input(g["x"], np.zeros(inst.n, np.uint8))
        vt = ctx.small_cycles_scan(3).walks()
        assert [n for _, n, _ in vt] == [tuple(r) for r in g["viol_tri_nodes"].tolist()]
        assert [s for _, _, s in vt] == [tuple(r) for r in g["viol_tri_stay"].tolist()]
        assert all(tuple(tri[i]) == n for i, n, _ in vt)  # src = index into the enumerated list
        if nh:
            ctx.small_cycles_set(holes[:20000])
            vh = ctx.small_cycles_scan(4).walks()
            assert [n for _, n, _ in vh] == [tuple(r) for r in g["viol_hole_nodes"].tolist()]
            assert [s for _, _, s in vh] == [tuple(r) for r in g["viol_hole_stay"].tolist()]
    finally:
        ctx.close()


@pytest.mark.gpu
def test_gpu_enumeration_on_generated_graphs(gpu, orc):
    """complete graph: C(c, 3) triangles, no hole (small_ccs_gtest.cpp:128-146); sparse / dense random graphs and tori
    against the restatement, list for list"""
    from sms_b200 import capi
    c = 40
    cases = [I.from_edges("K40", c, [(i, j) for i in range(c) for j in range(i + 1, c)]), I.random_graph(80, 700, seed=4),
             I.random_graph(3000, 12000, seed=11), I.torus2d(12), I.torus3d(5), I.random_graph(300, 9000, seed=2)]
    for inst in cases:
        tri, holes = orc.small_ccs(inst)
        ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid)
        try:
            nt, nh = ctx.small_cycles_enumerate(True)
            assert (nt, nh) == (len(tri), len(holes)), inst.name
            assert np.array_equal(ctx.small_cycles_get(3), tri) and np.array_equal(ctx.small_cycles_get(4), holes), inst.name
            nt2, nh2 = ctx.small_cycles_enumerate(False)  # triangles only: the holes handler is not always registered
            assert nt2 == nt and nh2 == 0
        finally:
            ctx.close()
        if inst.name == "K40":
            assert nt == c * (c - 1) * (c - 2) // 6 and nh == 0
