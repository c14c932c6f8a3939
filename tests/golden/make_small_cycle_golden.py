#!/usr/bin/env python
"""Generates tests/golden/smallcc_*.npz from the UNMODIFIED reference (oracle/_ref): the triangle / four-hole lists of
SmallChordlessCycles::computeSmallCCs (src/graph/small_ccs.cpp:18-43) on the reference's own fixtures -- the instances
its small_ccs_gtest.cpp pins (ca-netscience.mc: 921 triangles, ego-facebook.mc: 91, tripleSquareTriangle.mc: 2 + 2) and the
dense instances SURVEY.md 8(a9) names (be120.3.5.mc, bqp250-3.mc) -- plus what TriangleSeparator / HoleSeparator return on a
seeded x* (triangles.hpp:34-53, holes.hpp:38-54).  Run in the CPU container (needs /root/reference):

    python tests/golden/make_small_cycle_golden.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from sms_b200 import instances as I  # noqa: E402

DATA = "/root/reference/test/data"
FILES = ["ca-netscience.mc", "ego-facebook.mc", "tripleSquareTriangle.mc", "be120.3.5.mc", "bqp250-3.mc"]


def digest(a):
    return hashlib.blake2b(np.ascontiguousarray(a, np.int32).tobytes(), digest_size=8).hexdigest()


def main():
    from oracle import oracle_py
    oracle_py.build(ref=True)
    ref = oracle_py.ref()
    for f in FILES:
        inst, w = I.read_mc_weighted(os.path.join(DATA, f))
        tri, holes = ref.small_ccs(inst)
        x = I.xstar_lp_like(inst.m, seed=5) if "be120" in f else I.xstar_uniform(inst.m, seed=5)
        out = dict(n=np.int64(inst.n), eu=inst.eu, ev=inst.ev, x=x, ntri=np.int64(len(tri)), nholes=np.int64(len(holes)),
                   tri_hash=np.array(digest(tri)), holes_hash=np.array(digest(holes)))
        if tri.size <= 60000:
            out["tri"] = tri
        if holes.size <= 60000:
            out["holes"] = holes
        vt = ref.small_cycle_scan(inst, x, tri) if len(tri) else []
        out["viol_tri_nodes"] = np.asarray([n for _, n, _ in vt], np.int32).reshape(-1, 3)
        out["viol_tri_stay"] = np.asarray([s for _, _, s in vt], np.uint8).reshape(-1, 3)
        hs = holes[:20000]
        vh = ref.small_cycle_scan(inst, x, hs) if len(hs) else []
        out["viol_hole_nodes"] = np.asarray([n for _, n, _ in vh], np.int32).reshape(-1, 4)
        out["viol_hole_stay"] = np.asarray([s for _, _, s in vh], np.uint8).reshape(-1, 4)
        path = os.path.join(HERE, "smallcc_" + f.replace(".mc", "") + ".npz")
        np.savez_compressed(path, **out)
        print(f"{f}: n={inst.n} m={inst.m} triangles={len(tri)} holes={len(holes)} violated {len(vt)} / {len(vh)} -> {os.path.getsize(path)} B")


if __name__ == "__main__":
    main()
