#!/usr/bin/env python
"""Measured differences between the two sweep contracts:

  V  the reference's own sequential sweep (blacklist side effects carried from source to source)
  I  independent sources (every source sees the round's initial blacklist)

on configs 2-4 with iid U(0,1) x*, the synthetic LP-like mixture and GENUINE LP iterates of the HiGHS root cutting loop
(tests/golden/lpseq_*.npz).  Columns: raw walks and unique canonical cycles of each mode, V minus I, I minus V, the LP
bound after adding each set to the edge LP the point came from (HiGHS; lower = tighter), sweeps of the fixed-point scheme
the device uses for mode V, and how far canonical tie-breaking moves the set away from the reference's own tie order
(|canonical set  xor  reference-tie set|, same mode).  CPU only (oracle/): test infrastructure, never the product path.

    python scripts/mode_table.py [--quick] > profiles/r02_mode_table.md
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))

import cutting_loop as CL  # noqa: E402
from helpers import canonical_set  # noqa: E402
from sms_b200 import instances as I  # noqa: E402


def lp_bound(inst, w, base_cuts, extra):
    """LP bound of the edge LP with the cuts `base_cuts` + `extra` (sets of canonical walks)"""
    from scipy.optimize import linprog
    from scipy.sparse import csr_matrix
    src = np.repeat(np.arange(inst.n), np.diff(inst.rowptr))
    arc_of = {(int(u), int(v)): int(e) for u, v, e in zip(src, inst.col, inst.eid)}
    ri, rj, rv, rhs = [], [], [], []
    for k, c in enumerate(list(base_cuts) + list(extra)):
        eids, coef, r = CL.cut_row(inst, arc_of, c)
        ri.extend([k] * len(eids))
        rj.extend(eids)
        rv.extend(coef)
        rhs.append(r)
    A = csr_matrix((rv, (ri, rj)), shape=(len(rhs), inst.m)) if rhs else None
    res = linprog(-w, A_ub=A, b_ub=np.asarray(rhs) if rhs else None, bounds=(0.0, 1.0), method="highs")
    return -float(res.fun)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true", help="skip the 20k-node instance")
    args = ap.parse_args()
    from oracle import oracle_py
    orc = oracle_py.orc()
    rows = []
    cases = [("config 2: torus 100x100", I.torus2d(100), "lpseq_torus2d_100.npz"), ("config 3: torus 20^3", I.torus3d(20), "lpseq_torus3d_20.npz")]
    if not args.quick:
        cases.append(("config 4: random 20k / 80k", I.random_graph(20000, 80000, seed=7), "lpseq_random20k.npz"))
    cases.append(("config 1: pm1s_100.3.mc", None, "lpseq_pm1s_100.3.npz"))
    for title, inst, seqfile in cases:
        pts = []
        g = None
        p = os.path.join(ROOT, "tests", "golden", seqfile)
        if os.path.exists(p):
            g = np.load(p)
            if inst is None:
                inst = I.from_edges(title, int(g["n"]), np.stack([g["eu"], g["ev"]], axis=1))
        if inst is None:
            continue
        if "pm1s" not in title:
            pts += [("iid U(0,1)", I.xstar_uniform(inst.m, seed=1), None), ("LP-like 40/40/20", I.xstar_lp_like(inst.m, seed=1), None)]
        if g is not None:
            for r in g["pick"].tolist():
                if r == len(g["xs"]) - 1 and r > 0 and "pm1s" in title:
                    continue
                pts.append((f"root LP solve {r}", g["xs"][r], r))
        bl = inst.root_blacklist()
        for xname, x, r in pts:
            t0 = time.time()
            sV = orc.sweep(inst, x, blacklist=bl, mode="V", canonical=True)
            sI = orc.sweep(inst, x, blacklist=bl, mode="I", canonical=True)
            cV, cI = canonical_set(sV.walks()), canonical_set(sI.walks())
            rV = canonical_set(orc.sweep(inst, x, blacklist=bl, mode="V", canonical=False).walks())
            rI = canonical_set(orc.sweep(inst, x, blacklist=bl, mode="I", canonical=False).walks())
            # sweeps of the stamped fixed point (CPU statement, one window); costs sweeps x a whole round: small instances only,
            # the device's own counts for the large ones are in the bench lines (verbatim_sweeps)
            info = orc.sweep_fixpoint(inst, x, blacklist=bl, canonical=True)[1] if inst.n <= 1000 else {"sweeps": None}
            bV = bI = b0 = None
            if g is not None and r is not None and inst.n <= 10000:
                # the LP this point came from is not stored; the bound is taken over the cuts of THIS round only, added
                # to the bare edge LP started from x's objective -- comparable between V and I, not with the loop's bound
                w = g["w"]
                b0 = float(g["bounds"][r])
                bV = lp_bound(inst, w, [], cV)
                bI = lp_bound(inst, w, [], cI)
            rows.append((title, xname, sV.ncuts, sI.ncuts, len(cV), len(cI), len(cV - cI), len(cI - cV), info["sweeps"],
                         len(cV ^ rV), len(cI ^ rI), b0, bV, bI))
            print(f"# {title} / {xname}: {time.time() - t0:.1f} s", file=sys.stderr, flush=True)
    print("| instance | x* | raw walks V | raw walks I | unique V | unique I | V minus I | I minus V | sweeps to the fixed point (mode V) | "
          "canonical vs reference ties, V (sym. diff.) | same, I | LP bound of the point | bare edge LP + this round's V cuts | + this round's I cuts |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    for r in rows:
        f = lambda v: "-" if v is None else (f"{v:.3f}" if isinstance(v, float) else str(v))  # noqa: E731
        print("| " + " | ".join(f(v) for v in r) + " |")


if __name__ == "__main__":
    main()
