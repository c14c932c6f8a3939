#!/usr/bin/env python
"""Root cutting loop over the edge LP of MaxCut with HiGHS standing in for SCIP's LP (SURVEY.md 8d input D3; BASELINE.json
configs 1 and 4 "root-node separation" / "full root cutting loop", config 2 "one separation round per LP solve").

SCIP is not available on either box, so the loop the reference runs at the root
    solve LP -> CONSSEPALP -> ConshdlrOscw::sepaOddCycles (src/conshdlr/conshdlr_oscw.cpp:165-247) -> rows -> re-solve
is reproduced with scipy.optimize.linprog(method="highs"):

    maximise  sum_e w_e x_e   s.t.  0 <= x <= 1,   sum_{cross} x_e - sum_{stay} x_e <= #cross - 1   for every cut found so far

* one separation round per LP solve; x* is indexed by NetworKit edge id, the blacklist is {root} (the max-degree vertex,
  src/solver/basic_ocw.cpp:123-126), sources are the handler's list 0..n-1;
* every returned cut whose efficacy (violation / sqrt(len), SCIP's definition) exceeds 1e-6 (separating/minefficacyroot,
  src/solver/basic_ocw.cpp:244) is added -- cut SELECTION differs from SCIP's (no pool ageing, no parallelism filter); that
  step is "parity unpinned" in the reference as well;
* the loop ends when a round finds no efficacious cut or after `max_rounds`.

The separator is pluggable: "ref" (the unmodified reference, mode V -- what `sms` itself would do; CPU container only),
"orc" (the CPU restatement, mode I or V, canonical ties), "gpu" (libsmsgpu.so through the C ABI).  This file is harness
code: it imports oracle/ only for the "ref"/"orc" separators (tests and golden generation), never for "gpu".

    python scripts/cutting_loop.py --instance torus2d:10 --sep ref --rounds 50 --log /tmp/t10.npz
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from sms_b200 import instances as I  # noqa: E402

MIN_EFFICACY_ROOT = 1e-6  # src/solver/basic_ocw.cpp:244


def pm1_weights(m: int, seed: int = 1) -> np.ndarray:
    """+-1 couplings by edge-list position from mt19937(seed) (same rule as instances.write_mc)"""
    rng = np.random.Generator(np.random.MT19937(seed))
    return np.where(rng.integers(0, 2, size=m) == 1, 1.0, -1.0)


def weights_by_eid(inst, w_by_pos) -> np.ndarray:
    """edge-list position -> NetworKit edge id"""
    src = np.repeat(np.arange(inst.n), np.diff(inst.rowptr))
    # the arc u -> v of edge (eu, ev) at list position p: look the id up through the CSR
    key = {}
    for a in range(inst.col.shape[0]):
        key[(int(src[a]), int(inst.col[a]))] = int(inst.eid[a])
    out = np.zeros(inst.m)
    for p, (u, v) in enumerate(zip(inst.eu.tolist(), inst.ev.tolist())):
        out[key[(u, v)]] = w_by_pos[p]
    return out


def load_instance(spec: str):
    """'torus2d:L' | 'torus3d:L' | 'random:n:m:seed' | 'mc:path'  ->  (Instance, weights by edge id)"""
    kind, *rest = spec.split(":")
    if kind == "torus2d":
        inst = I.torus2d(int(rest[0]))
        return inst, weights_by_eid(inst, pm1_weights(inst.m, 1))
    if kind == "torus3d":
        inst = I.torus3d(int(rest[0]))
        return inst, weights_by_eid(inst, pm1_weights(inst.m, 1))
    if kind == "random":
        inst = I.random_graph(int(rest[0]), int(rest[1]), seed=int(rest[2]) if len(rest) > 2 else 7)
        return inst, weights_by_eid(inst, pm1_weights(inst.m, 1))
    if kind == "mc":
        inst, w = I.read_mc_weighted(rest[0])
        return inst, weights_by_eid(inst, w)
    raise ValueError(spec)


# ---------------------------------------------------------------------------------------------------------------------
# separators: x*, blacklist -> list of (nodes, stay) closed walks
# ---------------------------------------------------------------------------------------------------------------------
class OracleSeparator:
    def __init__(self, inst, which="ref", mode="V", canonical=False):
        from oracle import oracle_py  # harness only
        self.eng = oracle_py.ref() if which == "ref" else oracle_py.orc()
        self.inst, self.mode, self.canonical = inst, mode, canonical
        self.seconds = 0.0

    def __call__(self, x, bl):
        kw = dict(canonical=True) if self.canonical else {}
        s = self.eng.sweep(self.inst, x, blacklist=bl, mode=self.mode, **kw)
        self.seconds += s.seconds
        return [(n, st) for _, n, st in s.walks()]

    def close(self):
        pass


class GpuSeparator:
    def __init__(self, inst, mode=0, device=0):
        from sms_b200 import capi
        self.ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=device)
        self.ctx.set_sources(None, 0, 1)
        self.mode = mode
        self.seconds = 0.0
        self.stats = []

    def __call__(self, x, bl):
        t0 = time.perf_counter()
        self.ctx.set_round_input(x, bl)
        cuts = self.ctx.separate(dedup=1, mode=self.mode, min_efficacy=MIN_EFFICACY_ROOT)
        self.seconds += time.perf_counter() - t0
        self.stats.append(self.ctx.stats())
        return [(n, st) for _, n, st in cuts.walks()]

    def close(self):
        self.ctx.close()


# ---------------------------------------------------------------------------------------------------------------------
def canonical(nodes, stay):
    L = len(nodes)
    mn = min(nodes)
    best = None
    for p in range(L):
        if nodes[p] != mn:
            continue
        fwd = tuple((nodes[(p + k) % L], stay[(p + k) % L]) for k in range(L))
        bwd = tuple((nodes[(p - k) % L], stay[(p - k - 1) % L]) for k in range(L))
        for c in (fwd, bwd):
            if best is None or c < best:
                best = c
    return best


def cut_row(inst, arc_of, c):
    """canonical closed walk -> (edge ids, coefficients, rhs): sum_cross x - sum_stay x <= #cross - 1"""
    L = len(c)
    eids, coef, ncross = [], [], 0
    for i in range(L):
        u, v, stay = c[i][0], c[(i + 1) % L][0], c[i][1]
        eids.append(arc_of[(u, v)])
        coef.append(-1.0 if stay else 1.0)
        ncross += 0 if stay else 1
    return eids, coef, float(ncross - 1)


def run_loop(inst, w_by_eid, separator, max_rounds=50, min_eff=MIN_EFFICACY_ROOT, verbose=False):
    """Returns a dict: xs (list of x* per LP solve), bounds (LP objective per solve), found / added per round."""
    from scipy.optimize import linprog
    from scipy.sparse import csr_matrix
    src = np.repeat(np.arange(inst.n), np.diff(inst.rowptr))
    arc_of = {(int(u), int(v)): int(e) for u, v, e in zip(src, inst.col, inst.eid)}
    bl = inst.root_blacklist()
    rows_i, rows_j, rows_v, rhs = [], [], [], []
    have = set()
    xs, bounds, found, added = [], [], [], []
    nrows = 0
    for rnd in range(max_rounds):
        A = csr_matrix((rows_v, (rows_i, rows_j)), shape=(nrows, inst.m)) if nrows else None
        res = linprog(-w_by_eid, A_ub=A, b_ub=np.asarray(rhs) if nrows else None, bounds=(0.0, 1.0), method="highs")
        if res.status != 0:
            raise RuntimeError(f"HiGHS status {res.status}: {res.message}")
        x = np.ascontiguousarray(res.x, dtype=np.float64)
        xs.append(x)
        bounds.append(-float(res.fun))
        walks = separator(x, bl)
        xc = np.clip(x, 0.0, 1.0)
        new = 0
        for nodes, stay in walks:
            c = canonical(nodes, stay)
            if c in have:
                continue
            eids, coef, r = cut_row(inst, arc_of, c)
            act = float(np.dot(coef, xc[eids]))
            if (act - r) / np.sqrt(len(eids)) > min_eff:  # SCIPisCutEfficacious at the root
                have.add(c)
                rows_i.extend([nrows] * len(eids))
                rows_j.extend(eids)
                rows_v.extend(coef)
                rhs.append(r)
                nrows += 1
                new += 1
        found.append(len(walks))
        added.append(new)
        if verbose:
            print(f"round {rnd:3d}: LP bound {bounds[-1]:.6f}, {len(walks)} walks, {new} new cuts, {nrows} rows", flush=True)
        if new == 0:
            break
    return {"xs": xs, "bounds": bounds, "found": found, "added": added, "rows": nrows}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--instance", default="torus2d:10")
    ap.add_argument("--sep", default="ref", choices=["ref", "orc", "orcI", "gpu", "gpuV"])
    ap.add_argument("--rounds", type=int, default=50)
    ap.add_argument("--log", default=None, help="write the x* sequence (npz)")
    args = ap.parse_args()
    inst, w = load_instance(args.instance)
    if args.sep == "ref":
        sep = OracleSeparator(inst, "ref", "V")
    elif args.sep == "orc":
        sep = OracleSeparator(inst, "orc", "V", canonical=True)
    elif args.sep == "orcI":
        sep = OracleSeparator(inst, "orc", "I", canonical=True)
    else:
        sep = GpuSeparator(inst, mode=1 if args.sep == "gpuV" else 0)
    t0 = time.perf_counter()
    out = run_loop(inst, w, sep, args.rounds, verbose=True)
    print(f"{args.instance}: {len(out['xs'])} LP solves, final bound {out['bounds'][-1]:.6f}, {out['rows']} rows, "
          f"separation {sep.seconds:.3f} s of {time.perf_counter() - t0:.1f} s")
    sep.close()
    if args.log:
        np.savez_compressed(args.log, n=np.int64(inst.n), eu=inst.eu, ev=inst.ev, w=w, xs=np.stack(out["xs"]),
                            bounds=np.asarray(out["bounds"]), found=np.asarray(out["found"]), added=np.asarray(out["added"]))


if __name__ == "__main__":
    main()
