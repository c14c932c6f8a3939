#!/usr/bin/env python
"""Generates tests/golden/lpseq_*.npz: GENUINE root-LP iterates (SURVEY.md 8d input D3) and what the UNMODIFIED reference
separator returns on them.  Run in the CPU container (needs /root/reference for oracle/_ref and for pm1s_100.3.mc):

    python tests/golden/make_lp_golden.py [--from-logs DIR]

For every instance the root cutting loop of scripts/cutting_loop.py is run with HiGHS as the LP solver and the reference
separator in its own sequential mode (V) as the cut generator -- i.e. the x* sequence is the one `sms` itself would walk
through if SCIP's LP were HiGHS and every efficacious cut were kept.  Stored per instance:

  n, eu, ev, w          the instance (edge list in insertion order, weights by edge id)
  xs[K, m]              x* of every LP solve, by edge id, bit-exact float64 (bench.py --xstar replay feeds them in order)
  bounds[K]             LP objective per solve
  pick[P]               rounds selected for the parity tests
  for each picked round r (LP points are full of exact 0 / 1 / 0.5 values, i.e. of (dist, hops) ties whose order inside
  the reference's binary heap is an implementation artefact; only tie-INDEPENDENT facts of the reference are stored):
    I_has_walk_r[nsrc]  mode I: the source returned at least one walk
    I_min_len_r[nsrc]   mode I: length of the shortest walk the source returned (inf if none)
    counts_r            [walks mode I, walks mode V, unique canonical cycles mode I, unique mode V]

The fixtures are a few hundred KB each and are committed; the GPU box has no /root/reference.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))

import cutting_loop as CL  # noqa: E402
from helpers import canonical_set, walk_length  # noqa: E402

CASES = {
    # name: (instance spec, max LP solves)
    "torus2d_10": ("torus2d:10", 50),
    "pm1s_100.3": ("mc:/root/reference/test/data/pm1s_100.3.mc", 50),
    "torus2d_100": ("torus2d:100", 30),
    "torus3d_20": ("torus3d:20", 12),
    "random20k": ("random:20000:80000:7", 12),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--from-logs", default=None, help="directory with <name>.npz logs of scripts/cutting_loop.py --log (skips the loop)")
    ap.add_argument("--only", default=None)
    args = ap.parse_args()
    from oracle import oracle_py
    oracle_py.build(ref=True)
    ref = oracle_py.ref()
    for name, (spec, rounds) in CASES.items():
        if args.only and name not in args.only.split(","):
            continue
        inst, w = CL.load_instance(spec)
        log = os.path.join(args.from_logs, name + ".npz") if args.from_logs else None
        if log and os.path.exists(log):
            g = np.load(log)
            xs, bounds = g["xs"], g["bounds"]
            assert int(g["n"]) == inst.n and np.array_equal(g["eu"], inst.eu) and np.array_equal(g["w"], w)
        else:
            sep = CL.OracleSeparator(inst, "ref", "V")
            out = CL.run_loop(inst, w, sep, rounds)
            xs, bounds = np.stack(out["xs"]), np.asarray(out["bounds"])
        K = xs.shape[0]
        last = K - 2 if K >= 2 else 0  # the last solve that still produced cuts (the final one found none unless capped)
        pick = sorted({0, min(1, K - 1), min(2, K - 1), K // 2, max(last, 0), K - 1})
        if inst.n > 5000:
            pick = sorted({0, min(2, K - 1), K - 1})
        bl = inst.root_blacklist()
        src = inst.handler_sources()
        out = dict(n=np.int64(inst.n), eu=inst.eu, ev=inst.ev, w=w, xs=xs, bounds=bounds, pick=np.asarray(pick, np.int64), blacklist=bl)
        for r in pick:
            x = xs[r]
            sI = ref.sweep(inst, x, blacklist=bl, sources=src, mode="I")
            sV = ref.sweep(inst, x, blacklist=bl, sources=src, mode="V")
            has = np.bincount(sI.cut_src, minlength=len(src)) > 0
            mn = np.full(len(src), np.inf)
            for s, nodes, stay in sI.walks():
                mn[s] = min(mn[s], walk_length(inst, x, nodes, stay))
            out[f"I_has_walk_{r}"] = has
            out[f"I_min_len_{r}"] = mn
            out[f"counts_{r}"] = np.asarray([sI.ncuts, sV.ncuts, len(canonical_set(sI.walks())), len(canonical_set(sV.walks()))], np.int64)
            print(f"{name} round {r}: LP bound {bounds[r]:.4f}, integral {np.mean((x == 0) | (x == 1)):.3f}, "
                  f"walks I {sI.ncuts} / V {sV.ncuts}, unique I {out[f'counts_{r}'][2]} / V {out[f'counts_{r}'][3]}", flush=True)
        path = os.path.join(HERE, f"lpseq_{name}.npz")
        np.savez_compressed(path, **out)
        print(f"{name}: {K} LP solves, picked {pick} -> {os.path.getsize(path)} bytes", flush=True)


if __name__ == "__main__":
    main()
