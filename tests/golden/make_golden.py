#!/usr/bin/env python
"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libsmsref.so, compiled from
/root/reference by oracle/Makefile).  Run in the CPU container, where /root/reference exists:

    python tests/golden/make_golden.py

Each fixture holds one separation round of the reference's own OcwSeparator (src/graph/ocw_seperator.cpp:330-361,
driven like ConshdlrOscw::sepaOddCycles, src/conshdlr/conshdlr_oscw.cpp:188-195): the instance as an edge list, the
round's x* by edge id, the blacklist, the source list, and what the reference returned -- the raw walk list in
emission order (cut_off / cut_src / node / stay) in mode I (blacklist re-initialised before every source; the
contract of the CUDA path) and in mode V (the verbatim sequential sweep), plus the labels (dist, hops) the reference
held for every source after its search (mode I).  x* is tie-free (iid uniform doubles) so that the reference's
heap-internal tie order cannot matter; the LP-like fixtures contain exact 0/1 values and therefore ties: for those
only properties that do not depend on tie order are stored (the set of canonical cycles is NOT tie independent, the
per-source minimum walk length is).

The fixtures are small (a few hundred KB in total) and are committed; the GPU box has no /root/reference.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from sms_b200 import instances as I  # noqa: E402

CASES = {
    # name: (instance builder, x* builder, tie free?)
    "torus2d_10_uniform": (lambda: I.torus2d(10), lambda m: I.xstar_uniform(m, seed=1), True),
    "torus2d_16x9_uniform": (lambda: I.torus2d(16, 9), lambda m: I.xstar_uniform(m, seed=2), True),
    "torus3d_6_uniform": (lambda: I.torus3d(6), lambda m: I.xstar_uniform(m, seed=3), True),
    "random_400_uniform": (lambda: I.random_graph(400, 1600, seed=3), lambda m: I.xstar_uniform(m, seed=4), True),
    "dense_150_uniform": (lambda: I.random_graph(150, 1500, seed=5), lambda m: I.xstar_uniform(m, seed=5), True),
    "torus2d_30_uniform": (lambda: I.torus2d(30), lambda m: I.xstar_uniform(m, seed=6), True),
    # chord repair at work, but few enough good chords per cycle (<= 16) that the reference's std::sort of the chord list
    # (introsort: insertion sort up to 16 elements, i.e. stable) leaves the emission order well defined -- unlike dense_150
    "random_200_1300_uniform": (lambda: I.random_graph(200, 1300, seed=9), lambda m: I.xstar_uniform(m, seed=9), True),
    "torus2d_12_lp_like": (lambda: I.torus2d(12), lambda m: I.xstar_lp_like(m, seed=7), False),
    "random_400_lp_like": (lambda: I.random_graph(400, 1600, seed=3), lambda m: I.xstar_lp_like(m, seed=8), False),
}


def min_len_per_source(inst, x, sweep, nsrc):
    from helpers import walk_length
    out = np.full(nsrc, np.inf)
    for src, nodes, stay in sweep.walks():
        out[src] = min(out[src], walk_length(inst, x, nodes, stay))
    return out


def main():
    from oracle import oracle_py
    oracle_py.build(ref=True)
    ref = oracle_py.ref()
    for name, (mk, mkx, tie_free) in CASES.items():
        inst = mk()
        x = mkx(inst.m)
        bl = inst.root_blacklist()
        src = inst.handler_sources()
        sI = ref.sweep(inst, x, blacklist=bl, sources=src, mode="I", want_labels=True)
        sV = ref.sweep(inst, x, blacklist=bl, sources=src, mode="V")
        out = dict(n=np.int64(inst.n), eu=inst.eu, ev=inst.ev, x=x, blacklist=bl, sources=src,
                   tie_free=np.bool_(tie_free))
        if tie_free:
            out.update(I_cut_off=sI.cut_off, I_cut_src=sI.cut_src, I_node=sI.node, I_stay=sI.stay,
                       V_cut_off=sV.cut_off, V_cut_src=sV.cut_src, V_node=sV.node, V_stay=sV.stay,
                       )
            if len(sI.label_node) <= 20000:  # labels (mode I) only where they stay small
                out.update(label_off=sI.label_off, label_node=sI.label_node, label_dist=sI.label_dist,
                           label_hops=sI.label_hops)
        else:
            out.update(I_has_walk=np.bincount(sI.cut_src, minlength=len(src)) > 0,
                       I_min_len=min_len_per_source(inst, x, sI, len(src)))
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(f"{name}: n={inst.n} m={inst.m} walks I={sI.ncuts} V={sV.ncuts} -> {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()
