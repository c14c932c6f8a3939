"""Shared helpers of the parity tests (test infrastructure; may use oracle/)."""
from __future__ import annotations

import numpy as np


def canonical(nodes, stay):
    """Canonical form of a closed walk: among rotations starting at a minimum vertex, both directions, the
    lexicographically smallest sequence of (vertex, stay flag of the edge leaving it).  Same rule as K5
    (sms_b200/csrc/dedup_kernels.cuh)."""
    L = len(nodes)
    mn = min(nodes)
    best = None
    for p in range(L):
        if nodes[p] != mn:
            continue
        fwd = tuple((nodes[(p + k) % L], stay[(p + k) % L]) for k in range(L))
        bwd = tuple((nodes[(p - k) % L], stay[(p - k - 1) % L]) for k in range(L))
        for cand in (fwd, bwd):
            if best is None or cand < best:
                best = cand
    return best


def canonical_set(walks):
    """walks: iterable of (src, nodes, stay) -> set of canonical forms"""
    return {canonical(n, s) for _, n, s in walks}


def first_finder(walks):
    """canonical form -> smallest (src, emission index) that produced it; walks in (src, emission) order"""
    out = {}
    seq = {}
    for src, n, s in walks:
        k = seq.get(src, 0)
        seq[src] = k + 1
        c = canonical(n, s)
        if c not in out or (src, k) < out[c]:
            out[c] = (src, k)
    return out


def walk_length(inst, x, nodes, stay):
    """sum of stay ? x : 1-x along the closed walk (x clamped)"""
    xc = np.minimum(np.maximum(0.0, x), 1.0)
    tot = 0.0
    L = len(nodes)
    for i in range(L):
        u, v = nodes[i], nodes[(i + 1) % L]
        row = inst.col[inst.rowptr[u]:inst.rowptr[u + 1]]
        a = int(np.nonzero(row == v)[0][0]) + int(inst.rowptr[u])
        w = xc[inst.eid[a]]
        tot += w if stay[i] else 1 - w
    return tot
