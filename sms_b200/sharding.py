"""Host-side mirror of the multi-GPU plan of libsmsgpu (sms_b200/csrc/smsgpu.cu,
smsgpu_set_sources and the merge step of smsgpu_separate): which sources a rank owns, and how per-rank cut lists
combine into the list every rank returns.  Pure numpy -- used by bench.py / the torch.distributed harness to size and
check the sharded round, and exercised without a GPU by the world-size-2 gloo test (tests/test_sharding_gloo.py).

The reference has no counterpart: its sweep is one sequential loop over nodeRange()
(src/conshdlr/conshdlr_oscw.cpp:188-195).  Sharding is legal because sources are independent in mode I.
"""
from __future__ import annotations

import numpy as np


def local_count(nsrc: int, rank: int, nranks: int) -> int:
    """Number of sources rank owns: PAIRS (2p, 2p+1) of consecutive sources are dealt cyclically, so that the two sides
    of a vertex -- mirror-image searches -- stay on one GPU and share a warp there (smsgpu_set_sources)."""
    pairs_all = (nsrc + 1) // 2
    pairs_mine = (pairs_all - rank + nranks - 1) // nranks if pairs_all > rank else 0
    cnt = 2 * pairs_mine
    if pairs_mine > 0 and nsrc % 2 == 1 and (pairs_mine - 1) * nranks + rank == pairs_all - 1:
        cnt -= 1
    return cnt


def global_index(local_index, rank: int, nranks: int):
    """Position in the round's source list of a rank's local source index j: ((j >> 1) * nranks + rank) * 2 + (j & 1)."""
    j = np.asarray(local_index)
    return ((j >> 1) * nranks + rank) * 2 + (j & 1)


def local_sources(sources: np.ndarray, rank: int, nranks: int) -> np.ndarray:
    """The sources rank owns, in local index order (smsgpu_set_sources)."""
    if nranks < 1 or not 0 <= rank < nranks:
        raise ValueError(f"bad rank {rank} / {nranks}")
    src = np.asarray(sources)
    j = np.arange(local_count(len(src), rank, nranks))
    return np.ascontiguousarray(src[global_index(j, rank, nranks)])


def canonical_walk(nodes, stay):
    """Canonical representative of a closed walk, the rule of K5 (sms_b200/csrc/dedup_kernels.cuh): among the
    rotations that start at a minimum vertex, in both directions, the lexicographically smallest sequence of
    (vertex, stay flag of the edge leaving it)."""
    L = len(nodes)
    mn = min(nodes)
    best = None
    for p in range(L):
        if nodes[p] != mn:
            continue
        fwd = tuple((int(nodes[(p + k) % L]), int(stay[(p + k) % L])) for k in range(L))
        bwd = tuple((int(nodes[(p - k) % L]), int(stay[(p - k - 1) % L])) for k in range(L))
        for cand in (fwd, bwd):
            if best is None or cand < best:
                best = cand
    return best


def merge_walk_lists(per_rank, nranks: int):
    """per_rank[r] = list of (local source index, nodes, stay) in emission order, as rank r found them.
    Returns the merged, deduplicated list every rank ends up with after the all-gather: one entry per canonical
    cycle, (global source index of the first finder, nodes, stay) of that first finder, ordered by
    (first-finding global source index, emission index) -- the order of the single-GPU round."""
    tagged = []
    for r, walks in enumerate(per_rank):
        seq = {}
        for src, nodes, stay in walks:
            k = seq.get(src, 0)
            seq[src] = k + 1
            tagged.append(((int(global_index(src, r, nranks)), k), tuple(nodes), tuple(stay)))
    tagged.sort(key=lambda t: t[0])
    seen = set()
    out = []
    for (gsrc, _), nodes, stay in tagged:
        c = canonical_walk(nodes, stay)
        if c in seen:
            continue
        seen.add(c)
        out.append((gsrc, nodes, stay))
    return out
