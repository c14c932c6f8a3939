"""Host-side mirror of the reference's separator interface over the C ABI.

``GpuOcwSeparator`` has the public surface of ``sms::OcwSeparator``
(reference: include/sms/graph/ocw_seperator.hpp:14-41) -- updateWeights / updateEdgeWeight / updateBlacklist /
getViolatedOcws -- plus the batch call ``separateRound`` that ConshdlrOscw::sepaOddCycles
(src/conshdlr/conshdlr_oscw.cpp:165-247) uses instead of its per-vertex loop.  ``GpuTriangleSeparator`` mirrors
``sms::TriangleSeparator`` (include/sms/conshdlr/triangles.hpp:12-54) and ``sms::HoleSeparator``.
The C++ twin of this file (what a maintainer links into SMS) is sms_b200/host/gpu_ocw_separator.hpp.

No CPU path: every method ends in libsmsgpu.so.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .capi import Context, Cuts


@dataclass
class OddClosedWalk:
    """sms::OddClosedWalk (include/sms/graph/odd_closed_walk.hpp:21-123): nodes_[0..size] closed, isStay_[0..size)."""
    nodes: tuple
    stay: tuple
    eids: tuple = ()

    def size(self) -> int:
        return len(self.stay)

    def getIthNode(self, i: int) -> int:
        return self.nodes[i % len(self.stay)] if i == len(self.stay) else self.nodes[i]

    def ithEdgeIsStay(self, i: int) -> bool:
        return bool(self.stay[i])

    def numCrosses(self) -> int:
        return len(self.stay) - int(sum(self.stay))

    def isValid(self) -> bool:
        return self.numCrosses() % 2 == 1


def _walks(cuts: Cuts):
    o = cuts.off
    return [OddClosedWalk(tuple(cuts.node[o[k]:o[k + 1]].tolist()), tuple(cuts.is_stay[o[k]:o[k + 1]].tolist()),
                          tuple(cuts.eid[o[k]:o[k + 1]].tolist())) for k in range(cuts.ncuts)]


class GpuOcwSeparator:
    def __init__(self, n: int, rowptr, col, eid, device: int = 0):
        """OcwSeparator::OcwSeparator (src/graph/ocw_seperator.cpp:9-23): uploads the CSR once."""
        self.ctx = Context(n, rowptr, col, eid, device)
        self.n, self.m = self.ctx.n, self.ctx.m
        self._rowptr = np.asarray(rowptr)
        self._col = np.asarray(col)
        self._eid = np.asarray(eid)
        self.weights_ = np.zeros(self.m, dtype=np.float64)
        self.blacklist_ = np.zeros(self.n, dtype=np.uint8)
        self._dirty = True

    # --- the reference's per-round setters -------------------------------------------------------------------
    def updateWeights(self, u: int, v: int, w: float):
        """ocw_seperator.cpp:25-31 (edgeId = first adjacency entry of u holding v)"""
        row = self._col[self._rowptr[u]:self._rowptr[u + 1]]
        hit = np.nonzero(row == v)[0]
        if hit.size == 0:
            raise RuntimeError("Edge does not exist")
        self.updateEdgeWeight(int(self._eid[self._rowptr[u] + hit[0]]), w)

    def updateEdgeWeight(self, idx: int, ew: float):
        """ocw_seperator.cpp:33-35; the clamp itself happens on the device (K1)"""
        self.weights_[idx] = ew
        self._dirty = True

    def updateBlacklist(self, u: int, blacklisted: bool):
        """ocw_seperator.hpp:25-28"""
        self.blacklist_[u] = 1 if blacklisted else 0
        self._dirty = True

    def setRoundInput(self, x_by_eid, blacklist):
        """all m updateEdgeWeight + n updateBlacklist calls of conshdlr_oscw.cpp:175-185 at once"""
        self.weights_ = np.ascontiguousarray(x_by_eid, dtype=np.float64).copy()
        self.blacklist_ = np.ascontiguousarray(blacklist, dtype=np.uint8).copy()
        self._dirty = True

    def _flush(self):
        if self._dirty:
            self.ctx.set_round_input(self.weights_, self.blacklist_)
            self._dirty = False

    # --- queries ---------------------------------------------------------------------------------------------
    def getViolatedOcws(self, source: int):
        """ocw_seperator.cpp:330-361 for ONE source (doubled id), raw list in the reference's emission order"""
        self._flush()
        self.ctx.set_sources(np.array([source], dtype=np.int32))
        return _walks(self.ctx.separate(dedup=0))

    def separateRound(self, sources=None, dedup: bool = True, rank: int = 0, nranks: int = 1, **params) -> Cuts:
        """The whole loop of conshdlr_oscw.cpp:188-195 in one call (independent-source mode)."""
        self._flush()
        self.ctx.set_sources(sources, rank, nranks)
        return self.ctx.separate(dedup=1 if dedup else 0, **params)

    def close(self):
        self.ctx.close()


class GpuSmallCycleSeparator:
    """TriangleSeparator (triangles.hpp:12-54) for k = 3, HoleSeparator (holes.hpp) for k = 4."""

    def __init__(self, ocw: GpuOcwSeparator, k: int):
        if k not in (3, 4):
            raise ValueError("k must be 3 or 4")
        self.ocw, self.k = ocw, k
        self.cycles = []
        self._registered = False

    def addTriangle(self, u, v, w):
        assert self.k == 3
        self.cycles.append((u, v, w))
        self._registered = False

    def addFourHole(self, a, b, c, d):
        assert self.k == 4
        self.cycles.append((a, b, c, d))
        self._registered = False

    def setCycles(self, cycles):
        self.cycles = [tuple(int(v) for v in c) for c in np.asarray(cycles).reshape(-1, self.k)]
        self._registered = False

    def updateWeights(self, u, v, w):
        self.ocw.updateWeights(u, v, w)

    def getViolated(self):
        """getViolatedTriangles (triangles.hpp:34-53) / getViolatedFourHoles (holes.hpp:38-54)"""
        self.ocw._flush()
        if not self._registered:
            arr = np.asarray(self.cycles, dtype=np.int32).reshape(-1, self.k)
            self.ocw.ctx.small_cycles_set(arr)
            self._registered = True
        return self.ocw.ctx.small_cycles_scan(self.k)

    getViolatedTriangles = getViolated
    getViolatedFourHoles = getViolated
