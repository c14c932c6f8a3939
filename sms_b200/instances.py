"""Synthetic MaxCut separation instances of the shapes BASELINE.json names, in the reference's numbering.

Only numpy.  The separator never looks at the +-J couplings, only at the graph and the LP point x*
(reference: src/conshdlr/conshdlr_oscw.cpp:174-185), so an "instance" here is
(n, edge list in insertion order, x* by edge id, initial blacklist).

Numbering contract:
  * adjacency order of a vertex = order of the addEdge calls touching it
    (reference: extern/networkit/networkit/cpp/graph/Graph.cpp:472-520)
  * edge id = running counter over u ascending, adjacency position ascending, taken where u >= v
    (Graph.cpp:131-170, Graph::indexEdges)
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


@dataclass
class Instance:
    name: str
    n: int
    eu: np.ndarray  # int32[m], insertion order
    ev: np.ndarray  # int32[m]
    rowptr: np.ndarray = field(default=None)  # int64[n+1]
    col: np.ndarray = field(default=None)  # int32[2m]
    eid: np.ndarray = field(default=None)  # int32[2m]

    @property
    def m(self) -> int:
        return int(self.eu.shape[0])

    def root(self) -> int:
        """maxDegreeNode as BasicOcwSolver picks the root (reference: src/solver/basic_ocw.cpp:123-126)."""
        deg = np.diff(self.rowptr)
        return int(np.argmax(deg))

    def root_blacklist(self) -> np.ndarray:
        bl = np.zeros(self.n, dtype=np.uint8)
        bl[self.root()] = 1
        return bl

    def handler_sources(self) -> np.ndarray:
        """ConshdlrOscw iterates nodeRange() and passes the ORIGINAL id, which the separator reads as a
        double-cover id (reference: conshdlr_oscw.cpp:188-195)."""
        return np.arange(self.n, dtype=np.int32)


def csr_from_edge_list(n: int, eu: np.ndarray, ev: np.ndarray):
    """Vectorised NetworKit-order CSR + edge ids (see module docstring)."""
    eu = np.ascontiguousarray(eu, dtype=np.int64)
    ev = np.ascontiguousarray(ev, dtype=np.int64)
    m = eu.shape[0]
    if m and (np.any(eu == ev)):
        raise ValueError("self loops are not supported")
    src = np.concatenate([eu, ev])
    dst = np.concatenate([ev, eu])
    e_of = np.concatenate([np.arange(m), np.arange(m)])
    order = np.lexsort((e_of, src))  # by src, then by insertion index
    src_s, dst_s, e_s = src[order], dst[order], e_of[order]
    rowptr = np.zeros(n + 1, dtype=np.int64)
    np.add.at(rowptr, src_s + 1, 1)
    rowptr = np.cumsum(rowptr)
    owner = src_s >= dst_s
    ids = np.cumsum(owner) - 1
    id_of_edge = np.empty(m, dtype=np.int64)
    id_of_edge[e_s[owner]] = ids[owner]
    eid = id_of_edge[e_s]
    return rowptr.astype(np.int64), dst_s.astype(np.int32), eid.astype(np.int32)


def _finish(name: str, n: int, eu, ev) -> Instance:
    inst = Instance(name, n, np.ascontiguousarray(eu, dtype=np.int32), np.ascontiguousarray(ev, dtype=np.int32))
    inst.rowptr, inst.col, inst.eid = csr_from_edge_list(n, inst.eu, inst.ev)
    return inst


def from_edges(name: str, n: int, edges) -> Instance:
    e = np.asarray(edges, dtype=np.int32).reshape(-1, 2)
    return _finish(name, n, e[:, 0], e[:, 1])


def torus2d(L: int, W: int | None = None) -> Instance:
    """L x W toroidal grid, node id r*W+c; per node: edge to the right neighbour, then to the lower one."""
    W = L if W is None else W
    r, c = np.meshgrid(np.arange(L), np.arange(W), indexing="ij")
    u = (r * W + c).ravel()
    right = (r * W + (c + 1) % W).ravel()
    down = (((r + 1) % L) * W + c).ravel()
    eu = np.stack([u, u], axis=1).ravel()
    ev = np.stack([right, down], axis=1).ravel()
    return _finish(f"torus2d_{L}x{W}", L * W, eu, ev)


def torus3d(L: int) -> Instance:
    """L^3 toroidal grid, id (z*L+y)*L+x; per node edges +x, +y, +z."""
    z, y, x = np.meshgrid(np.arange(L), np.arange(L), np.arange(L), indexing="ij")
    u = ((z * L + y) * L + x).ravel()
    px = ((z * L + y) * L + (x + 1) % L).ravel()
    py = ((z * L + (y + 1) % L) * L + x).ravel()
    pz = ((((z + 1) % L) * L + y) * L + x).ravel()
    eu = np.stack([u, u, u], axis=1).ravel()
    ev = np.stack([px, py, pz], axis=1).ravel()
    return _finish(f"torus3d_{L}", L**3, eu, ev)


def random_graph(n: int, m: int, seed: int = 7) -> Instance:
    """m distinct uniform random vertex pairs (no loops, no multi-edges), kept in sampling order."""
    rng = np.random.Generator(np.random.PCG64(seed))
    have = np.empty((0, 2), dtype=np.int64)
    while have.shape[0] < m:
        need = m - have.shape[0]
        a = rng.integers(0, n, size=int(need * 1.1) + 16)
        b = rng.integers(0, n, size=a.shape[0])
        keep = a != b
        cand = np.stack([a[keep], b[keep]], axis=1)
        allp = np.concatenate([have, cand])
        key = np.minimum(allp[:, 0], allp[:, 1]) * n + np.maximum(allp[:, 0], allp[:, 1])
        _, first = np.unique(key, return_index=True)
        have = allp[np.sort(first)][:m]
    return _finish(f"random_n{n}_m{m}_s{seed}", n, have[:, 0], have[:, 1])


def read_mc(path: str) -> Instance:
    """'.mc' format: header 'n m', then 1-based 'u v w' lines (reference: src/io/io.cpp:268-273)."""
    with open(path) as f:
        rows = [ln.split() for ln in f if ln.strip() and not ln.startswith("#")]
    n, m = int(rows[0][0]), int(rows[0][1])
    e = np.array([[int(r[0]) - 1, int(r[1]) - 1] for r in rows[1 : 1 + m]], dtype=np.int64)
    return _finish(path.split("/")[-1], n, e[:, 0], e[:, 1])


def read_mc_weighted(path: str):
    """like read_mc, plus the weights by edge-list position: (Instance, float64[m])"""
    with open(path) as f:
        rows = [ln.split() for ln in f if ln.strip() and not ln.startswith("#")]
    n, m = int(rows[0][0]), int(rows[0][1])
    e = np.array([[int(r[0]) - 1, int(r[1]) - 1] for r in rows[1 : 1 + m]], dtype=np.int64)
    w = np.array([float(r[2]) for r in rows[1 : 1 + m]], dtype=np.float64)
    return _finish(path.split("/")[-1], n, e[:, 0], e[:, 1]), w


def write_mc(inst: Instance, path: str, weights=None, seed: int = 1) -> None:
    """Writes `inst` in the '.mc' format the reference's `sms` binary reads (src/io/io.cpp: header 'n m', then
    1-based 'u v w' lines; mcFileToGraph adds the edges in file order, so NetworKit's adjacency order and edge ids
    equal this instance's).  `weights` by edge-list position; default: +-1 couplings from mt19937_64(seed)."""
    if weights is None:
        rng = np.random.Generator(np.random.MT19937(seed))
        weights = np.where(rng.integers(0, 2, size=inst.m) == 1, 1, -1)
    weights = np.asarray(weights)
    if weights.shape[0] != inst.m:
        raise ValueError("one weight per edge")
    integral = np.all(weights == np.round(weights))
    with open(path, "w") as f:
        f.write(f"{inst.n} {inst.m}\n")
        for u, v, w in zip(inst.eu.tolist(), inst.ev.tolist(), weights.tolist()):
            f.write(f"{u + 1} {v + 1} {int(w) if integral else repr(float(w))}\n")


# ----------------------------------------------------------------------------------------------------
# LP points x* (indexed by edge id)
# ----------------------------------------------------------------------------------------------------
def xstar_uniform(m: int, seed: int = 1) -> np.ndarray:
    """D1: iid U(0,1) -- tie-free with probability 1."""
    return np.random.Generator(np.random.PCG64(seed)).random(m)


def xstar_lp_like(m: int, seed: int = 1, p0: float = 0.4, p1: float = 0.4) -> np.ndarray:
    """D2: 40 % exactly 0, 40 % exactly 1, 20 % U(0,1) -- tie-heavy like a real LP vertex."""
    rng = np.random.Generator(np.random.PCG64(seed))
    kind = rng.random(m)
    x = rng.random(m)
    x[kind < p0] = 0.0
    x[(kind >= p0) & (kind < p0 + p1)] = 1.0
    return x


def xstar_half_integral(m: int, seed: int = 1) -> np.ndarray:
    """values in {0, 0.5, 1} plus a few quarters: the hardest tie case."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.choice(np.array([0.0, 0.25, 0.5, 0.5, 0.75, 1.0]), size=m)


def xstar_near_cut(inst: Instance, seed: int = 1, noise: float = 0.35, flip: float = 0.08) -> np.ndarray:
    """x* close to a cut vector: x_e = [side(u) != side(v)] moved towards 1/2 by |N(0, noise)|, with a
    fraction `flip` of the edges flipped.  Gives few, long violated cycles (sparser than D1)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    side = rng.integers(0, 2, size=inst.n)
    # value per edge id: take the owner arcs (src >= dst) which enumerate edge ids 0..m-1 in order
    src = np.repeat(np.arange(inst.n), np.diff(inst.rowptr))
    owner = src >= inst.col
    cut = (side[src[owner]] != side[inst.col[owner]]).astype(np.float64)
    x = np.empty(inst.m)
    x[inst.eid[owner]] = cut
    f = rng.random(inst.m) < flip
    x[f] = 1.0 - x[f]
    d = np.minimum(np.abs(rng.normal(0.0, noise, size=inst.m)), 1.0)
    return np.where(x > 0.5, 1.0 - 0.5 * d, 0.5 * d)
