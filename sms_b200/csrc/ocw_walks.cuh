// K3/K4: candidate selection, walk extraction, simple-cycle split and chord repair for one source, run by the
// tile right after its search while the label table is still hot.
//   selectIntermediates  <- std::sort of intermediateVertices_ + "first 10" (ocw_seperator.cpp:115-117, 343-346)
//   buildWalk            <- getShortestPath + intermediateVertexToOcw      (ocw_seperator.cpp:120-188)
//   simpleOddCycles      <- OcwSeparator::simpleOddCycles                  (ocw_seperator.cpp:202-227)
//   extractCycles        <- OcwSeparator::extractCycles + OddClosedWalk::splitOnChord
//                                                   (ocw_seperator.cpp:229-328, odd_closed_walk.cpp:67-90)
// The predecessor chains are followed by lane 0 (buildWalk); everything that needs the graph afterwards is done with the
// lanes of the warp over independent items (extractCycles), because one warp visiting the positions of a 20-edge cycle
// one after the other pays two or three memory round trips per position.  The reference's O(n)
// scratch vectors (prevPos[n_], coveredNodes(n_)) become scans over the walk itself.
#pragma once
#include "ocw_search.cuh"

namespace smsgpu {

#ifdef SMSGPU_RUNS_CLOCKS // instrumentation build (ocw_search_runs.cuh): cycles of CTA 0 per phase
__device__ unsigned long long g_runClk[24];
#define WCLK(i_)                                                                         \
    do {                                                                                 \
        if (threadIdx.x == 0 && blockIdx.x == 0) {                                       \
            const long long now_ = clock64();                                            \
            atomicAdd(&g_runClk[i_], (unsigned long long)(now_ - wclkLast));             \
            wclkLast = now_;                                                             \
        }                                                                                \
    } while (0)
#define WCLK_DECL long long wclkLast = clock64()
#else
#define WCLK(i_) ((void)0)
#define WCLK_DECL ((void)0)
#endif

struct Emitter {
    const RawOut *R;
    int32_t srcIdx;
    int32_t seq;
    int32_t tier;
    // optional log of the cuts emitted (pairs engine: the twin source gets copies of them)
    long long *logCut = nullptr; // arena index of cut i, -1 if the arena was full
    int32_t *logLen = nullptr;
    int32_t nlog = 0, logCap = 0;
};

// reserve room for one cut; returns false (and still counts) when the arena is full
__device__ __forceinline__ bool emitReserve(Emitter &E, int len, long long &cut, long long &off) {
    cut = (long long)atomicAdd(&E.R->counters[0], 1ull);
    off = (long long)atomicAdd(&E.R->counters[1], (unsigned long long)len);
    const bool ok = cut < E.R->cutCap && off + len <= E.R->nodeCap;
    if (ok) {
        E.R->cutOff[cut] = (int32_t)off;
        E.R->cutLen[cut] = len;
        E.R->cutSrc[cut] = E.srcIdx;
        E.R->cutSeq[cut] = E.seq;
        E.R->cutTier[cut] = (int8_t)E.tier;
    }
    if (E.logCut) {
        if (E.nlog < E.logCap) {
            E.logCut[E.nlog] = ok ? cut : -1;
            E.logLen[E.nlog] = len;
        }
        E.nlog++;
    }
    E.seq++;
    return ok;
}

// ---- selection of the first maxInter intermediates by final (dist, hops), stable --------------------------------
// Writes the selected doubled ids to sel[0..nsel).  Uses the (now idle) queue memory as key scratch.
template <int TW>
__device__ int selectIntermediates(const Tile<TW> &T, TileMem &M, SrcState &S, int maxInter, uint32_t *sel) {
    const uint32_t H = S.histN;
    if (H > M.caps.heapCap) {
        S.overflow = 1;
        return 0;
    }
    for (uint32_t i = T.lane; i < H; i += TW) {
        const uint32_t t = M.hist[i];
        const int slot = findSlot(M, t >> 1);
        SMS_CHK(slot >= 0);
        const Entry e = loadEntry(&M.tab[slot]);
        HeapEnt k;
        k.hi = (unsigned long long)__double_as_longlong((t & 1u) ? e.d1 : e.d0);
        k.lo = ((unsigned long long)(((t & 1u) ? e.h1 : e.h0) & M.hopMask) << 32) | i; // history index: stable order
        M.heap[i] = k;
    }
    T.sync();
    const int nsel = (int)min((uint32_t)maxInter, H);
    for (int r = 0; r < nsel; ++r) {
        unsigned long long mhi = ~0ull, mlo = ~0ull;
        for (uint32_t i = T.lane; i < H; i += TW) {
            const HeapEnt k = M.heap[i];
            if (keyLess(k.hi, k.lo, mhi, mlo)) {
                mhi = k.hi;
                mlo = k.lo;
            }
        }
#pragma unroll
        for (int off = TW / 2; off > 0; off >>= 1) {
            const unsigned long long ohi = T.shflXor(mhi, off), olo = T.shflXor(mlo, off);
            if (keyLess(ohi, olo, mhi, mlo)) {
                mhi = ohi;
                mlo = olo;
            }
        }
        const uint32_t idx = (uint32_t)(mlo & 0xFFFFFFFFull);
        SMS_CHK(idx < H);
        if (T.lane == 0) {
            sel[r] = M.hist[idx];
            HeapEnt k;
            k.hi = ~0ull;
            k.lo = ~0ull;
            M.heap[idx] = k;
        }
        T.sync();
    }
    return nsel;
}

// ---- sequential helpers (lane 0) ------------------------------------------------------------------------------
// getShortestPath (:120-136): doubled ids t = p[0], previous_[p[0]] = p[1], ..., p[len] = source.  -1 on overflow.
__device__ inline int tracePath(const TileMem &M, uint32_t t, uint32_t src, int32_t *p) {
    int len = 0;
    uint32_t cur = t;
    p[0] = (int32_t)cur;
    while (cur != src) {
        if (len >= (int)M.caps.walkCap) return -1;
        const int slot = M.caps.dense ? (int)(cur >> 1) : findSlot(M, cur >> 1); // dense table: slot = vertex, nothing to probe
        SMS_CHK(slot >= 0);
        SMS_CHK(((cur & 1u) ? M.tab[slot].d1 : M.tab[slot].d0) < kUnlabelled);
        cur = predOf(M, slot, cur & 1u);
        p[++len] = (int32_t)cur;
    }
    return len;
}

// extractCycles (:229-328) on the closed walk c[0..len] (c[len] == c[0]) with stay flags st[0..len).
// Called by all 32 lanes of the warp with identical arguments (E and the serial parts belong to lane 0).  Everything that needs
// the graph is done with the lanes over INDEPENDENT items, so that a walk costs a handful of memory round trips instead of
// two or three per position (measured on the root-LP iterates of the 100x100 torus: 90 000 cycles per 18-edge walk, more
// than half of the round, profiles/README.md): (1) lanes over the positions: the arc of the edge entering the position, its
// weight and edge id, the row of the position's vertex; (2) lane 0: the reference's prefix sums (same order of additions);
// (3) lanes over ALL arcs of all positions, flattened: the chord test -- for every position i, every arc of c[i], the first
// earlier position of its end point -- with the candidates written in the reference's order (position, adjacency position,
// inner before outer) through a warp prefix sum.  Lane 0 then sorts and emits as before, from what (1) left in scratch.
__device__ __forceinline__ void extractCyclesBody(const GraphDev &G, TileMem &M, SrcState &S, Emitter &E, const int32_t *c,
                                                  const int32_t *st, const int len) {
    constexpr unsigned FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31;
    int32_t *warc = M.warc; // warc[i]: arc (in the row of c[i]) of the edge c[i-1] - c[i], i = 1..len
    int32_t *crs = M.crs;
    double *plen = M.plen;
    // scratch in the queue memory (idle once the intermediates are selected): per position the edge id of the entering
    // edge, the first arc of its vertex' row, and the exclusive prefix of the degrees of the positions 2 .. len-1
    if ((size_t)3 * (size_t)(len + 2) * sizeof(int32_t) > (size_t)M.caps.heapCap * sizeof(HeapEnt)) {
        S.overflow = 1;
        return;
    }
    int32_t *weid = reinterpret_cast<int32_t *>(M.heap);
    int32_t *rowB = weid + (len + 2);
    int32_t *degP = rowB + (len + 2);
    WCLK_DECL;
    // (1) lanes over the positions
    for (int i = 1 + lane; i <= len; i += 32) {
        const int32_t u = c[i], from = c[i - 1];
        const long long rb = G.rowptr[u], re = G.rowptr[u + 1];
        long long a = -1; // Graph::indexInOutEdgeArray (NetworKit Graph.cpp:119-127): first adjacency position of u holding `from`
        for (long long q = rb; q < re; ++q)
            if ((G.arcs[q].nbr & 0x7FFFFFFF) == from) {
                a = q;
                break;
            }
        SMS_CHK(a >= 0);
        SMS_CHK(i <= 2 * (int)M.caps.walkCap + 1);
        warc[i] = (int32_t)a;
        plen[i] = G.arcs[a].w; // the edge's weight for now, the prefix sum below
        weid[i] = G.eid[a];
        rowB[i] = (int32_t)rb; // 2m < 2^30 checked at create
        degP[i] = (i >= 2 && i < len) ? (int32_t)(re - rb) : 0;
    }
    __syncwarp(FULL);
    WCLK(16);
    // (2) prefix sums (:242-250), serially: the order of the additions is the reference's
    if (lane == 0) {
        int numCross = 0, arcsBefore = 0;
        double pathLength = 0.;
        crs[0] = 0;
        plen[0] = 0.;
        degP[0] = 0;
        for (int i = 1; i <= len; ++i) {
            numCross += st[i - 1] ? 0 : 1;
            const double w = plen[i];
            pathLength += st[i - 1] ? w : 1 - w;
            crs[i] = numCross;
            plen[i] = pathLength;
            const int dg = degP[i];
            degP[i] = arcsBefore;
            arcsBefore += dg;
        }
        degP[len + 1] = arcsBefore;
    }
    __syncwarp(FULL);
    WCLK(17);
    // (3) good chords (:254-306): lanes over the arcs of the positions 2 .. len-1, in (position, adjacency position) order
    int nch = 0;
    const double total = plen[len];
    const int T = degP[len + 1];
    for (int f0 = 0; f0 < T; f0 += 32) {
        const int f = f0 + lane;
        int f0c = -1, l0 = 0, f1c = -1, l1 = 0; // flags: bit0 crossing, bit1 inner
        int i = 0;
        long long a = 0;
        if (f < T) {
            int lo = 2, hi = len - 1; // the position of flat arc f: the largest i with degP[i] <= f
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (degP[mid] <= f)
                    lo = mid;
                else
                    hi = mid - 1;
            }
            i = lo;
            a = (long long)rowB[i] + (f - degP[i]);
            const Arc arc = G.arcs[a];
            const int32_t nb = arc.nbr & 0x7FFFFFFF;
            if (nb != c[i - 1] && nb != c[i + 1]) {
                int j = -1;
                for (int k = 0; k < i; ++k)
                    if (c[k] == nb) {
                        j = k;
                        break;
                    }
                if (j >= 0) {
                    const int innerCrossingDif = crs[i] - crs[j];
                    const double innerLen = plen[i] - plen[j];
                    const double outerLen = total - innerLen;
                    // weights_[edgeId(c[i], c[j])]: the first arc of c[i] towards c[j] -- this one, the graph is simple
                    const double x = arc.w;
                    if ((innerCrossingDif % 2) == 0) {
                        if (innerLen - x < 0) {
                            f0c = 1 | 2;
                            l0 = i - j + 1;
                        }
                        if (outerLen + x < 1) {
                            f1c = 0;
                            l1 = len - i + j + 1;
                        }
                    } else {
                        if (innerLen + x < 1) {
                            f0c = 2;
                            l0 = i - j + 1;
                        }
                        if (outerLen - x < 0) {
                            f1c = 1;
                            l1 = len - i + j + 1;
                        }
                    }
                    // the chord's start position j rides in the flags' high bits until the candidate is written
                    if (f0c >= 0) f0c |= j << 8;
                    if (f1c >= 0) f1c |= j << 8;
                }
            }
        }
        const int cnt = (f0c >= 0 ? 1 : 0) + (f1c >= 0 ? 1 : 0);
        int incl = cnt;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int o = __shfl_up_sync(FULL, incl, off);
            if (lane >= off) incl += o;
        }
        const int tot = __shfl_sync(FULL, incl, 31);
        if (tot == 0) continue;
        if (nch + tot > (int)M.caps.chordCap) {
            S.overflow = 1;
            return;
        }
        int pos = nch + incl - cnt;
        for (int q = 0; q < 2; ++q) {
            const int fl = q ? f1c : f0c;
            if (fl < 0) continue;
            int32_t *ch = M.chords + 4 * pos++;
            ch[0] = fl >> 8;
            ch[1] = i;
            ch[2] = q ? l1 : l0;
            ch[3] = (int32_t)((uint32_t)(fl & 3) | ((uint32_t)a << 2)); // 2m < 2^30 checked at create
        }
        nch += tot;
    }
    __syncwarp(FULL);
    WCLK(18);
    // stable sort by resulting length (:309; canonical: stable): a handful of candidates, lane 0
    if (lane == 0) {
        for (int a = 1; a < nch; ++a) {
            int32_t t0 = M.chords[4 * a], t1 = M.chords[4 * a + 1], t2 = M.chords[4 * a + 2], t3 = M.chords[4 * a + 3];
            int b = a - 1;
            while (b >= 0 && M.chords[4 * b + 2] > t2) {
                for (int q = 0; q < 4; ++q) M.chords[4 * (b + 1) + q] = M.chords[4 * b + q];
                --b;
            }
            M.chords[4 * (b + 1)] = t0;
            M.chords[4 * (b + 1) + 1] = t1;
            M.chords[4 * (b + 1) + 2] = t2;
            M.chords[4 * (b + 1) + 3] = t3;
        }
    }
    __syncwarp(FULL);
    WCLK(19);
    // greedy, non-overlapping (:312-323).  coveredNodes is READ by position and WRITTEN by node id (quirk A.7.3).  The decisions
    // are sequential over the candidates; the lanes share the test against the covered list and the writing of a new walk
    // (lane 0 owns the emitter and reserves its place).
    int ncov = 0, emitted = 0;
    const int covCap = 4 * (int)M.caps.walkCap + 8;
    for (int q = 0; q < nch; ++q) {
        const int32_t *ch = M.chords + 4 * q;
        const int start = ch[0], end = ch[1], wlen = ch[2];
        const uint32_t info = (uint32_t)ch[3];
        bool hit = false;
        for (int k = lane; k < min(ncov, covCap); k += 32) hit |= (M.covered[k] == start) || (M.covered[k] == end);
        if (__any_sync(FULL, hit)) continue;
        const bool crossing = info & 1u, inner = (info & 2u) != 0;
        const long long ca = (long long)(info >> 2);
        const int32_t chordEid = G.eid[ca];
        long long cut = 0, off = 0;
        int ok = 0;
        if (lane == 0) ok = emitReserve(E, wlen, cut, off) ? 1 : 0;
        ok = __shfl_sync(FULL, ok, 0);
        off = __shfl_sync(FULL, off, 0);
        // splitOnChord (odd_closed_walk.cpp:67-90): position p of the new walk
        for (int p = lane; p < wlen; p += 32) {
            int k = -1; // position in the old walk, -1: the chord
            if (inner) {
                if (p < end - start) k = start + p;
            } else {
                if (p < start)
                    k = p;
                else if (p > start)
                    k = end + (p - start - 1);
            }
            const int32_t node = k >= 0 ? c[k] : (inner ? c[end] : c[start]);
            const int32_t eid = k >= 0 ? weid[k + 1] : chordEid;
            const int32_t stay = k >= 0 ? st[k] : (crossing ? 0 : 1);
            if (ok) {
                E.R->node[off + p] = node;
                E.R->eid[off + p] = eid;
                E.R->stay[off + p] = (uint8_t)stay;
            }
            // covered: positions 1 .. size-2 of the new walk, by node id
            if (p >= 1 && p <= wlen - 2 && ncov + p - 1 < covCap) M.covered[ncov + p - 1] = node;
        }
        ncov += max(0, wlen - 2);
        if (ncov > covCap) {
            S.overflow = 1;
            return;
        }
        emitted++;
        __syncwarp(FULL); // the covered list is read by the next candidate
    }
    if (emitted == 0) { // :325-326: no chord cycle -> the cycle itself
        long long cut = 0, off = 0;
        int ok = 0;
        if (lane == 0) ok = emitReserve(E, len, cut, off) ? 1 : 0;
        ok = __shfl_sync(FULL, ok, 0);
        off = __shfl_sync(FULL, off, 0);
        if (ok)
            for (int k = lane; k < len; k += 32) {
                E.R->node[off + k] = c[k];
                E.R->eid[off + k] = weid[k + 1];
                E.R->stay[off + k] = (uint8_t)st[k];
            }
    }
    WCLK(20);
#ifdef SMSGPU_RUNS_CLOCKS
    if (threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(&g_runClk[21], 1ull);
#endif
}

// Two instantiations: inlined (kernels with registers to spare: the runs engine measured 58.5 against 53.1 rounds/s on the LP
// replay of the 100x100 torus) and out of line (the pairs kernel is compiled for 80 registers, and everything inlined into its
// finishing code moves the allocation of its hot loop: 503.8 against 472.2 ms on config 5).
template <bool OUTLINE>
struct ExtractCycles;
template <>
struct ExtractCycles<false> {
    static __device__ __forceinline__ void run(const GraphDev &G, TileMem &M, SrcState &S, Emitter &E, const int32_t *c, const int32_t *st,
                                               const int len) {
        extractCyclesBody(G, M, S, E, c, st, len);
    }
};
__device__ __noinline__ void extractCyclesOutOfLine(const GraphDev &G, TileMem &M, SrcState &S, Emitter &E, const int32_t *c, const int32_t *st,
                                                    const int len) {
    extractCyclesBody(G, M, S, E, c, st, len);
}
template <>
struct ExtractCycles<true> {
    static __device__ __forceinline__ void run(const GraphDev &G, TileMem &M, SrcState &S, Emitter &E, const int32_t *c, const int32_t *st,
                                               const int len) {
        extractCyclesOutOfLine(G, M, S, E, c, st, len);
    }
};

// simpleOddCycles (:202-227) on the closed walk wn[0..W], then extractCycles on every piece.  All 32 lanes, identical
// control flow (every lane reads the same walk); S.overflow is made uniform after every piece.
template <bool OUTLINE>
__device__ inline void simpleOddCyclesAndExtract(const GraphDev &G, TileMem &M, SrcState &S, Emitter &E, const int W) {
    constexpr unsigned FULL = 0xFFFFFFFFu;
    const int32_t *wn = M.wn, *ws = M.ws;
    int lastLeft = -1;
    bool any = false;
    const int lane = threadIdx.x & 31;
    const int32_t w32 = lane <= W ? wn[lane] : -1; // the first 32 positions in registers (most walks are shorter)
    for (int i = 0; i <= W; ++i) {
        const int32_t v = i < 32 ? __shfl_sync(FULL, w32, i) : wn[i];
        int pp = -1; // prevPos[v]: last earlier position of v -- the lanes share the scan
        for (int b = 0; b < i; b += 32) {
            const int k = b + lane;
            const int32_t wk = b == 0 ? w32 : (k < i ? wn[k] : -1);
            const unsigned m = __ballot_sync(FULL, k < i && wk == v);
            if (m) pp = b + 31 - __clz(m);
        }
        if (lastLeft < pp) {
            lastLeft = pp;
            if (i - lastLeft > 2) {
                ExtractCycles<OUTLINE>::run(G, M, S, E, wn + lastLeft, ws + lastLeft, i - lastLeft);
                __syncwarp(FULL);
                S.overflow = __shfl_sync(FULL, S.overflow, 0);
                if (S.overflow) return;
                any = true;
            }
        }
    }
    if (!any) {
        ExtractCycles<OUTLINE>::run(G, M, S, E, wn, ws, W);
        __syncwarp(FULL);
        S.overflow = __shfl_sync(FULL, S.overflow, 0);
    }
}

// intermediateVertexToOcw (:138-188) into M.wn / M.ws; returns the number of edges, -1 on overflow.
// The stem blacklisting side effect (:164-169) has no reader in independent-source mode and is performed in mode V only
// (G.blStampNext != nullptr): every vertex the reference blacklists gets the stamp `now` of this source.
__device__ inline int buildWalk(const GraphDev &G, const TileMem &M, uint32_t t, uint32_t src, int32_t now) {
    int32_t *pa = M.pa, *pb = M.pb;
    const int P = tracePath(M, t, src, pa);
    if (P < 0) return -1;
    const int Q = tracePath(M, t ^ 1u, src, pb);
    if (Q < 0) return -1;
    // path[i] = {u = orig(pa[i]), v = orig(pa[i+1]), stay = same side}, i < P; twinPath likewise from pb
    int W = 0;
    M.wn[0] = pa[0] >> 1;
    auto addEdge = [&](int32_t next, bool stay) {
        SMS_CHK(W < 2 * (int)M.caps.walkCap + 1);
        M.ws[W] = stay ? 1 : 0;
        M.wn[++W] = next;
    };
    if (Q == 0) {
        for (int i = 0; i < P; ++i) addEdge(pa[i + 1] >> 1, ((pa[i] ^ pa[i + 1]) & 1) == 0);
        return W;
    }
    int endPath = P - 1, endTwin = Q - 1;
    while (((pa[endPath + 1] >> 1) == (pb[endTwin + 1] >> 1)) && endPath > 0 && endTwin > 0) {
        endPath--;
        endTwin--;
        if (G.blStampNext) atomicMin(&G.blStampNext[pa[endPath + 1] >> 1], now); // blacklist_[path[endPath].v] = true (:168)
    }
    if ((pa[endPath + 1] >> 1) != (pb[endTwin + 1] >> 1)) {
        endPath++;
        endTwin++;
    }
    for (int i = 0; i <= endPath; ++i) addEdge(pa[i + 1] >> 1, ((pa[i] ^ pa[i + 1]) & 1) == 0);
    for (int i = endTwin; i >= 0; --i) addEdge(pb[i] >> 1, ((pb[i] ^ pb[i + 1]) & 1) == 0);
    return W;
}

// The walks of one source: intermediateVertexToOcw -> simpleOddCycles -> extractCycles for each selected intermediate
// (ocw_seperator.cpp:341-358).  All 32 lanes; lane 0 owns the emitter.  Leaves S.overflow uniform.
template <bool OUTLINE = false>
__device__ inline void walksOfSource(const GraphDev &G, TileMem &M, SrcState &S, Emitter &E, const uint32_t *sel, const int nsel,
                                     const uint32_t src, const int32_t now) {
    constexpr unsigned FULL = 0xFFFFFFFFu;
    WCLK_DECL;
    for (int r = 0; r < nsel; ++r) {
        int W = 0;
        if ((threadIdx.x & 31) == 0) W = buildWalk(G, M, sel[r], src, now);
        __syncwarp(FULL);
        W = __shfl_sync(FULL, W, 0);
        WCLK(11);
#ifdef SMSGPU_RUNS_CLOCKS
        if (threadIdx.x == 0 && blockIdx.x == 0 && W > 0) {
            atomicAdd(&g_runClk[13], (unsigned long long)W);
            atomicAdd(&g_runClk[14], 1ull);
            atomicMax(&g_runClk[15], (unsigned long long)W);
        }
#endif
        if (W < 0) {
            S.overflow = 1;
            return;
        }
        simpleOddCyclesAndExtract<OUTLINE>(G, M, S, E, W);
        WCLK(12);
        if (S.overflow) return;
    }
}

// Mode V, end of a source that completed in this tier: remember whether it found nothing, and if so do what
// `if (intermediateVertices_.empty()) blacklistIntReachable(originalId(source))` (ocw_seperator.cpp:113-114, 363-382)
// does at the source vertex itself: blacklist it unconditionally and blacklist its not-yet-blacklisted neighbours over
// integral edges (weight exactly 0.0 or 1.0).  The rest of that depth-first search -- the flood of this source's stamp
// through integral edges -- is floodStampsKernel's job (ocw_kernels.cuh), for all sources of the sweep at once.
template <int TW>
__device__ inline void verbatimSourceDone(const Tile<TW> &T, const GraphDev &G, const uint32_t src, const int32_t now, const bool empty) {
    if (T.lane == 0) G.srcEmpty[now] = empty ? 1 : 0;
    if (!empty) return;
    const uint32_t u = src >> 1;
    if (T.lane == 0) atomicMin(&G.blStampNext[u], now);
    for (long long a = G.rowptr[u] + T.lane; a < G.rowptr[u + 1]; a += TW) {
        const Arc arc = G.arcs[a];
        const uint32_t v = (uint32_t)arc.nbr & 0x7FFFFFFFu;
        if ((arc.w == 0.0 || arc.w == 1.0) && arc.nbr >= 0 && !stampedBefore(G, v, now)) atomicMin(&G.blStampNext[v], now);
    }
}

} // namespace smsgpu
