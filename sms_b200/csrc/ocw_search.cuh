// K2-K4 fused: bounded shortest paths on the implicit double cover + walk extraction + chord repair,
// one source per TILE of TW lanes (TW = 8, 16 or 32; 32/TW sources per warp run in lock step).
//
// Semantics = OcwSeparator::getViolatedOcws (reference src/graph/ocw_seperator.cpp:330-361) for one source
// against the round's INITIAL blacklist, with canonical tie-breaking (queue order (dist, hops, doubled id),
// stable sorts).  Restated on the CPU, function by function, in oracle/ocw_oracle.hpp (canonical = true);
// the parity tests compare the two walk for walk.
//
// What differs from the reference is only HOW the same sequence of events is produced:
//   * the priority queue is a TW-ary implicit heap whose sift-down examines one node's TW children with one
//     coalesced load + a shuffle reduction; entries are (dist, hops, id) keys, updated lazily (a stale entry
//     always pops after the node was settled through its fresher, smaller entry and is discarded);
//   * only nodes that can still be extracted (2*dist <= bestDist) enter the queue;
//   * distances live in a per-source open-addressing table keyed by original vertex (both sides in one
//     32-byte entry) instead of three dense arrays of 2n+1 -- cost proportional to the ball, not to n;
//   * a label that provably can never matter is not created ("label pruning").  A relaxation from
//     the node being settled (distance dc) offering d to a side that has no label is dropped iff
//     d + dc > bestDist AND d + dist[twin] > bestDist: it cannot set a record now, nor later (any later or
//     improved twin label is >= dc), it can never be extracted (2d > bestDist), and every later offer to the
//     same side that the reference would reject against it is dropped by the same test.  Labels that ARE
//     stored always equal the reference's distances_/hopDistances_;
//   * the arcs of the settled node are relaxed by the lanes in parallel (one arc per lane: stay, then cross);
//     the order-dependent part -- the running bestDist / intermediate history of ocw_seperator.cpp:71-82 -- is
//     replayed serially in adjacency order over the few lanes whose D = dist[t] + dist[t^1] can matter.
#pragma once
#include "ocw_types.cuh"

namespace smsgpu {

#ifdef SMSGPU_DEBUG_CHECKS
// value must be identical in all lanes of the tile
#define SMS_UNI(T, x)                                   \
    do {                                                \
        const auto uni_ref_ = (T).shfl((x), 0);         \
        SMS_CHK(uni_ref_ == (x));                       \
    } while (0)
#else
#define SMS_UNI(T, x) ((void)0)
#endif

template <int TW>
struct Tile {
    unsigned mask; // lanes of this tile inside the warp
    int lane;      // 0..TW-1
    int shift;     // first warp lane of the tile
    __device__ Tile() {
        const int wl = threadIdx.x & 31;
        lane = wl % TW;
        shift = wl - lane;
        mask = (TW == 32) ? 0xFFFFFFFFu : (((1u << TW) - 1u) << shift);
    }
    __device__ __forceinline__ void sync() const { __syncwarp(mask); }
    __device__ __forceinline__ unsigned ballot(bool p) const {
        return (__ballot_sync(mask, p) >> shift) & ((TW == 32) ? 0xFFFFFFFFu : ((1u << TW) - 1u));
    }
    template <class T>
    __device__ __forceinline__ T shfl(T v, int src) const {
        return __shfl_sync(mask, v, src, TW);
    }
    template <class T>
    __device__ __forceinline__ T shflXor(T v, int m) const {
        return __shfl_xor_sync(mask, v, m, TW);
    }
};

// One tile's working memory (generic pointers: shared memory in tier 0, global memory otherwise).
struct TileMem {
    Entry *tab;
    uint32_t *pred;    // [slot*2 + side] predecessor doubled id
    uint32_t *touched; // slots created by the current source
    HeapEnt *heap;
    uint32_t *hist;
    int32_t *pa, *pb;          // predecessor chains (doubled ids)
    int32_t *wn, *ws, *warc;   // walk nodes / stay flags / arc index of the edge ENTERING position i
    int32_t *crs;              // prefix #cross
    double *plen;              // prefix length
    int32_t *chords;           // 4 ints per chord: start, end, length, flags(bit0 crossing, bit1 inner) ; arc in high bits? no: see below
    int32_t *covered;
    TileCaps caps;
    uint32_t tabMask;
    // fast engine only (null / 0 in the heap engine)
    uint2 *pq;       // [slot*2 + side] {predecessor doubled id, queue position or kNoQpos}; aliases pred
    uint32_t *cl;    // bucket counters, two buckets per word: byte 0/2 = appended, byte 1/3 = live
    uint32_t *bm;    // presence bitmap of the label table, one bit per slot
    uint32_t hopMask; // hop words: all 32 bits, except in the pairs engine (hops | queue position << 16)
    uint4 *qrow;     // pairs engine: its queue, two uint4 per entry: {dist bits, hops << 32 | id}, {first arc, degree, -, -}
    // lanes engine (ocw_lanes_core.cuh; 0 / null elsewhere): a slot is occupied iff its flags word equals `stamp`; the
    // predecessor of a label sits in the queue entry its hop word points to
    uint32_t stamp;
    const uint4 *lq;
    uint32_t lqShift, lqMask; // queue entry of a label: (hop word >> lqShift) & lqMask
};

__device__ inline void carveTileMem(TileMem &M, unsigned char *base, const TileCaps &c, unsigned char *hot = nullptr) {
    M.caps = c;
    M.pq = nullptr;
    M.cl = nullptr;
    M.bm = nullptr;
    M.qrow = nullptr;
    M.stamp = 0;
    M.lq = nullptr;
    M.lqShift = 0;
    M.lqMask = 0;
    M.hopMask = c.pairs ? 0xFFFFu : 0xFFFFFFFFu;
    if (c.nbuckets) {
        M.cl = reinterpret_cast<uint32_t *>(hot);
        M.bm = reinterpret_cast<uint32_t *>(hot + (size_t)c.nbuckets * (c.pairs ? 1 : 2));
    }
    M.tabMask = c.tabCap - 1;
    unsigned char *p = base;
    M.tab = reinterpret_cast<Entry *>(p);
    p += (size_t)c.tabCap * sizeof(Entry);
    M.heap = reinterpret_cast<HeapEnt *>(p);
    p += (size_t)c.heapCap * sizeof(HeapEnt);
    M.plen = reinterpret_cast<double *>(p);
    p += (size_t)(2 * c.walkCap + 2) * sizeof(double);
    M.pred = reinterpret_cast<uint32_t *>(p);
    if (c.nbuckets) M.pq = reinterpret_cast<uint2 *>(p); // plen above keeps this 8-byte aligned
    p += (size_t)c.tabCap * 2 * sizeof(uint32_t) * (c.nbuckets ? 2 : 1);
    M.touched = reinterpret_cast<uint32_t *>(p);
    p += (size_t)c.touchedCap * sizeof(uint32_t);
    M.hist = reinterpret_cast<uint32_t *>(p);
    p += (size_t)c.histCap * sizeof(uint32_t);
    M.pa = reinterpret_cast<int32_t *>(p);
    p += (size_t)(c.walkCap + 2) * sizeof(int32_t);
    M.pb = reinterpret_cast<int32_t *>(p);
    p += (size_t)(c.walkCap + 2) * sizeof(int32_t);
    M.wn = reinterpret_cast<int32_t *>(p);
    p += (size_t)(2 * c.walkCap + 2) * sizeof(int32_t);
    M.ws = reinterpret_cast<int32_t *>(p);
    p += (size_t)(2 * c.walkCap + 2) * sizeof(int32_t);
    M.warc = reinterpret_cast<int32_t *>(p);
    p += (size_t)(2 * c.walkCap + 2) * sizeof(int32_t);
    M.crs = reinterpret_cast<int32_t *>(p);
    p += (size_t)(2 * c.walkCap + 2) * sizeof(int32_t);
    M.chords = reinterpret_cast<int32_t *>(p);
    p += (size_t)c.chordCap * 4 * sizeof(int32_t);
    M.covered = reinterpret_cast<int32_t *>(p);
    p += (size_t)(4 * c.walkCap + 8) * sizeof(int32_t);
    if (c.pairs) M.qrow = reinterpret_cast<uint4 *>(base + (((size_t)(p - base) + 31) & ~(size_t)31));
}

__device__ __forceinline__ Entry emptyEntry() {
    Entry e;
    e.d0 = kUnlabelled;
    e.d1 = kUnlabelled;
    e.key = kEmptyKey;
    e.h0 = kUnlabelledHops;
    e.h1 = kUnlabelledHops;
    e.flags = 0;
    return e;
}
__device__ __forceinline__ void storeEntry(Entry *p, const Entry &e) {
    uint4 *q = reinterpret_cast<uint4 *>(p);
    const uint4 *s = reinterpret_cast<const uint4 *>(&e);
    q[0] = s[0];
    q[1] = s[1];
}
__device__ __forceinline__ Entry loadEntry(const Entry *p) {
    Entry e;
    const uint4 *q = reinterpret_cast<const uint4 *>(p);
    uint4 *s = reinterpret_cast<uint4 *>(&e);
    s[0] = q[0];
    s[1] = q[1];
    return e;
}

__device__ __forceinline__ uint32_t homeSlot(const TileMem &M, uint32_t v) {
    return M.caps.dense ? v : ((v * 0x9E3779B1u) >> M.caps.tabShift);
}
// slot of original vertex v, or -1
__device__ __forceinline__ int findSlot(const TileMem &M, uint32_t v) {
    uint32_t s = homeSlot(M, v);
    if (M.stamp) { // lanes engine: occupied iff the entry carries the stamp of the search
        while (true) {
            const uint4 hb = reinterpret_cast<const uint4 *>(&M.tab[s])[1]; // {key, h0, h1, flags}
            if (hb.w != M.stamp) return -1;
            if (hb.x == v) return (int)s;
            s = (s + 1) & M.tabMask;
        }
    }
    if (M.bm) { // fast engine: a slot is occupied iff its bit is set; entries of free slots hold garbage
        while (true) {
            if (!((M.bm[s >> 5] >> (s & 31)) & 1u)) return -1;
            if (M.tab[s].key == v) return (int)s;
            s = (s + 1) & M.tabMask;
        }
    }
    SMS_CHK(s < M.caps.tabCap);
#ifdef SMSGPU_DEBUG_CHECKS
    uint32_t probes = 0;
#endif
    while (true) {
        const uint32_t k = M.tab[s].key;
        if (k == v) return (int)s;
        if (k == kEmptyKey) return -1;
        s = (s + 1) & M.tabMask;
        SMS_CHK(++probes <= M.caps.tabCap);
    }
}
// find or create; *isNew tells whether this call created it.  Safe against the other lanes of the tile
// inserting different keys at the same time.
__device__ __forceinline__ int findOrInsertSlot(const TileMem &M, uint32_t v, bool *isNew) {
    uint32_t s = homeSlot(M, v);
    *isNew = false;
    SMS_CHK(s < M.caps.tabCap);
#ifdef SMSGPU_DEBUG_CHECKS
    uint32_t probes = 0;
#endif
    while (true) {
        SMS_CHK(++probes <= M.caps.tabCap);
        uint32_t k = M.tab[s].key;
        if (k == v) return (int)s;
        if (k == kEmptyKey) {
            k = atomicCAS(&M.tab[s].key, kEmptyKey, v);
            if (k == kEmptyKey) {
                *isNew = true;
                return (int)s;
            }
            if (k == v) return (int)s;
        }
        s = (s + 1) & M.tabMask;
    }
}

__device__ __forceinline__ bool keyLess(unsigned long long ahi, unsigned long long alo, unsigned long long bhi,
                                        unsigned long long blo) {
    return ahi < bhi || (ahi == bhi && alo < blo);
}

__device__ __forceinline__ uint32_t predOf(const TileMem &M, int slot, uint32_t side) {
    if (M.lq) { // lanes engine: the hop word names the label's queue entry, which holds the predecessor
        const uint32_t hw = side ? M.tab[slot].h1 : M.tab[slot].h0;
        return M.lq[2 * (size_t)((hw >> M.lqShift) & M.lqMask) + 1].z;
    }
#ifndef SMSGPU_PAIRS_PRED_ARRAY
    if (M.qrow) { // pairs engine: the predecessor sits in the label's queue entry, in the array only when it has none
        const uint32_t q = (side ? M.tab[slot].h1 : M.tab[slot].h0) >> 16;
        if (q != 0xFFFFu) return M.qrow[2 * (size_t)q + 1].z;
    }
#endif
    return M.pq ? M.pq[2 * slot + side].x : M.pred[2 * slot + side];
}

// ---- TW-ary implicit heap, tile-cooperative ----------------------------------------------------------------
template <int TW>
__device__ __forceinline__ void heapPush(const Tile<TW> &T, HeapEnt *heap, uint32_t &heapN, unsigned long long hi,
                                         unsigned long long lo) {
    uint32_t i = heapN++;
    SMS_UNI(T, i);
    SMS_UNI(T, hi);
    SMS_UNI(T, lo);
    while (i > 0) {
        const uint32_t p = (i - 1) / TW;
        const HeapEnt pe = heap[p]; // same address in all lanes: broadcast
        if (!keyLess(hi, lo, pe.hi, pe.lo)) break;
        T.sync();
        if (T.lane == 0) heap[i] = pe;
        i = p;
    }
    T.sync();
    if (T.lane == 0) {
        HeapEnt e;
        e.hi = hi;
        e.lo = lo;
        heap[i] = e;
    }
    T.sync();
}

template <int TW>
__device__ __forceinline__ void heapPop(const Tile<TW> &T, HeapEnt *heap, uint32_t &heapN) {
    heapN--;
    if (heapN == 0) return;
    const HeapEnt last = heap[heapN];
    uint32_t i = 0;
    while (true) {
        const uint32_t c0 = i * TW + 1;
        if (c0 >= heapN) break;
        const uint32_t ci = c0 + T.lane;
        unsigned long long khi = ~0ull, klo = ~0ull;
        if (ci < heapN) {
            const HeapEnt ce = heap[ci];
            khi = ce.hi;
            klo = ce.lo;
        }
        unsigned long long mhi = khi, mlo = klo;
#pragma unroll
        for (int off = TW / 2; off > 0; off >>= 1) {
            const unsigned long long ohi = T.shflXor(mhi, off), olo = T.shflXor(mlo, off);
            if (keyLess(ohi, olo, mhi, mlo)) {
                mhi = ohi;
                mlo = olo;
            }
        }
        if (!keyLess(mhi, mlo, last.hi, last.lo)) break;
        const unsigned who = T.ballot(khi == mhi && klo == mlo);
        const int ml = __ffs(who) - 1;
        T.sync();
        if (T.lane == 0) {
            HeapEnt e;
            e.hi = mhi;
            e.lo = mlo;
            heap[i] = e;
        }
        i = c0 + ml;
    }
    T.sync();
    if (T.lane == 0) heap[i] = last;
    T.sync();
}

// ---- per-source state that is uniform across the tile -----------------------------------------------------
struct SrcState {
    double best;       // bestDist
    double bestHop;    // bestHopDist (a double in the reference, ocw_seperator.cpp:52)
    uint32_t histN, heapN, touchedN;
    long long arcs, settled;
    long long steps, loads; // extraction steps / bucket loads (fast engines; settled / steps = keys per step)
    int overflow;
};

// runDijsktra (ocw_seperator.cpp:37-112), canonical order.  Returns with S.overflow != 0 if a capacity was hit.
template <int TW>
__device__ void searchSource(const Tile<TW> &T, const GraphDev &G, TileMem &M, SrcState &S, const uint32_t src,
                             const double bestInit, const int32_t now) {
    S.best = bestInit;
    S.bestHop = 0.0;
    S.histN = 0;
    S.heapN = 0;
    S.touchedN = 0;
    S.arcs = 0;
    S.settled = 0;
    S.overflow = 0;
    const uint32_t target = src ^ 1u;

    // distances_[source] = 0, hopDistances_[source] = 0, heap.push(source)   (:54-57)
    if (T.lane == 0) {
        bool isNew;
        const int slot = findOrInsertSlot(M, src >> 1, &isNew);
        Entry e = emptyEntry();
        e.key = src >> 1;
        if (src & 1u) {
            e.d1 = 0.0;
            e.h1 = 0;
        } else {
            e.d0 = 0.0;
            e.h0 = 0;
        }
        storeEntry(&M.tab[slot], e);
        M.touched[0] = (uint32_t)slot;
    }
    S.touchedN = 1;
    T.sync();
    heapPush<TW>(T, M.heap, S.heapN, 0ull, (unsigned long long)src);

    while (true) {
        // ---- top of the queue, skipping stale entries ------------------------------------------------------
        uint32_t cur = 0, hc = 0;
        double dc = 0.0;
        int curSlot = -1;
        bool have = false;
        while (S.heapN > 0) {
            const HeapEnt top = M.heap[0];
            cur = (uint32_t)(top.lo & 0xFFFFFFFFull);
            hc = (uint32_t)(top.lo >> 32);
            dc = __longlong_as_double((long long)top.hi);
            curSlot = findSlot(M, cur >> 1);
            SMS_CHK(curSlot >= 0);
            SMS_UNI(T, S.heapN);
            SMS_UNI(T, cur);
            const uint32_t fl = M.tab[curSlot].flags;
            if (fl & (1u << (cur & 1u))) { // already settled through a fresher entry
                heapPop<TW>(T, M.heap, S.heapN);
                continue;
            }
            have = true;
            break;
        }
        // loop condition of ocw_seperator.cpp:88-90
        SMS_UNI(T, (int)have);
        SMS_UNI(T, S.best);
        SMS_UNI(T, dc);
        if (!have || cur == target || 2 * dc > S.best) break;
        heapPop<TW>(T, M.heap, S.heapN);
        // every lane has read this entry's flags in the staleness test above; heapPop returns without a barrier
        // when the queue becomes empty, so order those reads before lane 0 sets the settled bit
        T.sync();
        if (T.lane == 0) M.tab[curSlot].flags |= (1u << (cur & 1u));
        S.settled++;
        const uint32_t side = cur & 1u;
        const uint32_t u = cur >> 1;
        SMS_CHK((long long)u < G.n);
        const long long rb = G.rowptr[u], re = G.rowptr[u + 1];
        SMS_CHK(rb >= 0 && re >= rb && re <= 2 * G.m);
        const uint32_t hNew = hc + 1;

        for (long long base = rb; base < re; base += TW) {
            if (S.touchedN + TW > M.caps.touchedCap || S.heapN + 2 * TW > M.caps.heapCap) {
                S.overflow = 1;
                return;
            }
            const double best0 = S.best; // conservative (>= the running value) for pruning / queue admission
            const long long a = base + T.lane;
            bool act = a < re;
            uint32_t nbr = 0;
            double w = 0.0;
            if (act) {
                const Arc arc = G.arcs[a];
                nbr = (uint32_t)arc.nbr & 0x7FFFFFFFu;
                w = arc.w;
                SMS_CHK((long long)nbr < G.n && w >= 0.0 && w <= 1.0);
                act = arc.nbr >= 0; // blacklisted neighbours are skipped (:98-99)
                if (G.blStamp && act) act = !stampedBefore(G, nbr, now); // mode V: blacklisted by an earlier source
            }
            S.arcs += __popc(T.ballot(act));
            const double dS = dc + w;         // totalStayDist  (:104)
            const double dX = dc + (1.0 - w); // crossWeight = 1.0 - stayWeight; totalCrossDist (:107-108)
            // May a label be created for a side that has none?  Only if it could still matter (header comment).
            const bool creatS = dS + dc <= best0;
            const bool creatX = dX + dc <= best0;

            int slot = -1;
            bool isNew = false;
            if (act) {
                if (creatS || creatX)
                    slot = findOrInsertSlot(M, nbr, &isNew);
                else
                    slot = findSlot(M, nbr);
            }
            // touched list for the slots created in this chunk
            {
                const unsigned nm = T.ballot(isNew);
                SMS_CHK(S.touchedN + __popc(nm) <= M.caps.touchedCap);
                if (isNew) M.touched[S.touchedN + __popc(nm & ((1u << T.lane) - 1u))] = (uint32_t)slot;
                S.touchedN += __popc(nm);
                SMS_UNI(T, S.touchedN);
            }
            bool accS = false, accX = false, evS = false, evX = false;
            double DS = 0, DX = 0;
            uint32_t HS = 0, HX = 0;
            if (slot >= 0) {
                SMS_CHK((uint32_t)slot < M.caps.tabCap);
                Entry e = loadEntry(&M.tab[slot]);
                SMS_CHK(e.key == nbr);
                // label of the side the stay arc reaches (same side as cur) and of the other side
                double dA = side ? e.d1 : e.d0, dB = side ? e.d0 : e.d1;
                uint32_t hA = side ? e.h1 : e.h0, hB = side ? e.h0 : e.h1;
                // updatePaths(stayNeighbor, ...)   (:59-84)
                if (dA >= kUnlabelled)
                    accS = dS < kInfDist && (creatS || dS + dB <= best0);
                else
                    accS = (dA == dS) ? (hNew < hA) : (dS < dA);
                if (accS) {
                    dA = dS;
                    hA = hNew;
                    DS = dA + dB;
                    HS = hA + hB;
                    evS = DS <= best0;
                }
                // updatePaths(crossNeighbor, ...)
                if (dB >= kUnlabelled)
                    accX = dX < kInfDist && (creatX || dX + dA <= best0);
                else
                    accX = (dB == dX) ? (hNew < hB) : (dX < dB);
                if (accX) {
                    dB = dX;
                    hB = hNew;
                    DX = dB + dA;
                    HX = hB + hA;
                    evX = DX <= best0;
                }
                if (accS || accX) {
                    if (side) {
                        e.d1 = dA;
                        e.h1 = hA;
                        e.d0 = dB;
                        e.h0 = hB;
                    } else {
                        e.d0 = dA;
                        e.h0 = hA;
                        e.d1 = dB;
                        e.h1 = hB;
                    }
                    e.key = nbr;
                    storeEntry(&M.tab[slot], e);
                    if (accS) M.pred[2 * slot + side] = cur;
                    if (accX) M.pred[2 * slot + (side ^ 1u)] = cur;
                }
            }
            T.sync();
            // ---- replay of the record logic in adjacency order (:71-82) -------------------------------------
            unsigned evm = T.ballot(evS || evX);
            SMS_UNI(T, evm);
            while (evm) {
                const int l = __ffs(evm) - 1;
                evm &= evm - 1;
                const uint32_t nb = T.shfl(nbr, l);
                const int eS = T.shfl((int)evS, l), eX = T.shfl((int)evX, l);
                const double D0 = T.shfl(DS, l), D1 = T.shfl(DX, l);
                const uint32_t H0 = T.shfl(HS, l), H1 = T.shfl(HX, l);
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    if (!(k ? eX : eS)) continue;
                    const double D = k ? D1 : D0;
                    const double H = (double)(k ? H1 : H0);
                    bool push = false;
                    if (D < S.best || (D == S.best && H < S.bestHop)) {
                        S.best = D;
                        S.bestHop = H;
                        push = true;
                    } else if (D == S.best && H < 2 * S.bestHop) {
                        push = true;
                    }
                    if (push) {
                        if (S.histN >= M.caps.histCap) {
                            S.overflow = 1;
                            return;
                        }
                        if (T.lane == 0) M.hist[S.histN] = 2 * nb + (k ? (side ^ 1u) : side);
                        S.histN++;
                    }
                }
            }
            // ---- queue admission: only nodes that can still be extracted ------------------------------------
            const bool pushS = accS && 2 * dS <= best0, pushX = accX && 2 * dX <= best0;
            SMS_UNI(T, S.best);
            SMS_UNI(T, S.histN);
            unsigned pm = T.ballot(pushS || pushX);
            SMS_UNI(T, pm);
            SMS_UNI(T, S.heapN);
            while (pm) {
                const int l = __ffs(pm) - 1;
                pm &= pm - 1;
                const uint32_t nb = T.shfl(nbr, l);
                const int pS = T.shfl((int)pushS, l), pX = T.shfl((int)pushX, l);
                const double d0 = T.shfl(dS, l), d1 = T.shfl(dX, l);
                SMS_UNI(T, l);
                SMS_UNI(T, nb);
                SMS_UNI(T, pS);
                SMS_UNI(T, pX);
                SMS_UNI(T, d0);
                SMS_UNI(T, d1);
                SMS_UNI(T, hNew);
                SMS_UNI(T, side);
                SMS_CHK(S.heapN + 2 <= M.caps.heapCap);
                if (pS)
                    heapPush<TW>(T, M.heap, S.heapN, (unsigned long long)__double_as_longlong(d0),
                                 ((unsigned long long)hNew << 32) | (2 * nb + side));
                if (pX)
                    heapPush<TW>(T, M.heap, S.heapN, (unsigned long long)__double_as_longlong(d1),
                                 ((unsigned long long)hNew << 32) | (2 * nb + (side ^ 1u)));
            }
        }
    }
}

} // namespace smsgpu
