// "Fast" engine of K2 (tiers 1 and 2): the same search as searchSource (ocw_search.cuh) -- runDijsktra of the
// reference (src/graph/ocw_seperator.cpp:37-112) in the canonical order (dist, hops, doubled id), event for event --
// with the three things the ncu profile of the heap engine showed to dominate (profiles/r01_search_tier1_*.md)
// replaced:
//   * the 32-ary heap in HBM (sift loops = 22 % of all instructions, 3 dependent loads per pop, one serial push
//     per accepted label) becomes a MONOTONE BUCKET QUEUE: nbuckets buckets of equal width over [0, bestInit/2],
//     each a fixed array of bucketCap <= 32 keys.  Extraction order only needs the minimum of the first non-empty
//     bucket: one coalesced load + four REDUX steps.  Pushes are independent stores (position from one shared-memory
//     atomic), done by all lanes at once.  Keys are never moved: an extracted or superseded key is overwritten by a
//     tombstone, so a label can remember where its key sits and no stale key is ever extracted (the heap engine's
//     "settled" flag test, one dependent table probe per extraction, disappears);
//   * the label table stays in HBM, but which slots are occupied is a BITMAP in shared memory: a lookup of a vertex
//     that has no label (most arcs) costs no memory access, a new label is written without reading the slot first,
//     and nothing has to be cleaned up after a source except the bitmap.  A slot's entry carries the stamp of the
//     source that wrote it, so the garbage left in a slot by an earlier source can never be mistaken for a label;
//   * the per-chunk work is one pass: probe, relax both sides, tombstone a superseded key, push, store.
// Distances, hops, predecessors, the record logic and therefore the walks are bit-identical to the heap engine's
// (tests/test_gpu_parity.py runs every tier against the oracle).
#pragma once
#include "ocw_search.cuh"

namespace smsgpu {

constexpr uint32_t kNoQpos = 0xFFFFFFFFu;
constexpr unsigned long long kDeadKey = ~0ull;

// bucket counters: two buckets per 32-bit word; per bucket one byte "appended" and one byte "live"
__device__ __forceinline__ uint32_t clShift(uint32_t b) { return (b & 1u) * 16u; }

// returns the bucket the key sat in
__device__ __forceinline__ uint32_t killKey(const TileMem &M, uint32_t qpos) {
    M.heap[qpos].hi = kDeadKey;
    const uint32_t b = qpos / M.caps.bucketCap;
    atomicSub(&M.cl[b >> 1], 0x0100u << clShift(b));
    return b;
}

template <int TW>
__device__ void searchSourceFast(const Tile<TW> &T, const GraphDev &G, TileMem &M, SrcState &S, const uint32_t src,
                                 const double bestInit, const uint32_t stamp, const int32_t now) {
    static_assert(TW == 32, "the fast engine uses full-warp REDUX");
    S.best = bestInit;
    S.bestHop = 0.0;
    S.histN = 0;
    S.heapN = 0;
    S.touchedN = 0;
    S.arcs = 0;
    S.settled = 0;
    S.steps = 0;
    S.loads = 0;
    S.overflow = 0;
    const uint32_t target = src ^ 1u;
    const uint32_t NB = M.caps.nbuckets, C = M.caps.bucketCap;
    const uint32_t nclWords = NB >> 1, nbmWords = M.caps.tabCap >> 5;
    const double bscale = (double)NB / (0.5 * bestInit);

    // ---- reset the hot words, then label and enqueue the source (:54-57) ------------------------------------------
    for (uint32_t i = T.lane; i < nclWords; i += TW) M.cl[i] = 0;
    for (uint32_t i = T.lane; i < nbmWords; i += TW) M.bm[i] = 0;
    T.sync();
    if (T.lane == 0) {
        const uint32_t s = homeSlot(M, src >> 1);
        M.bm[s >> 5] = 1u << (s & 31);
        Entry e = emptyEntry();
        e.key = src >> 1;
        e.flags = stamp;
        if (src & 1u) {
            e.d1 = 0.0;
            e.h1 = 0;
        } else {
            e.d0 = 0.0;
            e.h0 = 0;
        }
        storeEntry(&M.tab[s], e);
        M.pq[2 * s + (src & 1u)] = make_uint2(src, 0u);
        HeapEnt k;
        k.hi = 0ull;
        k.lo = (unsigned long long)src;
        M.heap[0] = k;
        M.cl[0] = 0x0101u; // bucket 0: one appended, one live
    }
    S.touchedN = 1;
    T.sync();

    // The keys of the current bucket stay in registers (lane i holds position i) together with the row range of
    // their vertices, so an extraction normally costs four REDUX steps and no memory access; the bucket is
    // (re)loaded only when it is exhausted or a step appended to / killed a key in it (rare: that needs an arc
    // lighter than one bucket width).  Loading a bucket also prefetches the arc rows of all its keys into L2.
    uint32_t cb = 0;       // current bucket; extraction keys never decrease
    bool needLoad = true;  // registers do not mirror bucket cb
    uint32_t liveRegs = 0; // live keys among the registers
    unsigned long long khi = kDeadKey, klo = kDeadKey;
    long long rowB = 0, rowE = 0;
    while (true) {
        if (needLoad) {
            // ---- first non-empty bucket at or after cb ------------------------------------------------------------
            uint32_t b = NB, cnt = 0;
            for (uint32_t base = cb >> 1; base < nclWords; base += TW) {
                const uint32_t wi = base + T.lane;
                const uint32_t word = wi < nclWords ? M.cl[wi] : 0u;
                uint32_t live = word & 0xFF00FF00u;
                if (wi == (cb >> 1) && (cb & 1u)) live &= 0xFF000000u;
                const unsigned bal = T.ballot(live != 0u);
                if (bal) {
                    const int l = __ffs(bal) - 1;
                    const uint32_t lw = T.shfl(live, l), ww = T.shfl(word, l);
                    const uint32_t odd = (lw & 0x0000FF00u) ? 0u : 1u;
                    b = 2 * (base + l) + odd;
                    cnt = (ww >> (odd * 16u)) & 0xFFu;
                    break;
                }
            }
            if (b >= NB) break; // queue empty
            cb = b;
            khi = kDeadKey;
            klo = kDeadKey;
            if ((uint32_t)T.lane < cnt) {
                const HeapEnt k = M.heap[b * C + T.lane];
                khi = k.hi;
                klo = k.lo;
            }
            if (khi != kDeadKey) {
                const uint32_t v = (uint32_t)klo >> 1;
                rowB = G.rowptr[v];
                rowE = G.rowptr[v + 1];
                asm volatile("prefetch.global.L2 [%0];" ::"l"(G.arcs + rowB));
            }
            liveRegs = __popc(T.ballot(khi != kDeadKey));
            SMS_CHK(liveRegs > 0);
            needLoad = false;
        }
        // ---- minimum key of the bucket: (dist, hops, id) lexicographic, tombstones = all ones -------------------------
        const uint32_t k0 = (uint32_t)(khi >> 32), k1 = (uint32_t)khi, k2 = (uint32_t)(klo >> 32), k3 = (uint32_t)klo;
        const uint32_t m0 = __reduce_min_sync(T.mask, k0);
        bool mine = k0 == m0;
        const uint32_t m1 = __reduce_min_sync(T.mask, mine ? k1 : 0xFFFFFFFFu);
        mine = mine && k1 == m1;
        const uint32_t m2 = __reduce_min_sync(T.mask, mine ? k2 : 0xFFFFFFFFu);
        mine = mine && k2 == m2;
        const uint32_t m3 = __reduce_min_sync(T.mask, mine ? k3 : 0xFFFFFFFFu);
        mine = mine && k3 == m3 && khi != kDeadKey;
        SMS_CHK(m0 != 0xFFFFFFFFu);
        const uint32_t cur = m3, hc = m2;
        const double dc = __longlong_as_double((long long)(((unsigned long long)m0 << 32) | m1));
        // loop condition of ocw_seperator.cpp:88-90
        if (cur == target || 2 * dc > S.best) break;
        const int wl = __ffs(T.ballot(mine)) - 1;
        const long long rb = T.shfl(rowB, wl), re = T.shfl(rowE, wl);
        if (mine) { // extract_top: the key becomes a tombstone
            M.heap[cb * C + T.lane].hi = kDeadKey;
            khi = kDeadKey;
            atomicSub(&M.cl[cb >> 1], 0x0100u << clShift(cb));
        }
        liveRegs--;
        bool dirty = false; // a key was appended to / killed in bucket cb during this step
        T.sync();
        S.settled++;
        const uint32_t side = cur & 1u;
        const uint32_t u = cur >> 1;
        SMS_CHK((long long)u < G.n);
        const uint32_t hNew = hc + 1;

        for (long long base = rb; base < re; base += TW) {
            if (S.touchedN + TW > M.caps.touchedCap) {
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
                act = arc.nbr >= 0; // blacklisted neighbours are skipped (:98-99)
                if (G.blStamp && act) act = !stampedBefore(G, nbr, now); // mode V: blacklisted by an earlier source
            }
            S.arcs += __popc(T.ballot(act));
            const double dS = dc + w;         // totalStayDist  (:104)
            const double dX = dc + (1.0 - w); // totalCrossDist (:107-108)
            const bool creatS = dS + dc <= best0, creatX = dX + dc <= best0; // label pruning, see ocw_search.cuh

            // ---- label table: bitmap-arbitrated open addressing -------------------------------------------------
            int slot = -1;
            bool isNew = false;
            Entry e = emptyEntry();
            if (act) {
                uint32_t s = homeSlot(M, nbr);
                while (true) {
                    const uint32_t bit = 1u << (s & 31);
                    if (!(M.bm[s >> 5] & bit)) {
                        if (!(creatS || creatX)) break; // no label, and none may be created
                        if (!(atomicOr(&M.bm[s >> 5], bit) & bit)) {
                            isNew = true; // the slot is ours; whatever an earlier source left in it is never read
                            slot = (int)s;
                            break;
                        }
                        // another lane claimed it for its own (different) neighbour in this very step
                    } else {
                        e = loadEntry(&M.tab[s]);
                        if (e.key == nbr && e.flags == stamp) {
                            slot = (int)s;
                            break;
                        }
                    }
                    s = (s + 1) & M.tabMask;
                }
                if (isNew || slot < 0) e = emptyEntry();
            }
            S.touchedN += __popc(T.ballot(isNew));

            bool accS = false, accX = false, evS = false, evX = false, ovf = false;
            double DS = 0, DX = 0;
            uint32_t HS = 0, HX = 0;
            if (slot >= 0) {
                double dA = side ? e.d1 : e.d0, dB = side ? e.d0 : e.d1;
                uint32_t hA = side ? e.h1 : e.h0, hB = side ? e.h0 : e.h1;
                const bool hadA = dA < kUnlabelled, hadB = dB < kUnlabelled;
                // updatePaths(stayNeighbor, ...)   (:59-84)
                if (!hadA)
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
                const bool hadB2 = hadB;
                if (!hadB2)
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
                    e.flags = stamp;
                    storeEntry(&M.tab[slot], e);
#pragma unroll
                    for (int k = 0; k < 2; ++k) {
                        if (!(k ? accX : accS)) continue;
                        const uint32_t sd = k ? (side ^ 1u) : side;
                        const double d = k ? dX : dS;
                        // a superseded key leaves the queue now (the heap engine discards it when it surfaces)
                        if (k ? hadB2 : hadA) {
                            const uint32_t oldq = M.pq[2 * slot + sd].y;
                            if (oldq != kNoQpos) dirty |= killKey(M, oldq) == cb;
                        }
                        uint32_t qpos = kNoQpos;
                        if (2 * d <= best0) { // queue admission: only nodes that can still be extracted
                            const uint32_t bk = min((uint32_t)(d * bscale), NB - 1);
                            const uint32_t sh = clShift(bk);
                            const uint32_t pos = (atomicAdd(&M.cl[bk >> 1], 0x0101u << sh) >> sh) & 0xFFu;
                            dirty |= bk == cb;
                            if (pos >= C) {
                                ovf = true;
                            } else {
                                qpos = bk * C + pos;
                                asm volatile("prefetch.global.L2 [%0];" ::"l"(G.rowptr + nbr)); // read when its bucket loads
                                HeapEnt q;
                                q.hi = (unsigned long long)__double_as_longlong(d);
                                q.lo = ((unsigned long long)hNew << 32) | (2 * nbr + sd);
                                M.heap[qpos] = q;
                            }
                        }
                        M.pq[2 * slot + sd] = make_uint2(cur, qpos);
                    }
                }
            }
            if (T.ballot(ovf)) {
                S.overflow = 1;
                return;
            }
            if (T.ballot(dirty)) needLoad = true;
            // ---- replay of the record logic in adjacency order (:71-82) -------------------------------------------
            unsigned evm = T.ballot(evS || evX);
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
            T.sync(); // keys, labels and counters of this step are visible to the next extraction
        }
        if (liveRegs == 0) needLoad = true;
    }
}

} // namespace smsgpu
