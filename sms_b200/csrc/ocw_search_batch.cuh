// Batched variant of the fast engine (ocw_search_fast.cuh): several keys of the current bucket are extracted in ONE
// step when that is provably the same as extracting them one after the other.
//
// Why: with one key per step a warp relaxes ~8 arcs (average degree of the 1M-node workload) with 32 lanes and then
// waits for a chain of HBM round trips; the profile (profiles/r01_fast2_*.md) shows ~60 % long-scoreboard stalls at
// 20 warps / SM.  Packing the arcs of the k smallest keys of the bucket into the 32 lanes (key rank major, adjacency
// position minor) keeps all lanes busy and pays the latency chain once per k keys.
//
// Exactness.  Let k_0 < k_1 < ... be the keys of the current bucket in extraction order.  Processing k_j in the
// same step as k_0..k_{j-1} equals the sequential order iff
//   (a) no relaxation of k_0..k_{j-1} appends a key to / kills a key in the current bucket (only such a key could
//       be extracted before k_j, and only such a relaxation can change k_j's own label), and
//   (b) no arc of k_j leads to a vertex that an arc of k_0..k_{j-1} leads to (same 32-byte label entry), and
//   (c) the loop condition of ocw_seperator.cpp:88-90 holds for k_j with the bestDist left by the records of
//       k_0..k_{j-1}.
// All three are decided BEFORE anything is committed: every lane computes its relaxation into registers, (a) and
// (b) come from the lanes' own flags and one MATCH, (c) from the serial replay of the record logic, which runs in
// lane order = sequential order.  Keys from the first violating one on stay in the queue untouched (a slot they
// claimed keeps an entry without labels, which is equivalent to no entry).  k_0 is always valid, so every step
// makes progress; a bucket with one key, or a first key with more than 32 arcs, degenerates to the one-key step.
// The pruning / admission threshold best0 is the value at the start of the step for all keys of the batch; it is
// only required to be >= the running bestDist (header of ocw_search.cuh), which it is.
#pragma once
#include "ocw_search_fast.cuh"

namespace smsgpu {

constexpr int kMaxBatch = 8;

template <int TW>
__device__ void searchSourceBatch(const Tile<TW> &T, const GraphDev &G, TileMem &M, SrcState &S, const uint32_t src,
                                  const double bestInit, const uint32_t stamp) {
    static_assert(TW == 32, "the fast engine uses full-warp REDUX / MATCH");
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
    const uint32_t laneLt = (1u << T.lane) - 1u;

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

    uint32_t cb = 0;       // current bucket; extraction keys never decrease
    bool needLoad = true;  // registers do not mirror bucket cb
    uint32_t liveRegs = 0; // live keys among the registers
    unsigned long long khi = kDeadKey, klo = kDeadKey; // lane i: key at position i of bucket cb
    long long rowB = 0, rowE = 0;                      // ... and the arc range of its vertex
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
            S.loads++;
        }

        // ---- select the batch: smallest keys first, as long as their arcs fit into the 32 lanes --------------------
        // lane q keeps the description of key q (kCur, kD, kBase, kDeg); lanes [kBase, kBase + kDeg) relax its arcs
        bool taken = false;      // my register key is part of this batch ...
        uint32_t ownQ = 0;       // ... as key number ownQ
        uint32_t nk = 0, total = 0;
        uint32_t kCur = 0, kBase = 0, kDeg = 0;
        double kD = 0.0;
        uint32_t myQ = 0, curL = 0, hcL = 0; // the key my lane relaxes an arc of
        double dcL = 0.0;
        long long arcL = -1, arcEndL = 0;
        bool stop = false;
        for (int q = 0; q < kMaxBatch; ++q) {
            const bool avail = khi != kDeadKey && !taken;
            const uint32_t k0 = avail ? (uint32_t)(khi >> 32) : 0xFFFFFFFFu;
            const uint32_t m0 = __reduce_min_sync(T.mask, k0);
            if (m0 == 0xFFFFFFFFu) break; // no further live key in the bucket
            bool mine = avail && k0 == m0;
            const uint32_t m1 = __reduce_min_sync(T.mask, mine ? (uint32_t)khi : 0xFFFFFFFFu);
            mine = mine && (uint32_t)khi == m1;
            const uint32_t m2 = __reduce_min_sync(T.mask, mine ? (uint32_t)(klo >> 32) : 0xFFFFFFFFu);
            mine = mine && (uint32_t)(klo >> 32) == m2;
            const uint32_t m3 = __reduce_min_sync(T.mask, mine ? (uint32_t)klo : 0xFFFFFFFFu);
            mine = mine && (uint32_t)klo == m3;
            const double dq = __longlong_as_double((long long)(((unsigned long long)m0 << 32) | m1));
            if (q == 0 && (m3 == target || 2 * dq > S.best)) { // loop condition of ocw_seperator.cpp:88-90
                stop = true;
                break;
            }
            const int wl = __ffs(T.ballot(mine)) - 1;
            const long long rb = T.shfl(rowB, wl), re = T.shfl(rowE, wl);
            const uint32_t deg = (uint32_t)(re - rb);
            if (q > 0 && total + deg > (uint32_t)TW) break; // does not fit; a first key of any degree is chunked
            if (mine) {
                taken = true;
                ownQ = (uint32_t)q;
            }
            if (T.lane == q) {
                kCur = m3;
                kD = dq;
                kBase = total;
                kDeg = deg;
            }
            if ((uint32_t)T.lane >= total && (uint32_t)T.lane < total + deg) {
                myQ = (uint32_t)q;
                curL = m3;
                hcL = m2;
                dcL = dq;
                arcL = rb + (T.lane - total);
                arcEndL = re;
            }
            if (q == 0 && deg > (uint32_t)TW) { // one key, several chunks: lane l takes arcs rb + l, rb + l + 32, ...
                myQ = 0;
                curL = m3;
                hcL = m2;
                dcL = dq;
                arcL = rb + T.lane;
                arcEndL = re;
            }
            total += deg;
            nk++;
            if (total >= (uint32_t)TW) break;
        }
        if (stop) break;
        SMS_CHK(nk >= 1);
        const double best0 = S.best; // conservative (>= the running value) for pruning / queue admission
        uint32_t jstar = nk;         // keys [0, jstar) are extracted in this step
        bool anyDirty = false;

        for (uint32_t cbase = 0; cbase < max(total, 1u); cbase += TW) {
            if (S.touchedN + TW > M.caps.touchedCap) {
                S.overflow = 1;
                return;
            }
            // ---- every lane relaxes one arc of its key into registers (nothing is committed yet) -------------------
            const long long a = arcL + cbase;
            bool act = arcL >= 0 && a < arcEndL;
            uint32_t nbr = 0;
            double w = 0.0;
            if (act) {
                const Arc arc = G.arcs[a];
                nbr = (uint32_t)arc.nbr & 0x7FFFFFFFu;
                w = arc.w;
                act = arc.nbr >= 0; // blacklisted neighbours are skipped (:98-99)
            }
            const uint32_t side = curL & 1u, hNew = hcL + 1;
            const double dS = dcL + w;         // totalStayDist  (:104)
            const double dX = dcL + (1.0 - w); // totalCrossDist (:107-108)
            const bool creatS = dS + dcL <= best0, creatX = dX + dcL <= best0; // label pruning, see ocw_search.cuh

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

            bool accS = false, accX = false, evS = false, evX = false, hadA = false, hadB = false, dirty = false;
            double DS = 0, DX = 0;
            uint32_t HS = 0, HX = 0, oldqS = kNoQpos, oldqX = kNoQpos;
            if (slot >= 0) {
                double dA = side ? e.d1 : e.d0, dB = side ? e.d0 : e.d1;
                uint32_t hA = side ? e.h1 : e.h0, hB = side ? e.h0 : e.h1;
                hadA = dA < kUnlabelled;
                hadB = dB < kUnlabelled;
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
                if (!hadB)
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
                // would this relaxation append to / kill a key in the current bucket?
                if (accS) {
                    if (hadA) oldqS = M.pq[2 * slot + side].y;
                    dirty |= (oldqS != kNoQpos && oldqS / C == cb) || (2 * dS <= best0 && min((uint32_t)(dS * bscale), NB - 1) == cb);
                }
                if (accX) {
                    if (hadB) oldqX = M.pq[2 * slot + (side ^ 1u)].y;
                    dirty |= (oldqX != kNoQpos && oldqX / C == cb) || (2 * dX <= best0 && min((uint32_t)(dX * bscale), NB - 1) == cb);
                }
            }
            // ---- how many keys of the batch may be extracted together: conditions (a) and (b) -----------------------
            if (nk > 1) {
                const unsigned grp = __match_any_sync(T.mask, act ? nbr : (0x80000000u | (uint32_t)T.lane));
                const bool clash = act && (grp & laneLt) != 0u; // an earlier key reaches the same vertex
                const uint32_t cand = clash ? myQ : (dirty ? myQ + 1u : 0xFFFFFFFFu);
                jstar = min(jstar, __reduce_min_sync(T.mask, cand));
                SMS_CHK(jstar >= 1);
            }
            // ---- replay of the record logic (:71-82) key by key, in adjacency order; condition (c) -------------------
            for (uint32_t q = 0; q < jstar; ++q) {
                const uint32_t qc = T.shfl(kCur, (int)q);
                const double qd = T.shfl(kD, (int)q);
                if (q > 0 && (qc == target || 2 * qd > S.best)) {
                    jstar = q;
                    break;
                }
                unsigned evm = T.ballot((evS || evX) && myQ == q);
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
                            if (T.lane == 0) M.hist[S.histN] = 2 * nb + (k ? ((qc & 1u) ^ 1u) : (qc & 1u));
                            S.histN++;
                        }
                    }
                }
            }
            // ---- commit the relaxations of keys [0, jstar) ------------------------------------------------------------
            const bool commit = act && myQ < jstar;
            S.arcs += __popc(T.ballot(commit));
            bool ovf = false;
            if (slot >= 0) {
                if (commit && (accS || accX)) {
                    storeEntry(&M.tab[slot], e);
#pragma unroll
                    for (int k = 0; k < 2; ++k) {
                        if (!(k ? accX : accS)) continue;
                        const uint32_t sd = k ? (side ^ 1u) : side;
                        const double d = k ? dX : dS;
                        const uint32_t oldq = k ? oldqX : oldqS;
                        if (oldq != kNoQpos) killKey(M, oldq); // a superseded key leaves the queue now
                        uint32_t qpos = kNoQpos;
                        if (2 * d <= best0) { // queue admission: only nodes that can still be extracted
                            const uint32_t bk = min((uint32_t)(d * bscale), NB - 1);
                            const uint32_t sh = clShift(bk);
                            const uint32_t pos = (atomicAdd(&M.cl[bk >> 1], 0x0101u << sh) >> sh) & 0xFFu;
                            if (pos >= C) {
                                ovf = true;
                            } else {
                                qpos = bk * C + pos;
                                asm volatile("prefetch.global.L2 [%0];" ::"l"(G.rowptr + nbr)); // read when its bucket loads
                                HeapEnt qk;
                                qk.hi = (unsigned long long)__double_as_longlong(d);
                                qk.lo = ((unsigned long long)hNew << 32) | (2 * nbr + sd);
                                M.heap[qpos] = qk;
                            }
                        }
                        M.pq[2 * slot + sd] = make_uint2(curL, qpos);
                    }
                } else if (isNew) { // claimed but not committed: an entry without labels (equivalent to no entry)
                    Entry z = emptyEntry();
                    z.key = nbr;
                    z.flags = stamp;
                    storeEntry(&M.tab[slot], z);
                }
            }
            if (T.ballot(ovf)) {
                S.overflow = 1;
                return;
            }
            if (T.ballot(dirty && commit)) anyDirty = true;
            if (nk == 1) T.sync(); // several chunks of one key: the next chunk must see this chunk's labels
        }
        // ---- extract_top for the keys that went through: their queue entries become tombstones ----------------------
        if (taken && ownQ < jstar) {
            M.heap[cb * C + T.lane].hi = kDeadKey;
            khi = kDeadKey;
            atomicSub(&M.cl[cb >> 1], 0x0100u << clShift(cb));
        }
        liveRegs -= jstar;
        S.settled += jstar;
        S.steps++;
        T.sync(); // keys, labels and counters of this step are visible to the next one
        if (anyDirty || liveRegs == 0) needLoad = true;
    }
}

} // namespace smsgpu
