// "Pairs" engine of K2 (tier 1): TWO sources per warp, one per half-warp, advanced in LOCK STEP.
//
// Same search as searchSourceFast (ocw_search_fast.cuh) -- runDijsktra of the reference
// (src/graph/ocw_seperator.cpp:37-112) in the canonical order (dist, hops, doubled id), event for event, with the
// monotone bucket queue + tombstones and the shared-memory presence bitmap -- restructured around what the ncu
// profiles of the one-source-per-warp engine showed (profiles/r01_fast2_*.md): ~15 of 32 lanes active (a settled
// node has ~9 arcs), ~450 warp instructions and 3-4 dependent global loads per settled node, 20 sources in flight
// per SM.  Here
//   * a settled node's arcs are relaxed by the 16 lanes of a half-warp and the two halves of a warp work on two
//     different sources with ONE instruction stream: every warp-level primitive is executed by all 32 lanes
//     (ballot / shfl with the full mask, redux with the half's mask), the per-source control flow is predication
//     and a small state machine (idle -> running -> finished), so issue slots per settled node halve and twice as
//     many sources are in flight per resident warp;
//   * queue entries carry the row (first arc, degree) of their vertex -- read from an 8-byte per-vertex record next
//     to the arc, off the critical path -- so an extraction goes straight from the key to the arc load, and the
//     arc row is prefetched into L2 when the key is pushed;
//   * the bucket index is a monotone piecewise-linear 2^u map of the distance instead of a linear one: balls grow
//     geometrically with the radius, the map keeps the number of keys per bucket roughly constant, so half the
//     buckets (and shared memory) hold the same sources and a bucket never needs more than 16 slots;
//   * one REDUX usually decides the minimum (the high words of the distances differ); the full 128-bit
//     lexicographic reduction runs only when they tie.
// When a source ends (target reached, radius exceeded, queue empty, or a capacity hit) the whole warp finishes it
// with the 32-lane code shared with the other engines (selection of intermediates, walks, label export).
// Distances, hops, predecessors, the record logic and therefore the walks are bit-identical to the other engines'
// (tests/test_gpu_parity.py runs every engine against the oracle).
#pragma once
#include "ocw_search_fast.cuh"
#include "ocw_walks.cuh"

namespace smsgpu {

// The kernel's arguments once more in global memory, for the (cold, not inlined) finishing code: taking the
// address of by-value kernel parameters would copy all of them to every thread's stack.
struct PairsKArgs {
    GraphDev G;
    SearchParams P;
    RawOut R;
    TileCaps caps;
};

constexpr uint32_t kNoItem = 0xFFFFFFFFu;
constexpr uint32_t kPackedNone = 0xFFFFFFFFu; // label hop word of the pairs engine: no hops, no queue position
constexpr uint32_t kPairsSaveWords = 32 + 6 * 32; // shared-memory save area of one warp (state of its two searches)

struct PairsParams {
    const PairsKArgs *kargs;
    const uint2 *rowinfo; // [n] {first arc, degree} of every vertex (32-bit arc indices: 2m < 2^32)
    float scaleU;         // bucket map: u = (float)dist * scaleU; 0 selects the linear map
    float scaleB;         // bucket = (2^u - 1) * scaleB   (linear map: dist * scaleB)
    int mirror;           // 1: twin sources (v,0)/(v,1) are searched once when no id tie-break occurred
    uint32_t rowShift;    // Arc::row = first arc | degree << rowShift; a saturated degree field means "read rowinfo"
    const uint32_t *arcs12; // optional 12-byte copy of the arc records {nbr, w} (SMSGPU_ARCS12=1), else null
};

// The per-source working memory (label table, predecessors, queue) is touched once or twice per source and is far
// larger than L2 in total, while the graph (arcs + rows, ~136 MB at 1M/4M) is shared by every source: working-memory
// accesses carry the streaming (evict-first) hint so that they do not push the graph out of L2.
__device__ __forceinline__ Entry loadEntryCS(const Entry *p) {
    Entry e;
    const uint4 *q = reinterpret_cast<const uint4 *>(p);
    uint4 *s = reinterpret_cast<uint4 *>(&e);
#ifdef SMSGPU_PAIRS_PLAIN_LD // experiment: default loads (33.9 G isolated accesses/s against 30.0 G with ld.cs, profiles/r02_sector_flavors.md)
    s[0] = q[0];
    s[1] = q[1];
#else
    s[0] = __ldcs(q);
    s[1] = __ldcs(q + 1);
#endif
    return e;
}
__device__ __forceinline__ void storeEntryCS(Entry *p, const Entry &e) {
    uint4 *q = reinterpret_cast<uint4 *>(p);
    const uint4 *s = reinterpret_cast<const uint4 *>(&e);
    __stcs(q, s[0]);
    __stcs(q + 1, s[1]);
}

// Monotone (non-decreasing in d) bucket index; every operation below is monotone for d >= 0.
__device__ __forceinline__ uint32_t bucketOf(const double d, const float scaleU, const float scaleB, const uint32_t nbm1) {
    const float f = __double2float_rn(d);
    uint32_t b;
    if (scaleU > 0.f) {
        // piecewise-linear 2^u through the IEEE exponent field: bits = u * 2^23 + bias
        const float y = __int_as_float(__float2int_rz(f * scaleU * 8388608.0f) + 0x3F800000);
        b = __float2uint_rz((y - 1.0f) * scaleB);
    } else {
        b = __float2uint_rz(f * scaleB);
    }
    return min(b, nbm1);
}

// Everything after the search of one source, by all 32 lanes (same code path as ocwSearchKernel's tail).
// Returns 1 when the source completed in this tier (its counters count), 0 when it overflowed.
// `M` is the source's working memory as the engine left it (pairs engine: carved from its tile below; lanes engine: built
// by ocwLanesFinishKernel).
__device__ __forceinline__ int finishSourceBody(const PairsKArgs *K, TileMem &M, const double best, const uint32_t histN,
                                            const uint32_t touchedN, const uint32_t arcs, const uint32_t settled, const int overflow,
                                            const uint32_t src, const int32_t srcIdx, const int32_t twinIdx, const int mirrorOk) {
    // twinIdx >= 0: the next entry of the source list is this source's twin (other side of the same vertex).  Its results
    // are emitted from this search when mirrorOk; if this source overflows, the twin goes to the next tier with it (its
    // search is the mirror image and would hit the same capacity).  Returns 0 / 1 / 2 = sources disposed of... see below.
    const int32_t srcIdx2 = (twinIdx >= 0 && mirrorOk && !overflow) ? twinIdx : -1;
    const Tile<32> T;
    const GraphDev G = K->G;
    const TileCaps caps = M.caps;
    const RawOut &R = K->R;
    const SearchParams &P = K->P;
    SrcState S;
    S.best = best;
    S.bestHop = 0.0;
    S.histN = histN;
    S.heapN = 0;
    S.touchedN = touchedN;
    S.arcs = arcs;
    S.settled = settled;
    S.steps = 0;
    S.loads = 0;
    S.overflow = overflow;
    uint32_t sel[16];
    const int maxInter = min(P.maxInter, 16);
    if (!S.overflow && S.histN > 0) {
        const int nsel = selectIntermediates<32>(T, M, S, maxInter, sel);
        if (!S.overflow) { // all 32 lanes (the chord search is spread over them); lane 0 owns the emitter
#pragma unroll
            for (int r = 0; r < 16; ++r) sel[r] = T.shfl(sel[r], 0); // lane 0 holds the selection
            // srcIdx2 >= 0: the twin source (other side of the same vertex) gets the same walks -- its search is the mirror
            // image of this one, and a walk is a sequence of ORIGINAL vertices and stay/cross flags.  The cuts of the first
            // pass are logged and copied; only if there are more than the log holds, the walks are built a second time.
            constexpr int kLog = 64;
            long long logCut[kLog];
            int32_t logLen[kLog];
            int nlogged = 0;
            for (int pass = 0; pass < (srcIdx2 >= 0 ? 2 : 1) && !S.overflow; ++pass) {
                Emitter E;
                E.R = &R;
                E.srcIdx = pass ? srcIdx2 : srcIdx;
                E.seq = 0;
                E.tier = P.tier;
                if (pass == 0 && srcIdx2 >= 0) {
                    E.logCut = logCut;
                    E.logLen = logLen;
                    E.logCap = kLog;
                }
                if (pass == 1 && nlogged <= kLog) { // copy
                    if (T.lane == 0)
                        for (int i = 0; i < nlogged; ++i) {
                            long long cut, off;
                            if (emitReserve(E, logLen[i], cut, off) && logCut[i] >= 0) {
                                const long long so = R.cutOff[logCut[i]];
                                for (int k = 0; k < logLen[i]; ++k) {
                                    R.node[off + k] = R.node[so + k];
                                    R.eid[off + k] = R.eid[so + k];
                                    R.stay[off + k] = R.stay[so + k];
                                }
                            }
                        }
                    break;
                }
                walksOfSource<true>(G, M, S, E, sel, nsel, src, srcIdx);
                nlogged = T.shfl(E.nlog, 0);
            }
        }
        S.overflow = T.shfl(S.overflow, 0);
    }
    int counted = 0;
    if (!S.overflow && G.blStampNext) verbatimSourceDone<32>(T, G, src, srcIdx, histN == 0); // mode V (never mirrored)
    if (!S.overflow) {
        if (P.exportLabels) {
            for (uint32_t i = T.lane; i < caps.tabCap; i += 32) {
                if (M.bm ? !((M.bm[i >> 5] >> (i & 31)) & 1u) : (M.tab[i].flags != M.stamp)) continue;
                const Entry e = loadEntry(&M.tab[i]);
#pragma unroll
                for (int sd = 0; sd < 2; ++sd) {
                    const double d = sd ? e.d1 : e.d0;
                    if (d >= kUnlabelled) continue;
                    for (int pass = 0; pass < (srcIdx2 >= 0 ? 2 : 1); ++pass) { // the twin's labels: sides swapped
                        const long long li = (long long)atomicAdd(&R.counters[5], 1ull);
                        if (li < R.labCap) {
                            R.labSrc[li] = pass ? srcIdx2 : srcIdx;
                            R.labNode[li] = (int32_t)(2 * e.key + (sd ^ pass));
                            R.labDist[li] = d;
                            R.labHops[li] = (int32_t)((sd ? e.h1 : e.h0) & M.hopMask);
                        }
                    }
                }
            }
        }
        if (T.lane == 0) {
            for (int pass = 0; pass < (srcIdx2 >= 0 ? 2 : 1); ++pass) {
                const int32_t si = pass ? srcIdx2 : srcIdx;
                R.srcTier[si] = (int8_t)P.tier;
                if (R.srcWork) R.srcWork[si] = settled;
                if (R.srcArcs) {
                    R.srcArcs[si] = (long long)arcs;
                    R.srcSettled[si] = (long long)settled;
                }
            }
        }
        counted = 1;
        if (T.lane == 0) { // round totals
            const unsigned long long mult = srcIdx2 >= 0 ? 2ull : 1ull; // what the reference would have scanned / settled
            atomicAdd(&R.counters[2], mult * arcs);
            atomicAdd(&R.counters[3], mult * settled);
            atomicAdd(&R.counters[4], mult);
            if (srcIdx2 >= 0) { // ... of which obtained by symmetry, not scanned
                atomicAdd(&R.counters[6], 1ull);
                atomicAdd(&R.counters[7], (unsigned long long)arcs);
            }
        }
        if (srcIdx2 >= 0) counted = 2;
    } else {
        if (T.lane == 0) {
            const unsigned long long k = atomicAdd(P.ovfCount, twinIdx >= 0 ? 2ull : 1ull);
            P.ovfList[k] = srcIdx;
            if (twinIdx >= 0) P.ovfList[k + 1] = twinIdx; // adjacent: the next tier can pair them again
        }
        if (twinIdx >= 0) counted = 2;
    }
    T.sync();
    return counted; // 2: the twin is disposed of as well (mirrored, or sent to the next tier with this source)
}

__device__ __noinline__ int finishSourcePairs(const PairsKArgs *K, unsigned char *tile, unsigned char *hot, const double best,
                                              const uint32_t histN, const uint32_t touchedN, const uint32_t arcs,
                                              const uint32_t settled, const int overflow, const uint32_t src,
                                              const int32_t srcIdx, const int32_t twinIdx, const int mirrorOk) {
    TileMem M;
    carveTileMem(M, tile, K->caps, hot);
    return finishSourceBody(K, M, best, histN, touchedN, arcs, settled, overflow, src, srcIdx, twinIdx, mirrorOk);
}
// the same for working memory some other engine describes (lanes engine)
__device__ __noinline__ int finishSourceMem(const PairsKArgs *K, TileMem &M, const double best, const uint32_t histN,
                                            const uint32_t touchedN, const uint32_t arcs, const uint32_t settled, const int overflow,
                                            const uint32_t src, const int32_t srcIdx, const int32_t twinIdx, const int mirrorOk) {
    return finishSourceBody(K, M, best, histN, touchedN, arcs, settled, overflow, src, srcIdx, twinIdx, mirrorOk);
}

#ifndef SMSGPU_PAIRS_MIN_CTAS
#define SMSGPU_PAIRS_MIN_CTAS 6
#endif

#ifdef SMSGPU_PAIRS_CLOCKS // instrumentation build: cycles per phase of the lock-step loop, printed by one warp
#define PCLK(i_)                                  \
    do {                                          \
        const long long now_ = clock64();         \
        pclk[i_] += now_ - plast;                 \
        plast = now_;                             \
    } while (0)
#else
#define PCLK(i_) ((void)0)
#endif

// VERB: SMSGPU_MODE_VERBATIM (time-stamped blacklist); a template parameter so that the independent-source build of the hot
// loop carries none of its tests (measured: the run-time test cost the 1M-source round 1.5 %)
template <bool VERB>
__global__ void __launch_bounds__(128, SMSGPU_PAIRS_MIN_CTAS)
    ocwSearchPairsKernel(const GraphDev G, const PairsParams Q, const SearchParams P, const RawOut R, const TileCaps caps,
                         unsigned char *scratch) {
    extern __shared__ __align__(32) uint32_t smemw[];
    constexpr unsigned FULL = 0xFFFFFFFFu;
    constexpr uint32_t TW = 16;
    const int lane = threadIdx.x & 31;
    const int gbase = lane & 16;   // first warp lane of this half
    const uint32_t tl = lane & 15; // lane within the half
    const unsigned gmask = 0xFFFFu << gbase;
    const int warpInCta = threadIdx.x >> 5;
    const size_t tb = tileBytes(caps);
    const uint32_t hbw = (uint32_t)(hotBytes(caps) >> 2); // hot words of one half
    unsigned char *const tileB = scratch + (((size_t)blockIdx.x * (blockDim.x >> 5) + warpInCta) * 2 + (gbase >> 4)) * tb;

    // this half's working memory: offsets inside the tile as carveTileMem lays it out (uniform: functions of caps)
    uint32_t offPq, offHist, offQrow;
    {
        TileMem M;
        carveTileMem(M, tileB, caps, nullptr);
        offPq = (uint32_t)(reinterpret_cast<unsigned char *>(M.pq) - tileB);
        offHist = (uint32_t)(reinterpret_cast<unsigned char *>(M.hist) - tileB);
        offQrow = (uint32_t)(reinterpret_cast<unsigned char *>(M.qrow) - tileB);
    }
#define TABP(s_) (reinterpret_cast<Entry *>(tileB) + (s_))
#define PQP(i_) (reinterpret_cast<uint2 *>(tileB + offPq) + (i_))
#define QP(i_) (reinterpret_cast<uint4 *>(tileB + offQrow) + 2 * (i_)) /* queue entry i: [0] key, [1] row */
#define HISTP(i_) (reinterpret_cast<uint32_t *>(tileB + offHist) + (i_))
    const uint32_t NB = caps.nbuckets;
    constexpr uint32_t C = 16, CAP = 15; // queue positions per bucket (stride) / usable ones (the counter is a nibble)
    const uint32_t nclWords = NB >> 2, nbmWords = caps.tabCap >> 5;
    uint32_t *const cl = smemw + (uint32_t)(warpInCta * 2 + (gbase >> 4)) * hbw; // bucket counters (shared memory)
    uint32_t *const bm = cl + nclWords;                                          // presence bitmap (shared memory)
    const uint32_t tabMask = caps.tabCap - 1, tabShift = caps.tabShift;
    const uint32_t total = (uint32_t)(P.worklist ? P.nwork : P.nsrc);

    // State of the two searches of this warp.  Registers inside the stepping loop; saved to shared memory around the
    // (cold, register-hungry, not inlined) finishing call so that nothing of it is live across that call -- otherwise
    // the register allocator spills it at its definitions and uses, i.e. inside the hot loop (measured: half of the
    // L1->L2 traffic of the first version of this kernel was spill traffic).
    //   per half (16 words, written by its lane 0): st src wi best(2) bestHop histN touchedN arcsN settledN overflow
    //                                               cb lo liveRegs needLoad
    //   per lane (6 words): the current bucket's key and the row of its vertex
    uint32_t *const svU = smemw + (uint32_t)(blockDim.x >> 5) * 2 * hbw + (uint32_t)warpInCta * kPairsSaveWords + (gbase >> 4) * 16;
    uint32_t *const svK = smemw + (uint32_t)(blockDim.x >> 5) * 2 * hbw + (uint32_t)warpInCta * kPairsSaveWords + 32 + lane;
    if (tl == 0) {
        svU[0] = 0; // st: idle
        svU[14] = 1;
        svU[15] = 0;
    }
    __syncwarp(FULL);
#ifdef SMSGPU_PAIRS_CLOCKS
    long long pclk[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, plast = clock64(), piters = 0;
#endif

    while (true) {
        // per-source state, identical in the 16 lanes of a half
        int st = (int)svU[0]; // 0 idle (needs a source), 1 running, 2 ended (to be finished), 3 no work left
        uint32_t src = svU[1], wi = svU[2];
        double best = __hiloint2double((int)svU[4], (int)svU[3]);
        uint32_t bestHop = svU[5]; // bestHopDist: a double in the reference (:52) that only ever holds small integers
        uint32_t histN = svU[6], touchedN = svU[7], arcsN = svU[8], settledN = svU[9];
        int overflow = (int)svU[10];
        uint32_t cb = svU[11], lo = svU[12], liveRegs = svU[13]; // bucket in registers, bucket of the last extraction, live keys
        bool needLoad = svU[14] != 0;
        // the current bucket's keys, one per lane, with the row of their vertex
        unsigned long long khi = ((unsigned long long)svK[32] << 32) | svK[0], klo = ((unsigned long long)svK[96] << 32) | svK[64];
        uint32_t rowB = svK[128], deg = svK[160];
        uint32_t pend = svU[15] & 3u; // second source of the item: 0 none / done, 1 to be searched next, 2 mirror candidate
        bool tie = (svU[15] >> 2) != 0; // an extraction of the current search had to break a (dist, hops) tie by id
        bool freshB = false; // the keys in registers were loaded since the last extraction
        bool done = false;
        unsigned finb = 0;

      while (true) { // the stepping loop: leaves when a half has ended its source (or no work is left)
        finb = __ballot_sync(FULL, st == 2);
        if (finb) break;

        // ---- (2) idle halves fetch the next source ------------------------------------------------------------------
        if (__ballot_sync(FULL, st == 0)) {
            // A work item is a PAIR (2p, 2p+1) of consecutive entries of the source list -- in the handler's list the
            // two sides (v,0), (v,1) of one vertex.  Their searches are mirror images of each other under the side swap
            // (u,s) -> (u,1-s): same distances, hops, events and therefore walks, as long as no decision depends on the
            // doubled ids, i.e. as long as no extraction has to break a (dist, hops) tie by id.  So a half searches the
            // first source of its item; if that search saw no tie and the second source is the first one's twin, the
            // finishing code emits the second's results from the same labels (pend == 2 -> mirrored); otherwise the
            // second source is searched right afterwards on the same half (pend == 1).
            const bool second = st == 0 && pend == 1; // the second source of the item is due: no fetch
            uint32_t w32 = 0;
            if (st == 0 && !second && tl == 0) {
                const unsigned long long pr = atomicAdd(P.queueCounter, 1ull);
                w32 = 2 * pr < (unsigned long long)total ? (uint32_t)(2 * pr) : total;
            }
            w32 = __shfl_sync(FULL, w32, gbase);
            bool fresh = false;
            if (st == 0) {
                if (second) w32 = wi + 1;
                if (w32 >= total) {
                    st = 3;
                } else {
                    const int32_t i0 = P.worklist ? P.worklist[w32] : (int32_t)w32;
                    const uint32_t s0 = P.sources ? (uint32_t)P.sources[i0] : (uint32_t)(((((long long)(i0 >> 1)) * P.srcStride + P.srcBegin) << 1) | (i0 & 1));
                    // `if (blacklist_[currentSource_]) return {}` -- indexed with the DOUBLED id (ocw_seperator.cpp:334)
                    const bool skip0 = ((long long)s0 < G.n && (VERB ? stampedBefore(G, s0, i0) : P.blacklist[s0] != 0)) || (long long)(s0 >> 1) >= G.n;
                    if (second) { // skip test already passed when the item was fetched
                        wi = w32;
                        src = s0;
                        pend = 0;
                        fresh = true;
                    } else {
                        bool have1 = w32 + 1 < total;
                        uint32_t s1 = 0;
                        if (have1) {
                            const int32_t i1 = P.worklist ? P.worklist[w32 + 1] : (int32_t)(w32 + 1);
                            s1 = P.sources ? (uint32_t)P.sources[i1] : (uint32_t)(((((long long)(i1 >> 1)) * P.srcStride + P.srcBegin) << 1) | (i1 & 1));
                            have1 = !(((long long)s1 < G.n && (VERB ? stampedBefore(G, s1, i1) : P.blacklist[s1] != 0)) || (long long)(s1 >> 1) >= G.n);
                        }
                        if (!skip0) {
                            wi = w32;
                            src = s0;
                            pend = have1 ? ((s1 == (s0 ^ 1u) && Q.mirror) ? 2u : 1u) : 0u;
                            fresh = true;
                        } else if (have1) {
                            wi = w32 + 1;
                            src = s1;
                            pend = 0;
                            fresh = true;
                        } // else: both skipped, fetch again
                    }
                    if (fresh) {
                        tie = false;
                        for (uint32_t i = tl; i < nclWords + nbmWords; i += TW) cl[i] = 0; // counters, then the bitmap
                    }
                }
            }
            __syncwarp(FULL);
            if (fresh) {
                if (tl == 0) { // label and enqueue the source (:54-57)
                    const uint32_t s = ((src >> 1) * 0x9E3779B1u) >> tabShift;
                    bm[s >> 5] = 1u << (s & 31);
                    Entry e = emptyEntry();
                    e.key = src >> 1;
                    e.flags = P.stampBase + wi + 1u;
                    e.h0 = e.h1 = kPackedNone; // hop words of this engine: hops | queue position << 16
                    if (src & 1u) {
                        e.d1 = 0.0;
                        e.h1 = 0; // 0 hops, queue position 0
                    } else {
                        e.d0 = 0.0;
                        e.h0 = 0;
                    }
                    storeEntryCS(TABP(s), e);
                    __stcs(&PQP(2 * s + (src & 1u))->x, src);
                    const uint2 srow = Q.rowinfo[src >> 1];
                    __stcs(QP(0), make_uint4(0u, 0u, src, 0u)); // dist 0, id src, 0 hops
                    __stcs(QP(0) + 1, make_uint4(srow.x, srow.y, src, 0u)); // row, predecessor: itself (never followed)
                    cl[0] = 0x11u; // bucket 0: one appended, one live
                }
                st = 1;
                best = P.bestInit;
                bestHop = 0;
                histN = 0;
                touchedN = 1;
                arcsN = 0;
                settledN = 0;
                overflow = 0;
                cb = 0;
                lo = 0;
                liveRegs = 0;
                needLoad = true;
                khi = kDeadKey;
                klo = kDeadKey;
            }
            __syncwarp(FULL);
        }
        if (__ballot_sync(FULL, st == 1) == 0) {
            if (__ballot_sync(FULL, st != 3) == 0) { // both halves are out of work
                done = true;
                break;
            }
            continue; // a skipped source: fetch again
        }

        PCLK(0);
        // ---- (3) (re)load the current bucket into registers -----------------------------------------------------------
        const bool needL = (st == 1) && needLoad;
        if (needL) cb = lo; // registers are stale or empty: rescan from the bucket of the last extraction
        if (__ballot_sync(FULL, needL)) {
            uint32_t b = NB, cnt = 0;
            bool searching = needL;
            uint32_t base = cb >> 2;
            while (__ballot_sync(FULL, searching)) {
                const uint32_t wd = base + tl;
                const uint32_t word = (searching && wd < nclWords) ? cl[wd] : 0u;
                uint32_t live = word & 0xF0F0F0F0u; // four buckets per word: low nibble appended, high nibble live
                if (wd == (cb >> 2)) live &= 0xFFFFFFFFu << ((cb & 3u) * 8u);
                const unsigned gb = (__ballot_sync(FULL, live != 0u) >> gbase) & 0xFFFFu;
                const int l = gb ? (__ffs(gb) - 1) : 0;
                const uint32_t ww = __shfl_sync(FULL, word, gbase + l);
                if (searching) {
                    if (gb) {
                        uint32_t lw = ww & 0xF0F0F0F0u;
                        if ((base + l) == (cb >> 2)) lw &= 0xFFFFFFFFu << ((cb & 3u) * 8u);
                        const uint32_t byte = (uint32_t)(__ffs(lw) - 1) >> 3;
                        b = 4 * (base + l) + byte;
                        cnt = (ww >> (byte * 8u)) & 0xFu;
                        searching = false;
                    } else {
                        base += TW;
                        if (base >= nclWords) searching = false; // queue empty
                    }
                }
            }
            unsigned long long nhi = kDeadKey, nlo = kDeadKey;
            uint2 nrow = make_uint2(0u, 0u);
            if (needL && b < NB && tl < cnt) {
                const uint4 k = __ldcs(QP(b * C + tl)), r = __ldcs(QP(b * C + tl) + 1);
                nrow = make_uint2(r.x, r.y);
                nhi = ((unsigned long long)k.y << 32) | k.x;
                nlo = ((unsigned long long)k.w << 32) | k.z;
            }
            const unsigned lv = (__ballot_sync(FULL, nhi != kDeadKey) >> gbase) & 0xFFFFu;
            if (needL) {
                if (b >= NB) {
                    st = 2; // queue empty (:88)
                } else {
                    cb = b;
                    khi = nhi;
                    klo = nlo;
                    rowB = nrow.x;
                    deg = nrow.y;
                    liveRegs = __popc(lv);
                    needLoad = false;
                    freshB = true;
                    SMS_CHK(liveRegs > 0);
                }
            }
        }

        PCLK(1);
        // ---- (4) extract the minimum key of the bucket: (dist, hops, id) lexicographic -------------------------------
        bool run = (st == 1);
        const uint32_t k0 = (run && khi != kDeadKey) ? (uint32_t)(khi >> 32) : 0xFFFFFFFFu;
        const uint32_t m0 = __reduce_min_sync(gmask, k0);
        bool mine = run && k0 == m0 && khi != kDeadKey;
        unsigned gbm = (__ballot_sync(FULL, mine) >> gbase) & 0xFFFFu;
        if (__ballot_sync(FULL, (gbm & (gbm - 1)) != 0)) { // high words tie somewhere: full 128-bit reduction
            const uint32_t k1 = (uint32_t)khi, k2 = (uint32_t)(klo >> 32), k3 = (uint32_t)klo;
            const uint32_t m1 = __reduce_min_sync(gmask, mine ? k1 : 0xFFFFFFFFu);
            mine = mine && k1 == m1;
            const uint32_t m2 = __reduce_min_sync(gmask, mine ? k2 : 0xFFFFFFFFu);
            mine = mine && k2 == m2;
            const unsigned g2 = (__ballot_sync(FULL, mine) >> gbase) & 0xFFFFu;
            if (run && (g2 & (g2 - 1)) != 0) tie = true; // equal (dist, hops): the doubled id decides -> not mirror invariant
            const uint32_t m3 = __reduce_min_sync(gmask, mine ? k3 : 0xFFFFFFFFu);
            mine = mine && k3 == m3;
            gbm = (__ballot_sync(FULL, mine) >> gbase) & 0xFFFFu;
        }
        SMS_CHK(!run || gbm != 0);
        const int wl = gbase + (gbm ? (__ffs(gbm) - 1) : 0);
        const unsigned long long chi = __shfl_sync(FULL, khi, wl), clo = __shfl_sync(FULL, klo, wl);
        const uint32_t rb = __shfl_sync(FULL, rowB, wl), dg = __shfl_sync(FULL, deg, wl);
        const uint32_t cur = (uint32_t)clo;
        const double dc = __longlong_as_double((long long)chi);
        // loop condition of ocw_seperator.cpp:88-90
        if (run && (cur == (src ^ 1u) || 2 * dc > best)) {
            st = 2;
            run = false;
        }
        if (run) {
            if (lane == wl) { // extract_top: the key becomes a tombstone
                __stcs(reinterpret_cast<unsigned long long *>(QP(cb * C + tl)), kDeadKey);
                khi = kDeadKey;
                atomicSub(&cl[cb >> 2], 0x10u << ((cb & 3u) * 8u));
            }
            liveRegs--;
            settledN++;
        }
        __syncwarp(FULL); // the tombstone and the counter update are visible to the scan below
        // first extraction from a bucket: the rows of its other keys are needed within the next few steps -- start
        // them towards L2 now (the row of the key just extracted is being loaded anyway)
        if (freshB) {
            if (run && khi != kDeadKey && deg) {
#ifdef SMSGPU_PAIRS_ARCS12
                asm volatile("prefetch.global.L2 [%0];" ::"l"(Q.arcs12 + 3 * (size_t)rowB));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(Q.arcs12 + 3 * (size_t)(rowB + deg - 1)));
#else
                asm volatile("prefetch.global.L2 [%0];" ::"l"(G.arcs + rowB));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(G.arcs + rowB + deg - 1));
#endif
            }
            freshB = false;
        }
        const uint32_t side = cur & 1u;
        const uint32_t hNew = (uint32_t)(clo >> 32) + 1;
        const uint32_t stamp = P.stampBase + wi + 1u;
        int32_t nowIdx = 0; // mode V: position of this half's source in the source list (its time stamp)
        if (VERB && run) nowIdx = P.worklist ? P.worklist[wi] : (int32_t)wi;

        PCLK(2);
        // ---- (4b) bucket exhausted by this extraction: find the next one and issue its loads NOW, so that they are in
        // flight during the relax phase below instead of in front of the next extraction.  Nothing here waits for the
        // loaded keys (the number of live keys comes from the counters).  A push of the relax phase that lands at or
        // before the bucket loaded here makes the registers stale: `dirty` below, and the slow path (3) rescans from
        // `lo`, the bucket of the node being settled.
        lo = cb;
        const bool early = run && liveRegs == 0;
        if (__ballot_sync(FULL, early)) {
            uint32_t b = NB, cnt = 0, nlive = 0;
            bool searching = early;
            uint32_t base = cb >> 2;
            while (__ballot_sync(FULL, searching)) {
                const uint32_t wd = base + tl;
                const uint32_t word = (searching && wd < nclWords) ? cl[wd] : 0u;
                uint32_t live = word & 0xF0F0F0F0u; // four buckets per word: low nibble appended, high nibble live
                if (wd == (cb >> 2)) live &= 0xFFFFFFFFu << ((cb & 3u) * 8u);
                const unsigned gb = (__ballot_sync(FULL, live != 0u) >> gbase) & 0xFFFFu;
                const int l = gb ? (__ffs(gb) - 1) : 0;
                const uint32_t ww = __shfl_sync(FULL, word, gbase + l);
                if (searching) {
                    if (gb) {
                        uint32_t lw = ww & 0xF0F0F0F0u;
                        if ((base + l) == (cb >> 2)) lw &= 0xFFFFFFFFu << ((cb & 3u) * 8u);
                        const uint32_t byte = (uint32_t)(__ffs(lw) - 1) >> 3;
                        b = 4 * (base + l) + byte;
                        cnt = (ww >> (byte * 8u)) & 0xFu;
                        nlive = (ww >> (byte * 8u + 4u)) & 0xFu;
                        searching = false;
                    } else {
                        base += TW;
                        if (base >= nclWords) searching = false; // nothing queued right now
                    }
                }
            }
            if (early) {
                if (b >= NB) {
                    needLoad = true; // empty unless this step pushes something: (3) decides after the relax phase
                } else {
                    khi = kDeadKey;
                    klo = kDeadKey;
                    if (tl < cnt) {
                        const uint4 k = __ldcs(QP(b * C + tl)), r = __ldcs(QP(b * C + tl) + 1);
                        khi = ((unsigned long long)k.y << 32) | k.x;
                        klo = ((unsigned long long)k.w << 32) | k.z;
                        rowB = r.x;
                        deg = r.y;
                    }
                    cb = b;
                    liveRegs = nlive;
                    freshB = true;
                }
            }
        }

        PCLK(3);
        // ---- (5) relax the arcs of the settled node, 16 per pass ------------------------------------------------------
        for (uint32_t c0 = 0; __ballot_sync(FULL, run && c0 < dg); c0 += TW) {
            if (run && touchedN + TW > caps.touchedCap) {
                if (tl == 0) atomicAdd(&R.counters[12], 1ull); // overflow reasons (diagnostics): label table
                overflow = 1;
                st = 2;
                run = false;
            }
            const double best0 = best; // conservative (>= the running value) for pruning / queue admission
            bool act = run && (c0 + tl < dg);
            uint32_t nbr = 0;
#ifdef SMSGPU_PAIRS_ROW_IN_ARC
            uint32_t rowp = 0;
#endif
            double w = 0.0;
            uint2 nrow = make_uint2(0u, 0u);
            if (act) {
#ifdef SMSGPU_PAIRS_ARCS12 // experiment: the 12-byte copy of the arc records (96 MB at config 5: inside the 126 MB of L2)
                const uint32_t *ap = Q.arcs12 + 3 * (size_t)(rb + c0 + tl);
                Arc arc;
                arc.nbr = (int32_t)__ldg(ap);
                arc.w = __hiloint2double((int)__ldg(ap + 2), (int)__ldg(ap + 1));
#else
                const Arc arc = G.arcs[rb + c0 + tl];
#endif
                nbr = (uint32_t)arc.nbr & 0x7FFFFFFFu;
#ifdef SMSGPU_PAIRS_ROW_IN_ARC
                rowp = arc.row;
#endif
                w = arc.w;
                act = arc.nbr >= 0;    // blacklisted neighbours are skipped (:98-99)
                if (VERB && act) act = !stampedBefore(G, nbr, nowIdx); // mode V: blacklisted by an earlier source
            }
            arcsN += __popc((__ballot_sync(FULL, act) >> gbase) & 0xFFFFu);
            PCLK(4);
            const double dS = dc + w;         // totalStayDist  (:104)
            const double dX = dc + (1.0 - w); // totalCrossDist (:107-108)
            const bool creatS = dS + dc <= best0, creatX = dX + dc <= best0; // label pruning, see ocw_search.cuh

            // ---- label table: bitmap-arbitrated open addressing ---------------------------------------------------
            int slot = -1;
            bool isNew = false;
            double dA = kUnlabelled, dB = kUnlabelled; // label of the side the stay arc reaches / of the other side
            uint32_t pA = kPackedNone, pB = kPackedNone; // hops | queue position << 16 of the two sides
            if (act) {
                uint32_t s = (nbr * 0x9E3779B1u) >> tabShift;
#ifndef SMSGPU_PAIRS_PROBE1 // default: two slots per round trip (-11 % on the 1M/4M round); PROBE1 = one slot
                // two slots per round trip: the 32 lanes wait for the slowest probe sequence of the pass
                while (true) {
                    const uint32_t bit = 1u << (s & 31);
                    if (!(bm[s >> 5] & bit)) {
                        if (!(creatS || creatX)) break;
                        if (!(atomicOr(&bm[s >> 5], bit) & bit)) {
                            isNew = true;
                            slot = (int)s;
                            break;
                        }
                        s = (s + 1) & tabMask;
                        continue;
                    }
                    const uint32_t s1 = (s + 1) & tabMask;
                    const bool occ1 = (bm[s1 >> 5] >> (s1 & 31)) & 1u;
                    Entry e = loadEntryCS(TABP(s));
                    if (occ1) {
                        const Entry e1 = loadEntryCS(TABP(s1));
                        if (!(e.key == nbr && e.flags == stamp) && e1.key == nbr && e1.flags == stamp) {
                            e = e1;
                            s = s1;
                        }
                    }
                    if (e.key == nbr && e.flags == stamp) {
                        slot = (int)s;
                        dA = side ? e.d1 : e.d0;
                        dB = side ? e.d0 : e.d1;
                        pA = side ? e.h1 : e.h0;
                        pB = side ? e.h0 : e.h1;
                        break;
                    }
                    s = occ1 ? ((s1 + 1) & tabMask) : s1;
                }
#else
                while (true) {
                    const uint32_t bit = 1u << (s & 31);
                    if (!(bm[s >> 5] & bit)) {
                        if (!(creatS || creatX)) break; // no label, and none may be created
                        if (!(atomicOr(&bm[s >> 5], bit) & bit)) {
                            isNew = true; // the slot is ours; whatever an earlier source left in it is never read
                            slot = (int)s;
                            break;
                        }
                        // another lane claimed it for its own (different) neighbour in this very step
                    } else {
                        const Entry e = loadEntryCS(TABP(s));
                        if (e.key == nbr && e.flags == stamp) {
                            slot = (int)s;
                            dA = side ? e.d1 : e.d0;
                            dB = side ? e.d0 : e.d1;
                            pA = side ? e.h1 : e.h0;
                            pB = side ? e.h0 : e.h1;
                            break;
                        }
                    }
                    s = (s + 1) & tabMask;
                }
#endif
            }
            touchedN += __popc((__ballot_sync(FULL, isNew) >> gbase) & 0xFFFFu);
            PCLK(5);

            bool accS = false, accX = false, evS = false, evX = false;
            const bool hadA = dA < kUnlabelled, hadB = dB < kUnlabelled;
            uint32_t hA = pA & 0xFFFFu, hB = pB & 0xFFFFu;
            double DS = 0, DX = 0;
            uint32_t HS = 0, HX = 0;
            if (slot >= 0) {
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
            }
            // ---- replay of the record logic in adjacency order (:71-82), before the stores so that its operands die ----
            unsigned evm = (__ballot_sync(FULL, evS || evX) >> gbase) & 0xFFFFu;
            if (!run) evm = 0;
            while (__ballot_sync(FULL, evm != 0)) {
                const bool have = evm != 0;
                const int l = gbase + (have ? (__ffs(evm) - 1) : 0);
                evm &= evm - 1; // 0 stays 0
                const uint32_t nb = __shfl_sync(FULL, nbr, l);
                const int eS = __shfl_sync(FULL, (int)evS, l), eX = __shfl_sync(FULL, (int)evX, l);
                const double D0 = __shfl_sync(FULL, DS, l), D1 = __shfl_sync(FULL, DX, l);
                const uint32_t H0 = __shfl_sync(FULL, HS, l), H1 = __shfl_sync(FULL, HX, l);
                if (have) {
#pragma unroll
                    for (int k = 0; k < 2; ++k) {
                        if (!(k ? eX : eS)) continue;
                        const double D = k ? D1 : D0;
                        const uint32_t H = k ? H1 : H0;
                        bool push = false;
                        if (D < best || (D == best && H < bestHop)) {
                            best = D;
                            bestHop = H;
                            push = true;
                        } else if (D == best && H < 2 * bestHop) {
                            push = true;
                        }
                        if (push && !overflow) {
                            if (histN >= caps.histCap) {
                                if (tl == 0) atomicAdd(&R.counters[14], 1ull); // history of intermediates
                                overflow = 1;
                                st = 2;
                                run = false;
                                evm = 0;
                            } else {
                                if (tl == 0) *HISTP(histN) = 2 * nb + (k ? (side ^ 1u) : side);
                                histN++;
                            }
                        }
                    }
                }
            }
            PCLK(6);
            // ---- store the label, retire superseded keys, push the new ones ---------------------------------------
            bool ovf = false, dirty = false;
            // the row of the neighbour travels with its key: one 8-byte read per PUSHING lane (about one per settled node).
            // The arc record carries the same information packed (Arc::row), but keeping it live across the probe costs the
            // hot loop registers: measured 490 -> 520 ms per round (profiles/README.md, round 2), so the row record is read.
            if ((accS && 2 * dS <= best0) || (accX && 2 * dX <= best0)) {
#ifdef SMSGPU_PAIRS_ROW_IN_ARC
                nrow = make_uint2(rowp & ((1u << Q.rowShift) - 1u), rowp >> Q.rowShift);
                if (nrow.y == (0xFFFFFFFFu >> Q.rowShift)) nrow = Q.rowinfo[nbr];
#else
                nrow = Q.rowinfo[nbr];
#endif
            }
            if (accS || accX) {
                if (hNew >= 0xFFFFu) ovf = true; // hops share a word with the queue position in this engine
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    if (!(k ? accX : accS)) continue;
                    const uint32_t sd = k ? (side ^ 1u) : side;
                    const double d = k ? dX : dS;
                    // a superseded key leaves the queue now (its position sits in the label's hop word: no extra load)
                    const uint32_t oldq = (k ? pB : pA) >> 16;
                    if ((k ? hadB : hadA) && oldq != 0xFFFFu) {
                        __stcs(reinterpret_cast<unsigned long long *>(QP(oldq)), kDeadKey);
                        const uint32_t ob = oldq / C;
                        atomicSub(&cl[ob >> 2], 0x10u << ((ob & 3u) * 8u));
                        dirty |= ob == cb;
                    }
                    uint32_t qpos = 0xFFFFu;
                    if (2 * d <= best0) { // queue admission: only nodes that can still be extracted
                        const uint32_t bk = bucketOf(d, Q.scaleU, Q.scaleB, NB - 1);
                        const uint32_t sh = (bk & 3u) * 8u;
                        const uint32_t pos = (atomicAdd(&cl[bk >> 2], 0x11u << sh) >> sh) & 0xFu;
                        dirty |= bk <= cb; // at or before the bucket held in registers (which may be ahead of this node's)
                        SMS_CHK(bk >= lo);
                        if (pos >= CAP) { // the carry of a full nibble corrupts the counters of a source that is given up anyway
                            ovf = true;
                        } else {
                            qpos = bk * C + pos;
                            const unsigned long long dbits = (unsigned long long)__double_as_longlong(d);
                            __stcs(QP(qpos), make_uint4((uint32_t)dbits, (uint32_t)(dbits >> 32), 2 * nbr + sd, hNew));
#ifdef SMSGPU_PAIRS_PRED_ARRAY
                            __stcs(QP(qpos) + 1, make_uint4(nrow.x, nrow.y, 0u, 0u));
#else
                            // the predecessor rides in the queue entry (never overwritten: tombstones touch its first 8 bytes
                            // only), which is a full-sector store anyway; a 4-byte store into the predecessor array costs the
                            // memory system a sector read and a sector write
                            __stcs(QP(qpos) + 1, make_uint4(nrow.x, nrow.y, cur, 0u));
#endif
                        }
                    }
                    if (k)
                        pB = (hNew & 0xFFFFu) | (qpos << 16);
                    else
                        pA = (hNew & 0xFFFFu) | (qpos << 16);
#ifdef SMSGPU_PAIRS_PRED_ARRAY
                    __stcs(&PQP(2 * slot + sd)->x, cur); // predecessor
#else
                    if (qpos == 0xFFFFu) __stcs(&PQP(2 * slot + sd)->x, cur); // predecessor of a label that has no queue entry
#endif
                }
                Entry e;
                e.d0 = side ? dB : dA;
                e.d1 = side ? dA : dB;
                e.h0 = side ? pB : pA;
                e.h1 = side ? pA : pB;
                e.key = nbr;
                e.flags = stamp;
                storeEntryCS(TABP(slot), e);
            }
            if ((__ballot_sync(FULL, ovf) >> gbase) & 0xFFFFu) {
                if (tl == 0) {
                    atomicAdd(&R.counters[13], 1ull); // a bucket is full
                    atomicAdd(&R.counters[15], (unsigned long long)lo); // diagnostics: where in the queue (bucket being settled)
                }
                overflow = 1;
                st = 2;
                run = false;
            }
            if ((__ballot_sync(FULL, dirty) >> gbase) & 0xFFFFu) needLoad = true;
            __syncwarp(FULL); // keys, labels and counters of this pass are visible to the next extraction
            PCLK(7);
        }
#ifdef SMSGPU_PAIRS_CLOCKS
        piters++;
#endif
        if (run && liveRegs == 0) needLoad = true;
      } // stepping loop
        if (done) break;

        // ---- (1) finish ended sources: state to shared memory, then the whole warp works for one half at a time ---------
        if (tl == 0) {
            svU[0] = st == 2 ? 0u : (uint32_t)st; // an ended half is idle afterwards
            svU[1] = src;
            svU[2] = wi;
            svU[3] = (uint32_t)__double2loint(best);
            svU[4] = (uint32_t)__double2hiint(best);
            svU[5] = bestHop;
            svU[6] = histN;
            svU[7] = touchedN;
            svU[8] = arcsN;
            svU[9] = settledN;
            svU[10] = (uint32_t)overflow;
            svU[11] = cb;
            svU[12] = lo;
            svU[13] = liveRegs;
            svU[14] = needLoad ? 1u : 0u;
            svU[15] = pend | (tie ? 4u : 0u);
        }
        svK[0] = (uint32_t)khi;
        svK[32] = (uint32_t)(khi >> 32);
        svK[64] = (uint32_t)klo;
        svK[96] = (uint32_t)(klo >> 32);
        svK[128] = rowB;
        svK[160] = deg;
        __syncwarp(FULL);
#pragma unroll 1
        for (int gg = 0; gg < 2; ++gg) {
            if (!(finb & (1u << (gg * 16)))) continue; // warp-uniform
            uint32_t *u = smemw + (uint32_t)(blockDim.x >> 5) * 2 * hbw + (uint32_t)warpInCta * kPairsSaveWords + gg * 16;
            const uint32_t fwi = u[2];
            const int32_t fidx = P.worklist ? P.worklist[fwi] : (int32_t)fwi;
            const uint32_t fpend = u[15] & 3u;
            // the twin's results come from the same labels iff this search never broke a tie by id (and did not overflow)
            const int32_t fidx2 = fpend == 2u ? (P.worklist ? P.worklist[fwi + 1] : (int32_t)(fwi + 1)) : -1;
            const int mirrorOk = (u[15] >> 2) == 0u;
            const int ok = finishSourcePairs(Q.kargs, scratch + (((size_t)blockIdx.x * (blockDim.x >> 5) + warpInCta) * 2 + gg) * tileBytes(caps),
                                             reinterpret_cast<unsigned char *>(smemw + (uint32_t)(warpInCta * 2 + gg) * hbw),
                                             __hiloint2double((int)u[4], (int)u[3]), u[6], u[7], u[8], u[9], (int)u[10], u[1], fidx, fidx2, mirrorOk);
            __syncwarp(FULL);
            if (lane == 0) u[15] = (fpend != 0u && ok != 2) ? 1u : 0u; // 1: the second source is searched next
        }
        __syncwarp(FULL);
    }
#undef TABP
#undef PQP
#undef QP
#undef HISTP
}

} // namespace smsgpu
