// "Lanes" engine of K2 (tier 1, in front of the pairs engine on long source lists): ONE SOURCE PER THREAD.
//
// Same search as every other engine -- runDijsktra of the reference (src/graph/ocw_seperator.cpp:37-112) in the
// canonical order (dist, hops, doubled id), event for event -- written as plain sequential code: the statement of the
// reference, with the reference's dense arrays replaced by a hashed label table and its heap by a monotone bucket queue.
// Why it was built: the warp-cooperative engines spend ~400 warp instructions per settled node on ballots, shuffles and
// predication to keep 16 lanes busy with the ~9 arcs of a node, and have 7 000 searches in flight
// (profiles/r02_pairs_tier1_random1m_256k.md).  A thread per source needs no cross-lane instruction and puts up to 100 000
// searches in flight.
// What was measured (config 5, B200, profiles/r02_lanes_experiments.md): it is SLOWER -- 824 ms per round at 98 304
// searches in flight against 475 ms for the pairs engine -- so it is OFF by default (SMSGPU_LANES=1 turns it on; the parity
// tests run it).  Without shared memory there is no presence bitmap and no bucket counter on chip: every probe, every
// bucket word is a DRAM access, 3.6 KB of DRAM traffic per settled node against 1.0 KB (every isolated access costs a
// 128-byte line, scripts/sector_flavors.cu), and the 32 lanes of a warp take 32 different paths (6.7 of 32 threads active
// per instruction).  More searches in flight did not help beyond ~65 000 (profiles/r02_lanes_search_v3.md).
//
// Working memory of one search (a "tile", thread private, in HBM; bytes: lanesTileBytes -- what limits the searches in
// flight is memory CAPACITY, so the tile is compact):
//   tab   Entry[tabCap]        label table, open addressing; a slot is occupied iff its `flags` word carries the stamp of
//                              the search (nothing is cleared between searches);
//                              hop words = hops (12 bits) | queue entry << 12 (13 bits) | position in its bucket << 25
//   q     32 B x qCap          queue entries {dist, id, hops | first arc, degree, PREDECESSOR, -}, allocated in push order,
//                              written once with two 16-byte stores (a full sector: no read-modify-write in the memory
//                              system), never moved -- the entry a label points to also holds its predecessor, so there is
//                              no predecessor array.  Labels that may not enter a bucket (2 d > bestDist) get an entry too.
//   bk    32 B x NB            one sector per bucket: word 0 = entries appended << 16 | dead mask (bit i: position i
//                              extracted / superseded), then 14 half-words: the queue entries of its positions
//   hist  uint32[histCap]      intermediateVertices_ in push order
// This header is the engine's sequential core and compiles for the host as well (tests/cpp/lanes_host.cpp runs it on
// the CPU against the oracle); the kernels around it are in ocw_search_lanes.cuh.
#pragma once
#include "ocw_types.cuh"
#include <cstring>

namespace smsgpu {

constexpr uint32_t kLanesC = 14;               // queue positions per bucket
constexpr uint32_t kLanesHopBits = 12, kLanesHopMask = 0xFFFu;
constexpr uint32_t kLanesQBits = 13, kLanesQMask = 0x1FFFu; // queue entry of a label: (hop word >> 12) & mask
constexpr uint32_t kLanesNoPos = 15;           // hop word >> 25: the label's entry sits in no bucket
constexpr uint32_t kLanesPackedNone = 0xFFFFFFFFu;

struct LanesCfg {
    uint32_t tabCap, tabShift;  // label table slots (power of two), 32 - log2(tabCap)
    uint32_t touchedCap;        // labels a search may create
    uint32_t histCap;
    uint32_t nbuckets;          // NB (multiple of 4)
    uint32_t qCap;              // queue entries (<= 8192)
    float scaleU, scaleB;       // bucket map of the pairs engine (ocw_search_pairs.cuh: bucketOf)
    uint32_t rowShift;          // Arc::row = first arc | degree << rowShift
};

// result record of one tile's search, read by the finishing warp
struct __align__(16) LanesRec {
    uint32_t src;       // doubled id of the source searched
    int32_t srcIdx;     // local source index; < 0: the tile has nothing to finish
    int32_t twinIdx;    // second source of the work item (local index), or -1
    uint32_t flags;     // bit 0: a capacity was hit (the source goes to the next engine); bit 1: an extraction broke a
                        // (dist, hops) tie by id (the twin is not the mirror image); bit 2: the second source is the twin
    double best;
    uint32_t histN, touchedN, arcs, settled;
    uint32_t stamp, qN;
    uint32_t pad[4];
};
static_assert(sizeof(LanesRec) == 64, "LanesRec is two sectors");

struct LanesTile {
    Entry *tab;
    uint4 *q;       // two uint4 per entry
    uint4 *bk;      // two uint4 per bucket
    uint32_t *hist;
    LanesRec *rec;
};

__host__ __device__ inline size_t lanesTileBytes(const LanesCfg &c) {
    size_t b = (size_t)c.tabCap * sizeof(Entry);
    b += (size_t)c.qCap * 32;
    b += (size_t)c.nbuckets * 32;
    b += (size_t)c.histCap * sizeof(uint32_t);
    b = (b + 31) & ~(size_t)31;
    b += sizeof(LanesRec);
    return (b + 255) & ~(size_t)255;
}
__host__ __device__ inline LanesTile lanesCarve(unsigned char *base, const LanesCfg &c) {
    LanesTile T;
    unsigned char *p = base;
    T.tab = reinterpret_cast<Entry *>(p);
    p += (size_t)c.tabCap * sizeof(Entry);
    T.q = reinterpret_cast<uint4 *>(p);
    p += (size_t)c.qCap * 32;
    T.bk = reinterpret_cast<uint4 *>(p);
    p += (size_t)c.nbuckets * 32;
    T.hist = reinterpret_cast<uint32_t *>(p);
    p += (size_t)c.histCap * sizeof(uint32_t);
    p = base + (((size_t)(p - base) + 31) & ~(size_t)31);
    T.rec = reinterpret_cast<LanesRec *>(p);
    return T;
}

// ---- memory access: streaming hints on the thread-private working memory (it is touched once or twice and is far larger
// than L2, the graph is shared by all searches), read-only path for the graph ----------------------------------------
__host__ __device__ __forceinline__ uint4 lnLoad4(const uint4 *p) {
#ifdef __CUDA_ARCH__
    return __ldcs(p);
#else
    return *p;
#endif
}
__host__ __device__ __forceinline__ void lnStore4(uint4 *p, const uint4 v) {
#ifdef __CUDA_ARCH__
    __stcs(p, v);
#else
    *p = v;
#endif
}
__host__ __device__ __forceinline__ uint4 lnLoadRO4(const uint4 *p) {
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}
__host__ __device__ __forceinline__ double lnBitsToDouble(const uint32_t lo, const uint32_t hi) {
#ifdef __CUDA_ARCH__
    return __hiloint2double((int)hi, (int)lo);
#else
    const unsigned long long b = ((unsigned long long)hi << 32) | lo;
    double d;
    memcpy(&d, &b, 8);
    return d;
#endif
}
__host__ __device__ __forceinline__ unsigned long long lnDoubleBits(const double d) {
#ifdef __CUDA_ARCH__
    return (unsigned long long)__double_as_longlong(d);
#else
    unsigned long long b;
    memcpy(&b, &d, 8);
    return b;
#endif
}
// Monotone (non-decreasing in d >= 0) bucket index: the pairs engine's map (ocw_search_pairs.cuh: bucketOf).  Only
// monotonicity matters for the result; host and device need not agree bit for bit.
__host__ __device__ __forceinline__ uint32_t lnBucketOf(const double d, const float scaleU, const float scaleB, const uint32_t nbm1) {
#ifdef __CUDA_ARCH__
    const float f = __double2float_rn(d);
    uint32_t b;
    if (scaleU > 0.f) {
        const float y = __int_as_float(__float2int_rz(f * scaleU * 8388608.0f) + 0x3F800000);
        b = __float2uint_rz((y - 1.0f) * scaleB);
    } else {
        b = __float2uint_rz(f * scaleB);
    }
    return min(b, nbm1);
#else
    const float f = (float)d;
    uint32_t b;
    if (scaleU > 0.f) {
        const int bits = (int)(f * scaleU * 8388608.0f) + 0x3F800000;
        float y;
        memcpy(&y, &bits, 4);
        b = (uint32_t)((y - 1.0f) * scaleB);
    } else {
        b = (uint32_t)(f * scaleB);
    }
    return b < nbm1 ? b : nbm1;
#endif
}

struct LanesEntryRegs { // one label-table entry in registers
    uint4 a, b;         // a = {d0 lo, d0 hi, d1 lo, d1 hi}, b = {key, h0, h1, flags}
};
__host__ __device__ __forceinline__ LanesEntryRegs lnLoadEntry(const Entry *p) {
    LanesEntryRegs e;
    e.a = lnLoad4(reinterpret_cast<const uint4 *>(p));
    e.b = lnLoad4(reinterpret_cast<const uint4 *>(p) + 1);
    return e;
}
// The bucket sectors are read and updated at L2 only (ld.cg / st.cg / atomics): an atomic does not update a line in L1.
__host__ __device__ __forceinline__ uint4 lnLoadBk(const uint4 *p) {
#ifdef __CUDA_ARCH__
    return __ldcg(p);
#else
    return *p;
#endif
}
__host__ __device__ __forceinline__ void lnStoreBk(uint4 *p, const uint4 v) {
#ifdef __CUDA_ARCH__
    __stcg(p, v);
#else
    *p = v;
#endif
}
__host__ __device__ __forceinline__ void lnStoreBkHalf(unsigned short *p, const unsigned short v) {
#ifdef __CUDA_ARCH__
    __stcg(p, v);
#else
    *p = v;
#endif
}
__host__ __device__ __forceinline__ uint32_t lnCntAdd(uint32_t *p, const uint32_t v) { // returns the old word
#ifdef __CUDA_ARCH__
    return atomicAdd(p, v);
#else
    const uint32_t o = *p;
    *p = o + v;
    return o;
#endif
}
__host__ __device__ __forceinline__ void lnCntOr(uint32_t *p, const uint32_t v) { // no result: a reduction, nothing to wait for
#ifdef __CUDA_ARCH__
    atomicOr(p, v);
#else
    *p |= v;
#endif
}

#ifndef SMSGPU_LANES_GROUP
#define SMSGPU_LANES_GROUP 4
#endif
constexpr int kLanesG = SMSGPU_LANES_GROUP; // arcs per group: their loads, probes and queue-position atomics travel together
                                            // (measured: 4 beats 8 -- 163 against 232 registers -- and 2)

// The search of one source.  `srow` = {first arc, degree} of the source's vertex.  Fills `rec` (src, flags bits 0-1,
// best, histN, touchedN, arcs, settled, stamp, qN); srcIdx / twinIdx / flag bit 2 belong to the caller.
//
// A thread is a chain of dependent memory round trips, so the arcs of a settled node are handled in GROUPS of kLanesG:
//   (1) the group's arc records, one round trip; (2) the label entries at their home slots, one round trip (further probes
//   of a collision chain on demand); (3) the reference's sequential logic over the group, in registers: accept tests, running
//   bestDist / history (:59-84), and for every accepted label its queue entry (stored at once: its place is the push
//   counter), the ISSUE of the atomic that appends it to its bucket (the result is not looked at yet) and of the reduction
//   that kills a superseded key; (4) one wait for the atomics, then the bucket's index half-words and the labels.
__host__ __device__ inline void lanesSearch(const Arc *__restrict__ arcs, const uint2 *__restrict__ rowinfo, const LanesCfg &C,
                                            const LanesTile &T, const uint32_t src, const uint2 srow, const uint32_t stamp,
                                            const double bestInit, LanesRec &rec) {
    const uint32_t NB = C.nbuckets, tabMask = C.tabCap - 1u;
    // ---- reset the bucket sectors, label and enqueue the source (:54-57) -------------------------------------------
    {
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        for (uint32_t i = 2; i < 2 * NB; ++i) lnStoreBk(T.bk + i, z);
        lnStoreBk(T.bk, make_uint4(1u << 16, 0u, 0u, 0u)); // bucket 0: one entry appended, position 0 = queue entry 0
        lnStoreBk(T.bk + 1, z);
        const uint32_t s = ((src >> 1) * 0x9E3779B1u) >> C.tabShift;
        const unsigned long long un = lnDoubleBits(kUnlabelled);
        uint4 a = make_uint4((uint32_t)un, (uint32_t)(un >> 32), (uint32_t)un, (uint32_t)(un >> 32));
        uint4 b = make_uint4(src >> 1, kLanesPackedNone, kLanesPackedNone, stamp);
        if (src & 1u) {
            a.z = 0u;
            a.w = 0u;
            b.z = 0u; // 0 hops, queue entry 0, position 0
        } else {
            a.x = 0u;
            a.y = 0u;
            b.y = 0u;
        }
        lnStore4(reinterpret_cast<uint4 *>(T.tab + s), a);
        lnStore4(reinterpret_cast<uint4 *>(T.tab + s) + 1, b);
        lnStore4(T.q, make_uint4(0u, 0u, src, 0u));             // dist 0, id src, 0 hops
        lnStore4(T.q + 1, make_uint4(srow.x, srow.y, src, 0u)); // row, predecessor: itself (never followed)
    }
    double best = bestInit;
    uint32_t bestHop = 0; // bestHopDist: a double in the reference (:52) that only ever holds small integers
    uint32_t histN = 0, touchedN = 1, arcsN = 0, settledN = 0, qN = 1;
    bool overflow = false, tie = false;
    uint32_t lo = 0; // bucket of the last extraction: keys never decrease

    while (!overflow) {
        // ---- first bucket at or after `lo` with a live key: four bucket sectors per round trip ------------------------
        uint32_t b = NB, live = 0;
        uint32_t ix[7] = {0, 0, 0, 0, 0, 0, 0}; // its 14 index half-words
        for (uint32_t b0 = lo; b0 < NB && b >= NB; b0 += 4) {
            uint4 ra[4], rb[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                ra[j] = make_uint4(0u, 0u, 0u, 0u);
                rb[j] = make_uint4(0u, 0u, 0u, 0u);
                if (b0 + j < NB) {
                    ra[j] = lnLoadBk(T.bk + 2 * (size_t)(b0 + j));
                    rb[j] = lnLoadBk(T.bk + 2 * (size_t)(b0 + j) + 1);
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t ap = ra[j].x >> 16, lv = ((1u << ap) - 1u) & ~ra[j].x;
                if (b >= NB && lv) {
                    b = b0 + j;
                    live = lv;
                    ix[0] = ra[j].y;
                    ix[1] = ra[j].z;
                    ix[2] = ra[j].w;
                    ix[3] = rb[j].x;
                    ix[4] = rb[j].y;
                    ix[5] = rb[j].z;
                    ix[6] = rb[j].w;
                }
            }
        }
        if (b >= NB) break; // queue empty (:88)
        // ---- its minimum live key, (dist, hops, id) lexicographic; four entries per round trip ---------------------
        unsigned long long mhi = ~0ull, mlo = ~0ull;
        uint32_t mpos = 0, mfirst = 0, mdeg = 0;
        bool mtie = false;
#pragma unroll
        for (int p0 = 0; p0 < 16; p0 += 4) {
            if (!((live >> p0) & 0xFu)) continue;
            uint4 k[4], r[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                k[j] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
                r[j] = make_uint4(0u, 0u, 0u, 0u);
                if (p0 + j < (int)kLanesC && ((live >> (p0 + j)) & 1u)) {
                    const uint32_t n = (ix[(p0 + j) >> 1] >> (((p0 + j) & 1) * 16)) & 0xFFFFu;
                    k[j] = lnLoad4(T.q + 2 * (size_t)n);
                    r[j] = lnLoad4(T.q + 2 * (size_t)n + 1);
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (p0 + j >= (int)kLanesC || !((live >> (p0 + j)) & 1u)) continue;
                const unsigned long long hi = ((unsigned long long)k[j].y << 32) | k[j].x;
                const unsigned long long lw = ((unsigned long long)k[j].w << 32) | k[j].z; // hops << 32 | id
                if (hi == mhi && (uint32_t)(lw >> 32) == (uint32_t)(mlo >> 32)) mtie = true; // equal (dist, hops): the id decides
                if (hi < mhi || (hi == mhi && lw < mlo)) {
                    if (hi != mhi || (uint32_t)(lw >> 32) != (uint32_t)(mlo >> 32)) mtie = false;
                    mhi = hi;
                    mlo = lw;
                    mpos = (uint32_t)(p0 + j);
                    mfirst = r[j].x;
                    mdeg = r[j].y;
                }
            }
        }
        if (mtie) tie = true; // -> the twin's search is not the mirror image of this one
        const uint32_t cur = (uint32_t)mlo, hc = (uint32_t)(mlo >> 32);
        const double dc = lnBitsToDouble((uint32_t)mhi, (uint32_t)(mhi >> 32));
        // loop condition of ocw_seperator.cpp:88-90
        if (cur == (src ^ 1u) || 2 * dc > best) break;
        lnCntOr(reinterpret_cast<uint32_t *>(T.bk + 2 * (size_t)b), 1u << mpos); // extract_top: the position is dead from now on
        settledN++;
        lo = b;
        const uint32_t side = cur & 1u, hNew = hc + 1u;
        if (hNew >= kLanesHopMask) { // hops share a word with the queue entry
            overflow = true;
            break;
        }

        // ---- relax the arcs of the settled node in adjacency order, kLanesG per group -------------------------------
        for (uint32_t a0 = 0; a0 < mdeg && !overflow; a0 += kLanesG) {
            const uint32_t nb = (mdeg - a0) < (uint32_t)kLanesG ? (mdeg - a0) : (uint32_t)kLanesG;
            uint4 av[kLanesG];
            LanesEntryRegs pe[kLanesG];
            // what phase (4) needs of every arc
            uint32_t slotOf[kLanesG];  // its label's slot, 0xFFFFFFFF: nothing to store
            uint4 labA[kLanesG];       // the new distances {d0, d1} as stored
            uint32_t hopA[kLanesG], hopB[kLanesG]; // new hop words of the stay side / the other side (bucket position pending)
            uint32_t retA[kLanesG], retB[kLanesG]; // old bucket word returned by the append
            uint32_t bkA[kLanesG], bkB[kLanesG];   // the bucket appended to
            uint32_t how[kLanesG];     // bit 0 / 1: stay / cross side accepted; bit 2 / 3: appended to a bucket
            // (1) the arc records
#pragma unroll
            for (int j = 0; j < kLanesG; ++j) {
                av[j] = make_uint4(0x80000000u, 0u, 0u, 0u);
                if ((uint32_t)j < nb) av[j] = lnLoadRO4(reinterpret_cast<const uint4 *>(arcs + mfirst + a0 + j));
            }
            // (2) the labels at their home slots (speculative: nothing of this group has been stored yet)
#pragma unroll
            for (int j = 0; j < kLanesG; ++j) {
                pe[j].a = make_uint4(0u, 0u, 0u, 0u);
                pe[j].b = make_uint4(0u, 0u, 0u, 0u);
                slotOf[j] = 0xFFFFFFFFu;
                how[j] = 0u;
                if ((uint32_t)j < nb && !(av[j].x & 0x80000000u))
                    pe[j] = lnLoadEntry(T.tab + (((av[j].x & 0x7FFFFFFFu) * 0x9E3779B1u) >> C.tabShift));
            }
            // (3) the reference's sequential logic over the group
#pragma unroll
            for (int j = 0; j < kLanesG; ++j) {
                if ((uint32_t)j >= nb || overflow) continue;
                if (av[j].x & 0x80000000u) continue; // blacklisted neighbours are skipped (:98-99)
                const uint32_t nbr = av[j].x & 0x7FFFFFFFu;
                const double w = lnBitsToDouble(av[j].z, av[j].w);
                arcsN++;
                if (touchedN + 1 > C.touchedCap) {
                    overflow = true;
                    continue;
                }
                const double dS = dc + w;         // totalStayDist  (:104)
                const double dX = dc + (1.0 - w); // totalCrossDist (:107-108)
                // label pruning (header of ocw_search.cuh) against the running bestDist
                const bool creatS = dS + dc <= best, creatX = dX + dc <= best;
                // ---- label table: slots claimed by earlier arcs of this group are not in memory yet ------------------
                uint32_t s = (nbr * 0x9E3779B1u) >> C.tabShift;
                LanesEntryRegs e = pe[j];
                bool found = false, isNew = false;
                while (true) {
                    bool claimed = false;
#pragma unroll
                    for (int jj = 0; jj < kLanesG; ++jj)
                        if (jj < j && slotOf[jj] == s) claimed = true; // by another vertex (a row holds a vertex once)
                    if (!claimed) {
                        if (e.b.w != stamp) { // free slot
                            isNew = creatS || creatX;
                            break;
                        }
                        if (e.b.x == nbr) {
                            found = true;
                            break;
                        }
                    }
                    s = (s + 1u) & tabMask;
                    e = lnLoadEntry(T.tab + s);
                }
                if (!found && !isNew) continue; // no label, and none may be created
                double dA = kUnlabelled, dB = kUnlabelled;             // label of the side the stay arc reaches / of the other side
                uint32_t pA = kLanesPackedNone, pB = kLanesPackedNone; // hop words
                if (found) {
                    const double e0 = lnBitsToDouble(e.a.x, e.a.y), e1 = lnBitsToDouble(e.a.z, e.a.w);
                    dA = side ? e1 : e0;
                    dB = side ? e0 : e1;
                    pA = side ? e.b.z : e.b.y;
                    pB = side ? e.b.y : e.b.z;
                }
                const bool hadA = dA < kUnlabelled, hadB = dB < kUnlabelled;
                uint32_t hA = pA & kLanesHopMask, hB = pB & kLanesHopMask;
                uint32_t hw = 0u;
                // the row of the neighbour, for its queue entries
                uint32_t first = av[j].y & ((1u << C.rowShift) - 1u), deg = av[j].y >> C.rowShift;
#pragma unroll
                for (int k = 0; k < 2; ++k) { // updatePaths(stayNeighbor, ...) then updatePaths(crossNeighbor, ...)  (:59-84)
                    const double d = k ? dX : dS;
                    const bool had = k ? hadB : hadA;
                    const double dOld = k ? dB : dA, dTw = k ? dA : dB;
                    const uint32_t hOld = k ? hB : hA, hTw = k ? hA : hB;
                    bool acc;
                    if (!had)
                        acc = d < kInfDist && ((k ? creatX : creatS) || d + dTw <= best);
                    else
                        acc = (dOld == d) ? (hNew < hOld) : (d < dOld);
                    if (!acc || overflow) continue;
                    if (qN >= C.qCap) {
                        overflow = true;
                        continue;
                    }
                    hw |= 1u << k;
                    const uint32_t sd = k ? (side ^ 1u) : side;
                    // a superseded key leaves the queue (bucket from its distance, position from the hop word): a reduction, no wait
                    const uint32_t oldPos = (k ? pB : pA) >> 25;
                    if (had && oldPos < kLanesC)
                        lnCntOr(reinterpret_cast<uint32_t *>(T.bk + 2 * (size_t)lnBucketOf(dOld, C.scaleU, C.scaleB, NB - 1u)), 1u << oldPos);
                    // the new queue entry; only nodes that can still be extracted are appended to a bucket
                    const uint32_t n = qN++;
                    const unsigned long long db = lnDoubleBits(d);
                    uint32_t ret = 0, bkt = 0;
                    if (2 * d <= best) {
                        if (deg == (0xFFFFFFFFu >> C.rowShift)) { // saturated degree field: read the row record
                            const uint2 ri = rowinfo[nbr];
                            first = ri.x;
                            deg = ri.y;
                        }
                        bkt = lnBucketOf(d, C.scaleU, C.scaleB, NB - 1u);
                        ret = lnCntAdd(reinterpret_cast<uint32_t *>(T.bk + 2 * (size_t)bkt), 1u << 16); // looked at in (4)
                        hw |= 4u << k;
                    }
                    lnStore4(T.q + 2 * (size_t)n, make_uint4((uint32_t)db, (uint32_t)(db >> 32), 2u * nbr + sd, hNew));
                    lnStore4(T.q + 2 * (size_t)n + 1, make_uint4(first, deg, cur, 0u));
                    const uint32_t pw = hNew | (n << kLanesHopBits) | (kLanesNoPos << 25);
                    if (k) {
                        dB = d;
                        hB = hNew;
                        pB = pw;
                        retB[j] = ret;
                        bkB[j] = bkt;
                    } else {
                        dA = d;
                        hA = hNew;
                        pA = pw;
                        retA[j] = ret;
                        bkA[j] = bkt;
                    }
                    // record logic (:71-82)
                    const double D = d + dTw;
                    const uint32_t H = hNew + hTw;
                    bool push = false;
                    if (D < best || (D == best && H < bestHop)) {
                        best = D;
                        bestHop = H;
                        push = true;
                    } else if (D == best && H < 2u * bestHop) {
                        push = true;
                    }
                    if (push) {
                        if (histN >= C.histCap) {
                            overflow = true;
                            continue;
                        }
                        T.hist[histN++] = 2u * nbr + sd;
                    }
                }
                if (hw) {
                    const unsigned long long a0b = lnDoubleBits(side ? dB : dA), a1b = lnDoubleBits(side ? dA : dB);
                    labA[j] = make_uint4((uint32_t)a0b, (uint32_t)(a0b >> 32), (uint32_t)a1b, (uint32_t)(a1b >> 32));
                    hopA[j] = pA;
                    hopB[j] = pB;
                    how[j] = hw;
                    slotOf[j] = s;
                    if (isNew) touchedN++;
                }
            }
            // (4) the appends have answered: index half-words of the buckets, labels
#pragma unroll
            for (int j = 0; j < kLanesG; ++j) {
                if (slotOf[j] == 0xFFFFFFFFu) continue;
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    if (!((how[j] >> (2 + k)) & 1u)) continue;
                    const uint32_t pos = (k ? retB[j] : retA[j]) >> 16;
                    if (pos >= kLanesC) {
                        overflow = true; // the bucket is full
                        continue;
                    }
                    const uint32_t pw = k ? hopB[j] : hopA[j];
                    lnStoreBkHalf(reinterpret_cast<unsigned short *>(T.bk + 2 * (size_t)(k ? bkB[j] : bkA[j])) + 2 + pos,
                                  (unsigned short)((pw >> kLanesHopBits) & kLanesQMask));
                    const uint32_t nw = (pw & ~(15u << 25)) | (pos << 25);
                    if (k)
                        hopB[j] = nw;
                    else
                        hopA[j] = nw;
                }
                lnStore4(reinterpret_cast<uint4 *>(T.tab + slotOf[j]), labA[j]);
                lnStore4(reinterpret_cast<uint4 *>(T.tab + slotOf[j]) + 1,
                         make_uint4(av[j].x & 0x7FFFFFFFu, side ? hopB[j] : hopA[j], side ? hopA[j] : hopB[j], stamp));
            }
        }
    }
    rec.src = src;
    rec.flags = (overflow ? 1u : 0u) | (tie ? 2u : 0u);
    rec.best = best;
    rec.histN = histN;
    rec.touchedN = touchedN;
    rec.arcs = arcsN;
    rec.settled = settledN;
    rec.stamp = stamp;
    rec.qN = qN;
}

} // namespace smsgpu
