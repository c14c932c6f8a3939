// Shared device/host types of the odd-cycle separation engine (libsmsgpu.so).
// Product code: nothing under oracle/ is included or linked here.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

// Debug build (-DSMSGPU_DEBUG_CHECKS, sms_b200/build.py --debug): device-side bounds / invariant checks that stop the
// kernel with file:line.  compute-sanitizer is not available on the GPU pool, so this is the memory checker.
#ifdef SMSGPU_DEBUG_CHECKS
#undef NDEBUG
#include <cassert>
#define SMS_CHK(cond) assert(cond)
#else
#define SMS_CHK(cond) ((void)0)
#endif

namespace smsgpu {

// One CSR adjacency entry in HBM, rebuilt every round by K1 (dc_weights): 16 B, one 128-bit load per arc scan.
//   nbr : neighbour's original vertex id, bit 31 set when the neighbour is blacklisted this round
//         (folds the `if (blacklist_[neighbor]) continue` of ocw_seperator.cpp:98-99 into the arc)
//   row : the NEIGHBOUR's adjacency row, packed: first arc | degree << rowShift (degree saturates at the field's
//         maximum, which then means "read the row record").  A relaxation that pushes the neighbour into the queue
//         needs its row for the queue entry; carried by the arc it costs no dependent load (pairs engine).
//   w   : stay weight = clamp(x*[eid], 0, 1) (ocw_seperator.cpp:25-35); the cross weight 1.0 - w is computed
// The NetworKit edge id of the arc is needed only when a walk is written out; it lives in GraphDev::eid.
struct __align__(16) Arc {
    int32_t nbr;
    uint32_t row;
    double w;
};

// Label-table entry: both double-cover copies (v,0) and (v,1) of one original vertex share one 32-byte
// entry (one HBM sector), so the twin test dist[t] + dist[t^1] of ocw_seperator.cpp:72-74 costs no probe.
struct __align__(16) Entry {
    double d0, d1;     // distances_ of side 0 / side 1; kUnlabelled when the side has no label
    uint32_t key;      // original vertex id, kEmptyKey when the slot is free
    uint32_t h0, h1;   // hopDistances_
    uint32_t flags;    // bit s: side s settled (extracted from the queue)
};
static_assert(sizeof(Entry) == 32, "Entry must be one 32-byte sector");

// Priority-queue entry; (hi, lo) compared as one 128-bit unsigned number = the canonical strict order
// (dist, hops, doubled id).  dist >= +0.0, so the IEEE bit pattern is order preserving.
struct __align__(16) HeapEnt {
    unsigned long long hi; // __double_as_longlong(dist)
    unsigned long long lo; // hops << 32 | doubled id
};

constexpr uint32_t kEmptyKey = 0xFFFFFFFFu;
constexpr double kUnlabelled = 4.0;       // > any reachable distance (< 1.5); never compares as an improvement
constexpr uint32_t kUnlabelledHops = 0x3FFFFFFFu;
constexpr double kInfDist = 1.0 + 10e-4;  // kInfDist_, include/sms/graph/ocw_seperator.hpp:45: offers >= this are never labelled

// std::min(std::max(0., w), 1.) exactly as the reference evaluates it (ocw_seperator.cpp:28,34)
__host__ __device__ __forceinline__ double clamp01(double w) {
    const double a = (0. < w) ? w : 0.; // std::max(0., w)
    return (1. < a) ? 1. : a;           // std::min(a, 1.)
}

struct GraphDev {
    int64_t n, m;
    const int64_t *rowptr; // n+1
    const Arc *arcs;       // 2m, adjacency order of the reference graph
    const int32_t *eid;    // 2m, NetworKit edge id of every adjacency entry (walk output only)
    // SMSGPU_MODE_VERBATIM only (all three null in independent-source mode): the reference's blacklist as a function of
    // time.  stamp[v] = position in the source list of the FIRST source that blacklists v (-1: blacklisted from the
    // start, kNeverStamp: never); the source at position `now` sees v blacklisted iff stamp[v] < now.
    const int32_t *blStamp; // [n] stamps of the previous sweep (read)
    int32_t *blStampNext;   // [n] stamps this sweep produces (atomicMin)
    uint8_t *srcEmpty;      // [nsrc] 1: the source ran and found no intermediate (its stamp floods integral edges)
};
constexpr int32_t kNeverStamp = 0x7FFFFFFF;
// `if (blacklist_[neighbor]) continue` (ocw_seperator.cpp:98-99) / `if (blacklist_[currentSource_])` (:334) in mode V
__device__ __forceinline__ bool stampedBefore(const GraphDev &G, const uint32_t v, const int32_t now) {
    return G.blStamp[v] < now;
}

// Capacities of one tile's private working memory (shared memory in tier 0, HBM in tiers 1 and 2).
struct TileCaps {
    uint32_t tabCap;     // label-table slots (power of two unless dense)
    uint32_t tabShift;   // 32 - log2(tabCap)
    uint32_t dense;      // 1: slot = vertex id, no hashing (tier 2)
    uint32_t touchedCap; // max entries a source may create (<= tabCap / 2 unless dense)
    uint32_t heapCap;    // queue entries
    uint32_t histCap;    // intermediate-vertex history entries
    uint32_t walkCap;    // max edges of one path / walk
    uint32_t chordCap;   // max good chords of one cycle
    // "fast" engine (tiers 1 and 2, ocw_search_fast.cuh): monotone bucket queue instead of the heap and a
    // presence bitmap that arbitrates the label table.  nbuckets == 0 selects the heap engine.
    uint32_t nbuckets;   // buckets over [0, bestInit / 2]; heapCap == nbuckets * bucketCap
    uint32_t bucketCap;  // entries per bucket (<= 32: one coalesced load per pop)
    uint32_t pairs;      // pairs engine (ocw_search_pairs.cuh): the queue is an array of 32-byte entries {key, row}
};

// bytes of one tile's working memory
__host__ __device__ inline size_t tileBytes(const TileCaps &c) {
    size_t b = 0;
    b += (size_t)c.tabCap * sizeof(Entry);         // table
    b += (size_t)c.tabCap * 2 * sizeof(uint32_t) * (c.nbuckets ? 2 : 1); // pred (+ queue position) per (slot, side)
    b += (size_t)c.touchedCap * sizeof(uint32_t);  // touched slots
    b += (size_t)c.heapCap * sizeof(HeapEnt);      // queue (re-used as sort scratch after the search)
    b += (size_t)c.histCap * sizeof(uint32_t);     // history of intermediates
    // walk scratch: pa, pb (walkCap+2 each), wn (2*walkCap+2), ws (2*walkCap+2), warc (2*walkCap+2), crs (2*walkCap+2)
    b += (size_t)(2 * (c.walkCap + 2) + 4 * (2 * c.walkCap + 2)) * sizeof(int32_t);
    b += (size_t)(2 * c.walkCap + 2) * sizeof(double); // plen
    b += (size_t)c.chordCap * 4 * sizeof(int32_t);     // chords
    b += (size_t)c.chordCap * 0 + (size_t)(4 * c.walkCap + 8) * sizeof(int32_t); // covered list
    b = (b + 31) & ~(size_t)31;
    if (c.pairs) b += (size_t)c.heapCap * 32; // pairs engine: queue entries {dist, hops|id, first arc, degree}, one sector each
    return (b + 31) & ~(size_t)31;
}

// bytes of the fast engine's hot words of one tile: bucket counters (2 bytes per bucket) + table presence bitmap.
// Shared memory in tier 1, HBM (L2 resident) in tier 2.
__host__ __device__ inline size_t hotBytes(const TileCaps &c) {
    if (!c.nbuckets) return 0;
    return (((size_t)c.nbuckets * (c.pairs ? 1 : 2) + (size_t)c.tabCap / 8) + 31) & ~(size_t)31;
}

// Raw (pre-dedup) output arena, filled by the search kernel with atomic bump allocation.
struct RawOut {
    int32_t *node, *eid;   // [nodeCap]
    uint8_t *stay;         // [nodeCap]
    int32_t *cutOff;       // [cutCap]
    int32_t *cutLen;       // [cutCap]
    int32_t *cutSrc;       // [cutCap] local source index
    int32_t *cutSeq;       // [cutCap] emission index within the source
    int8_t *cutTier;       // [cutCap] tier that emitted it (valid iff == srcTier[src])
    unsigned long long *counters; // [0] cuts, [1] nodes, [2] arcs scanned, [3] settled, [4] sources run, [5] labels
    long long cutCap, nodeCap;
    // label export (parity tests)
    int32_t *labSrc, *labNode, *labHops;
    double *labDist;
    long long labCap;
    // per-source
    int8_t *srcTier;       // [nsrc] tier the source finished in, -1 = not finished / skipped
    long long *srcArcs, *srcSettled; // nullable
    uint32_t *srcWork;     // [nsrc] settled nodes of the source (always written): next round's longest-first order
};

struct SearchParams {
    const int32_t *sources; // doubled ids; nullptr -> ((i >> 1) * srcStride + srcBegin) * 2 + (i & 1)  (pairs dealt cyclically)
    long long nsrc;         // local number of sources
    long long srcBegin, srcStride;
    const int32_t *worklist; // nullptr in the first tier, else local source indices that overflowed
    long long nwork;
    unsigned long long *queueCounter; // work-stealing cursor
    int32_t *ovfList;       // overflowed local source indices
    unsigned long long *ovfCount;
    const uint8_t *blacklist; // [n] the round's initial blacklist (source skip test, ocw_seperator.cpp:334)
    double bestInit;
    int32_t maxInter;
    int32_t tier;
    int32_t exportLabels;
    int32_t debugFlags;     // bit 0: only the first tile of every warp works (debugging aid)
    uint32_t stampBase;     // fast engine: entry stamp of work item wi is stampBase + wi + 1 (unique per scratch)
    int32_t hotInSmem;      // fast engine: bucket counters + presence bitmap in shared memory (tier 1) or HBM (tier 2)
    int32_t runSlot;        // RawOut::counters slot that counts the sources run (4; another one for an overlapped launch)
};

} // namespace smsgpu
