// libsmsgpu.so -- C ABI (include/smsgpu.h) over the sm_100a kernels.  Host side of the separation round:
// uploads the CSR once (the ConshdlrOscw constructor's job in the reference, conshdlr_oscw.cpp:33-45 ->
// OcwSeparator::OcwSeparator), ships x* / blacklist per round, runs K1 -> K2-K4 (tier 0, 1, 2) -> K5 on one
// stream and returns the cut list in pinned host memory.
// Product code: no oracle, no CPU fallback -- every entry point fails with SMSGPU_ERR_CUDA without a device.
#include "../../include/smsgpu.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "dedup_kernels.cuh"
#include "ocw_kernels.cuh"
#include "small_cycles.cuh"

#ifndef SMSGPU_PAIRS_BUCKETS
#define SMSGPU_PAIRS_BUCKETS 2048
#endif
#ifndef SMSGPU_TIER1_BUCKETS
#define SMSGPU_TIER1_BUCKETS 4096
#endif

using namespace smsgpu;

namespace {

thread_local std::string g_createError;

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            want = bytes;
            e = cudaMalloc(&p, want);
        }
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T>
    T *as() const {
        return static_cast<T *>(p);
    }
};
struct PinBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        const size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
    template <class T>
    T *as() const {
        return static_cast<T *>(p);
    }
};

constexpr int kNumTiers = 4; // 0 shared memory, 1 HBM hashed (medium), 2 HBM hashed (large), 3 HBM dense
struct TierCfg {
    TileCaps caps{};
    long long ntiles = 0;   // tiles that own HBM working memory (tiers 1, 2)
    int ctas = 0;           // grid size
    size_t smemBytes = 0;   // dynamic shared memory per CTA (tier 0: whole tiles, tier 1: hot words)
    bool hotInSmem = false; // fast engine: where the bucket counters and the presence bitmap live
    bool pairs = false;     // pairs engine: two tiles per warp
    bool runs = false;      // runs engine (ocw_search_runs.cuh): one tile per CTA of kRunThreads threads
    uint64_t stampNext = 0; // fast engine: stamps handed out so far for this scratch (ocw_search_fast.cuh)
    DevBuf scratch;
    bool ready = false;
};

// host mirror of a cut list (pinned) + device copy
struct CutListDev {
    DevBuf off, node, eid, stay, ncross, src, seq, viol;
    long long ncuts = 0, nnodes = 0;
};
struct CutListHost {
    PinBuf off, node, eid, stay, ncross, src, viol;
    long long ncuts = 0, nnodes = 0;
};

} // namespace

struct smsgpu_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr; // the overlapped tier-2 launch of the sources known to be heavy
    cudaEvent_t ev[8]{};
    cudaEvent_t tierEv[2 * 5]{}; // start / end of each tier's search launch; [8], [9]: the overlapped tier-2 launch
    cudaEvent_t joinEv = nullptr, heavyEv = nullptr;
    int smCount = 148;
    long long n = 0, m = 0;
    std::string err;

    // graph (uploaded once)
    DevBuf dRowptr, dCol, dEid, dArcs;
    DevBuf dArcs12;           // optional 12-byte copy of the arc records (experiment build -DSMSGPU_PAIRS_ARCS12)
    DevBuf dPairsArgs;        // PairsKArgs of the launch in flight
    DevBuf dRowinfo;          // uint2[n] {first arc, degree}: pairs engine (tier 1)
    DevBuf dRowpack;          // uint32[2m] Arc::row of every adjacency entry (static; K1 copies it into the arc records)
    uint32_t rowShift = 31;   // Arc::row = first arc | degree << rowShift
    bool pairs = false;       // tier 1 runs the pairs engine (ocw_search_pairs.cuh)
    float pairsScaleU = 0.f;  // bucket map of the pairs engine (PairsParams)
    float pairsScaleB = 0.f;
    bool pairsMirror = true;  // twin sources searched once (SMSGPU_MIRROR=0 turns it off)
    // lanes engine (ocw_search_lanes.cuh): one source per THREAD, in front of the pairs engine on long source lists
    int lanesMode = 0;                    // SMSGPU_LANES: 0 off, 1 on for every list, 2 on for lists of >= lanesMinSources entries
    long long lanesMinSources = 200000;
    long long lanesMaxTiles = 40960;      // SMSGPU_LANES_TILES: tiles (searches in flight) over all sets
    int lanesSetsWanted = 2;              // SMSGPU_LANES_SETS: tile sets = streams that alternate batches
    int lanesSets = 0;
    int lanesTabLg = 0, lanesBuckets = SMSGPU_PAIRS_BUCKETS, lanesQCap = 8192; // SMSGPU_LANES_TABLG / _NB / _QCAP
    long long lanesTilesPerSet = 0, lanesAllocWant = 0;
    LanesCfg lanesCfg{};
    TileCaps lanesFcaps{};
    size_t lanesTileStride = 0, lanesWarpStride = 0;
    int lanesFinishCtas = 0;
    uint32_t lanesStamp = 0;
    DevBuf lanesWarpScratch, dLanesArgs, ovfD;
    void *lanesScratch = nullptr;
    size_t lanesScratchBytes = 0;
    cudaStream_t laneStream[4]{};
    cudaEvent_t laneEv[4]{}, lanesForkEv = nullptr, lanesT0 = nullptr, lanesT1 = nullptr;
    double lanesMs = 0.0;
    // round input
    DevBuf dX, dBl;
    const double *xPtr = nullptr;   // device pointer in use (own buffer or caller's)
    const uint8_t *blPtr = nullptr;
    PinBuf hX, hBl;
    bool haveInput = false, arcsDirty = true;
    // sources
    DevBuf dSources;
    bool explicitSources = false;
    long long nsrcGlobal = 0, nsrcLocal = 0;
    int rank = 0, nranks = 1;

    int tileWidth = 0; // resolved
    TierCfg tier[kNumTiers];

    // raw arena
    DevBuf rNode, rEid, rStay, rCutOff, rCutLen, rCutSrc, rCutSeq, rCutTier, rValid;
    DevBuf labSrc, labNode, labHops, labDist;
    DevBuf srcTier, srcArcs, srcSettled;
    // history used by the next round (DESIGN.md "issue order"): per-source work, the longest-first order built from it,
    // and whether the shared-memory tier is worth running
    DevBuf srcWork, srcOrder, orderBins;
    bool haveOrder = false, useHistory = true, overlapHeavy = true;
    long long heavyPairs = 0; // srcOrder[0 .. 2 heavyPairs): pairs with a member that finished in tier >= 2 last round
    int skipTier0 = 0; // > 0: rounds for which tier 0 is still skipped (it completed < 1/4 of its sources last time)
    long long cutCap = 0, nodeCap = 0, labCap = 0;
    // 24 x u64: see RawOut::counters; [8] queue cursor, [9] ovf count A, [10] ovf count B, [11] collisions;
    // [16] queue cursor, [17] ovf count, [18] sources run of the overlapped tier-2 launch
    DevBuf counters;
    PinBuf hCounters, hHeavy;
    bool heavyPending = false; // hHeavy is being written behind heavyEv
    DevBuf ovfA, ovfB, ovfC;
    std::vector<int32_t> hSources; // the local explicit source list as uploaded (smsgpu_set_sources)

    // dedup work
    DevBuf wCnode, wCeid, wCstay, wHash, wRank, wSlot, wKeep, wNcross, wViol, wTabKey, wTabRank, wTabIdx;
    DevBuf wCnt, wCursor, wBase, wOrder, wOutLen, wBlockSums;
    CutListDev fin;
    CutListHost host;
    bool fetched = false;

    // labels / per-source stats on the host
    PinBuf hLabOff, hLabNode, hLabDist, hLabHops, hSrcArcs, hSrcSettled, hSrcTier;
    long long nLabels = 0;
    bool haveLabels = false, haveSrcStats = false;

    // small cycles (k = 3, 4)
    DevBuf scNodes[2], scEids[2], scCode, scXc;
    cudaEvent_t scEv[4]{};
    double scMs[3] = {0, 0, 0}; // last scan: scan kernel, (count scan +) gather kernels, D2H of the walks
    long long scCount[2] = {0, 0};
    CutListHost scHost[2];
    CutListDev scFin;
    PinBuf scTmp, scListHost[2];
    // enumeration of the small cycles (smsgpu_small_cycles_enumerate)
    std::vector<int64_t> hRowptr; // host copy of the CSR (enumeration set-up)
    std::vector<int32_t> hCol;
    DevBuf dArcSrc, dColSorted, eHoleBase;
    bool enumReady = false;

    smsgpu_params lastParams{};
    smsgpu_stats stats{};

    // SMSGPU_MODE_VERBATIM: two stamp arrays (previous / current sweep), per-source "found nothing" flags, work list
    DevBuf vStampA, vStampB, vStampC, vIota, vEmpty, vEmptyB0, vActive, vFlags;
    long long vActiveCount = 0;
    long long vWindow = 0; // mode V: window length chosen by the previous round (0: default rule)

    // multi-GPU
    void *ncclComm = nullptr;
    int commRank = 0, commSize = 1;
    DevBuf gCounts, pBlock, mBlocks;
    PinBuf hGCounts;
    DevBuf mNode, mEid, mStay, mOff, mLen, mSrc, mSeq, mValid;
};

namespace {

int fail(smsgpu_ctx *c, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c)
        c->err = buf;
    else
        g_createError = buf;
    return code;
}

#define CU(call)                                                                                           \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) {                                                                           \
            (void)cudaGetLastError();                                                                      \
            return fail(ctx, e_ == cudaErrorMemoryAllocation ? SMSGPU_ERR_NOMEM : SMSGPU_ERR_CUDA,         \
                        "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));                \
        }                                                                                                  \
    } while (0)

inline unsigned gridFor(long long n, int block) { return (unsigned)std::max<long long>(1, (n + block - 1) / block); }

uint32_t log2ceil(uint64_t v) {
    uint32_t l = 0;
    while ((1ull << l) < v) ++l;
    return l;
}

} // namespace

// =====================================================================================================================
// life cycle
// =====================================================================================================================
namespace {

// Three instantiations of the search kernel (one source per WARP; sub-warp tiles were measured to corrupt each
// other on sm_100a, DESIGN.md "tile width"):
//   <32, true,  false>  tier 0: heap engine, everything in shared memory
//   <32, false, true >  tiers 1, 2: fast engine (bucket queue + presence bitmap), working memory in HBM
//   <32, false, false>  tier 3: heap engine on a dense table in HBM
template <bool SMEM, bool FAST>
cudaError_t launchSearchT(smsgpu_ctx *c, const TierCfg &t, const GraphDev &G, const SearchParams &P, const RawOut &R, cudaStream_t st,
                          int ctas) {
    auto kern = ocwSearchKernel<32, SMEM, FAST>;
    if (t.smemBytes) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t.smemBytes);
        if (e != cudaSuccess) return e;
    }
    if (FAST) {
        // This kernel may run beside the pairs kernel (heavy sources, second stream), whose CTAs need 36 KB of shared memory
        // each: an SM configured for a large L1 by a resident CTA of this kernel takes none of them until it drains
        // (measured: one such CTA on every SM cost the pairs kernel 32 ms).  Ask for the shared-memory-heavy split.
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
        if (e != cudaSuccess) return e;
    }
    kern<<<ctas, 128, t.smemBytes, st>>>(G, P, R, t.caps, t.scratch.as<unsigned char>());
    return cudaGetLastError();
}
// `ctasCap` > 0 launches fewer CTAs than the tier has tiles for (a short work list beside another launch)
cudaError_t launchSearch(smsgpu_ctx *c, int tierIdx, const GraphDev &G, const SearchParams &P, const RawOut &R, cudaStream_t st,
                         int ctasCap = 0) {
    const TierCfg &t = c->tier[tierIdx];
    const int ctas = ctasCap > 0 ? std::min(ctasCap, t.ctas) : t.ctas;
    if (tierIdx == 0) return launchSearchT<true, false>(c, t, G, P, R, st, ctas);
    if (t.pairs) {
        auto kern = G.blStamp ? ocwSearchPairsKernel<true> : ocwSearchPairsKernel<false>;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t.smemBytes);
        if (e != cudaSuccess) return e;
        PairsKArgs K{G, P, R, t.caps};
        e = c->dPairsArgs.ensure(sizeof K);
        if (e != cudaSuccess) return e;
        e = cudaMemcpyAsync(c->dPairsArgs.p, &K, sizeof K, cudaMemcpyHostToDevice, st); // pageable source: staged before return
        if (e != cudaSuccess) return e;
        const PairsParams Q{c->dPairsArgs.as<PairsKArgs>(), c->dRowinfo.as<uint2>(), c->pairsScaleU, c->pairsScaleB, c->pairsMirror ? 1 : 0, c->rowShift,
                            c->dArcs12.as<uint32_t>()};
        kern<<<ctas, 128, t.smemBytes, st>>>(G, Q, P, R, t.caps, t.scratch.as<unsigned char>());
        return cudaGetLastError();
    }
    if (t.caps.nbuckets) return launchSearchT<false, true>(c, t, G, P, R, st, ctas);
    if (t.runs) {
        cudaError_t e = cudaFuncSetAttribute(ocwSearchRunsKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RunSmem));
        if (e != cudaSuccess) return e;
        ocwSearchRunsKernel<<<ctas, kRunThreads, sizeof(RunSmem), st>>>(G, P, R, t.caps, t.scratch.as<unsigned char>());
        return cudaGetLastError();
    }
    return launchSearchT<false, false>(c, t, G, P, R, st, ctas);
}
template <bool SMEM, bool FAST>
int occupancyT(size_t smemBytes) {
    int nb = 0;
    auto kern = ocwSearchKernel<32, SMEM, FAST>;
    if (smemBytes) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, 128, smemBytes) != cudaSuccess) {
        (void)cudaGetLastError();
        nb = 1;
    }
    return std::max(nb, 1);
}

// Working-memory capacities of the three tiers (DESIGN.md "tiers").
void configureTiers(smsgpu_ctx *c) {
    const int tw = c->tileWidth;
    const int tilesPerCta = 128 / tw;
    // tier 0: everything in shared memory; sized for the balls of sparse lattices with fractional x*
    {
        TierCfg &t = c->tier[0];
        t.caps = TileCaps{128, 32 - 7, 0, 64 + (uint32_t)tw, 64 + 2 * (uint32_t)tw, 32, 16, 16};
        t.smemBytes = tileBytes(t.caps) * tilesPerCta;
        t.ctas = c->smCount * occupancyT<true, false>(t.smemBytes);
        t.ready = true;
    }
    // tiers 1 and 2: per-warp working memory in HBM, hashed table.  Tier 2 has 16x the capacities and catches the
    // heavy tail of ball sizes at full occupancy.  Capacities are clamped to what the graph can ever need.
    {
        const uint64_t n = (uint64_t)c->n, m = (uint64_t)c->m;
        const uint32_t touchedMax = (uint32_t)n + (uint32_t)tw;                                // one entry per vertex
        const uint32_t pushMax = (uint32_t)std::min<uint64_t>(4 * m + 64 + 2 * tw, 1u << 30); // one push per arc side
        const uint32_t walkMax = (uint32_t)std::min<uint64_t>(2 * n + 2, 1u << 28);
        for (int ti = 1; ti <= 2; ++ti) {
            TierCfg &t = c->tier[ti];
            const uint32_t lg = ti == 1 ? 13 : 17; // touched-entry capacity 2^lg
            const uint32_t touched = std::min<uint32_t>(1u << lg, touchedMax);
            const uint32_t tabLg = std::max<uint32_t>(log2ceil((uint64_t)2 * touched), 5);
            // fast engine: bucket queue of nbuckets x 32 keys; the hot words (2 B per bucket + 1 bit per table slot)
            // live in shared memory in tier 1 (10 KB per warp) and behind the tile in HBM in tier 2
            const uint32_t nb = ti == 1 ? (uint32_t)SMSGPU_TIER1_BUCKETS : 16384u, bc = 32u;
            t.caps = TileCaps{1u << tabLg, 32 - tabLg, 0, touched, nb * bc,
                              std::min<uint32_t>(1u << (lg - 1), pushMax), std::min<uint32_t>(ti == 1 ? 256 : 4096, walkMax),
                              ti == 1 ? 1024u : (1u << 14), nb, bc};
            t.hotInSmem = ti == 1;
            if (ti == 1 && c->pairs) {
                // pairs engine: two sources per warp, 16-slot buckets under the geometric bucket map
                const uint32_t pnb = (uint32_t)SMSGPU_PAIRS_BUCKETS, pbc = 16u;
                t.caps.nbuckets = pnb;
                t.caps.bucketCap = pbc;
                t.caps.heapCap = pnb * pbc;
                t.caps.pairs = 1;
                // the table may fill to 3/4 (the other engines stop at 1/2): only the heaviest sources get there, and a longer
                // probe sequence in tier 1 is far cheaper than a second run in tier 2 (measured: 2/3 of tier 1's overflows)
                t.caps.touchedCap = std::min<uint32_t>(t.caps.tabCap / 4 * 3, touchedMax);
                t.pairs = true;
                const int pairTiles = 2 * (128 / 32);
                t.smemBytes = hotBytes(t.caps) * pairTiles + (128 / 32) * kPairsSaveWords * sizeof(uint32_t);
                int nb2 = 0;
                cudaFuncSetAttribute(ocwSearchPairsKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t.smemBytes);
                if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb2, ocwSearchPairsKernel<false>, 128, t.smemBytes) != cudaSuccess) {
                    (void)cudaGetLastError();
                    nb2 = 1;
                }
                t.ctas = c->smCount * std::max(nb2, 1);
                t.ntiles = (long long)t.ctas * pairTiles;
                continue;
            }
            t.smemBytes = t.hotInSmem ? hotBytes(t.caps) * tilesPerCta : 0;
            t.ctas = c->smCount * occupancyT<false, true>(t.smemBytes);
            t.ntiles = (long long)t.ctas * tilesPerCta;
        }
    }
    // tier 3: dense table indexed by vertex, capacities that cannot overflow before the graph is exhausted
    {
        TierCfg &t = c->tier[3];
        const uint64_t n = (uint64_t)c->n, m = (uint64_t)c->m;
        const uint32_t big = (uint32_t)std::min<uint64_t>(4 * m + 64 + 2 * tw, 1u << 30);
        t.caps = TileCaps{(uint32_t)n, 0, 1, (uint32_t)n + (uint32_t)tw, big, big,
                          (uint32_t)std::min<uint64_t>(2 * n + 2, 1u << 28),
                          (uint32_t)std::min<uint64_t>(4 * m + 16, 1u << 22)};
        // the runs engine (a CTA per source, whole runs of equal (dist, hops) settled at once); SMSGPU_RUNS=0 puts the
        // one-key-at-a-time heap engine back (kept for comparison and tested)
        const char *er = std::getenv("SMSGPU_RUNS");
        t.runs = !(er && er[0] == '0');
    }
}

int ensureTierScratch(smsgpu_ctx *ctx, int ti) {
    TierCfg &t = ctx->tier[ti];
    if (t.ready) return 0;
    const size_t tb = tileBytes(t.caps) + (t.hotInSmem ? 0 : hotBytes(t.caps)) + (t.runs ? runScratchBytes(ctx->n) : 0); // tile stride
    if (t.runs) { // one tile per CTA: as many CTAs as can be resident, as far as a quarter of the free memory goes
        size_t freeB = 0, totalB = 0;
        CU(cudaMemGetInfo(&freeB, &totalB));
        int perSm = 1;
        cudaFuncSetAttribute(ocwSearchRunsKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RunSmem));
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, ocwSearchRunsKernel, kRunThreads, sizeof(RunSmem)) != cudaSuccess) {
            (void)cudaGetLastError();
            perSm = 1;
        }
        long long nt = std::min<long long>((long long)((freeB / 4) / tb), (long long)ctx->smCount * std::max(perSm, 1));
        if (nt < 1) return fail(ctx, SMSGPU_ERR_NOMEM, "runs tier needs %zu bytes per tile, %zu free", tb, freeB);
        t.ctas = (int)nt;
        t.ntiles = nt;
    } else if (ti >= 2) { // large tiles: as many as fit in a quarter of the free memory, at most the resident warps
        size_t freeB = 0, totalB = 0;
        CU(cudaMemGetInfo(&freeB, &totalB));
        long long nt = (long long)((freeB / 4) / tb);
        const int tilesPerCta = 128 / ctx->tileWidth;
        // dense tier: as many tiles as warps can be resident (LP-like x* with its exact distance ties sends whole rounds
        // here: no bucket holds a zero-weight component; measured 2 -> 4 CTAs per SM: see profiles/README.md)
        nt = std::min<long long>(nt, ti == 2 ? t.ntiles : (long long)ctx->smCount * std::max(1, occupancyT<false, false>(0)) * tilesPerCta);
        if (nt < 1) return fail(ctx, SMSGPU_ERR_NOMEM, "dense tier needs %zu bytes per tile, %zu free", tb, freeB);
        t.ctas = (int)std::max<long long>(1, nt / tilesPerCta);
        t.ntiles = (long long)t.ctas * tilesPerCta;
    }
    CU(t.scratch.ensure((size_t)t.ntiles * tb));
    initTablesKernel<<<ctx->smCount * 8, 256, 0, ctx->stream>>>(t.scratch.as<unsigned char>(), tb, t.ntiles, t.caps.tabCap);
    CU(cudaGetLastError());
    t.stampNext = 0;
    t.ready = true;
    return 0;
}


// ---- lanes engine: working memory and one pass over a work list -----------------------------------------------------------
// Tiles for `items` searches (at most lanesMaxTiles, at most 60 % of the free memory), in lanesSets sets.
int ensureLanes(smsgpu_ctx *ctx, long long items) {
    const TileCaps &pc = ctx->tier[1].caps;
    LanesCfg C{};
    C.tabCap = pc.tabCap;
    C.tabShift = pc.tabShift;
    C.touchedCap = pc.touchedCap;
    if (ctx->lanesTabLg > 0 && (1u << ctx->lanesTabLg) < pc.tabCap) { // a smaller table: more tiles fit, larger balls go to the pairs engine
        C.tabCap = 1u << ctx->lanesTabLg;
        C.tabShift = 32 - (uint32_t)ctx->lanesTabLg;
        C.touchedCap = std::min<uint32_t>(pc.touchedCap, C.tabCap / 4 * 3);
    }
    C.histCap = pc.histCap;
    C.nbuckets = (uint32_t)ctx->lanesBuckets;
    C.qCap = (uint32_t)ctx->lanesQCap;
    C.scaleU = ctx->pairsScaleU;
    C.scaleB = ctx->pairsScaleB * ((float)ctx->lanesBuckets / (float)SMSGPU_PAIRS_BUCKETS); // the pairs engine's map over this engine's buckets
    C.rowShift = ctx->rowShift;
    ctx->lanesCfg = C;
    TileCaps F{};
    F.tabCap = C.tabCap;
    F.tabShift = C.tabShift;
    F.dense = 0;
    F.touchedCap = C.touchedCap;
    F.heapCap = pc.histCap; // selectIntermediates' key scratch
    F.histCap = pc.histCap;
    F.walkCap = pc.walkCap;
    F.chordCap = pc.chordCap;
    ctx->lanesFcaps = F;
    const size_t stride = lanesTileBytes(C);
    long long want = std::max<long long>(1, std::min<long long>(items, ctx->lanesMaxTiles));
    int sets = (want >= 512) ? std::max(1, std::min(ctx->lanesSetsWanted, 4)) : 1;
    long long perSet = (want + sets - 1) / sets;
    perSet = (perSet + 127) & ~127ll; // whole CTAs of the search kernel
    if (ctx->lanesScratch && ctx->lanesTileStride == stride && want <= ctx->lanesAllocWant) return 0; // what is there is enough (or all there can be)
    if (ctx->lanesScratch) {
        CU(cudaDeviceSynchronize());
        cudaFree(ctx->lanesScratch);
        ctx->lanesScratch = nullptr;
        ctx->lanesScratchBytes = 0;
    }
    size_t freeB = 0, totalB = 0;
    CU(cudaMemGetInfo(&freeB, &totalB));
    const long long fit = (long long)((freeB / 100 * 60) / stride);
    if (fit < 128) return fail(ctx, SMSGPU_ERR_NOMEM, "lanes engine needs %zu bytes per tile, %zu free", stride, freeB);
    ctx->lanesAllocWant = want;
    if (sets * perSet > fit) {
        perSet = (fit / sets) & ~127ll;
        if (perSet < 128) {
            sets = 1;
            perSet = fit & ~127ll;
        }
        ctx->lanesAllocWant = ctx->lanesMaxTiles; // all there can be
    }
    if (want >= ctx->lanesMaxTiles) ctx->lanesAllocWant = 1ll << 62;
    const size_t bytes = (size_t)sets * (size_t)perSet * stride;
    CU(cudaMalloc(&ctx->lanesScratch, bytes));
    ctx->lanesScratchBytes = bytes;
    CU(cudaMemsetAsync(ctx->lanesScratch, 0, bytes, ctx->stream)); // no entry carries a stamp
    ctx->lanesStamp = 0;
    ctx->lanesSets = sets;
    ctx->lanesTilesPerSet = perSet;
    ctx->lanesTileStride = stride;
    // finishing warps: their walk scratch (table and history live in the tiles)
    TileCaps sc = F;
    sc.tabCap = 0;
    sc.touchedCap = 0;
    sc.histCap = 0;
    ctx->lanesWarpStride = (tileBytes(sc) + 255) & ~(size_t)255;
    ctx->lanesFinishCtas = (int)std::min<long long>((perSet + 3) / 4, (long long)ctx->smCount * 8);
    CU(ctx->lanesWarpScratch.ensure((size_t)sets * (size_t)ctx->lanesFinishCtas * 4 * ctx->lanesWarpStride));
    for (int k = 0; k < sets; ++k) {
        if (!ctx->laneStream[k]) CU(cudaStreamCreateWithFlags(&ctx->laneStream[k], cudaStreamNonBlocking));
        if (!ctx->laneEv[k]) CU(cudaEventCreateWithFlags(&ctx->laneEv[k], cudaEventDisableTiming));
    }
    if (!ctx->lanesForkEv) CU(cudaEventCreateWithFlags(&ctx->lanesForkEv, cudaEventDisableTiming));
    if (!ctx->lanesT0) CU(cudaEventCreate(&ctx->lanesT0));
    if (!ctx->lanesT1) CU(cudaEventCreate(&ctx->lanesT1));
    if (std::getenv("SMSGPU_DEBUG_OVF"))
        std::fprintf(stderr, "[smsgpu] lanes engine: %d sets x %lld tiles x %zu bytes = %.1f GB\n", sets, perSet, stride, (double)bytes / 1e9);
    return 0;
}

// One pass of the lanes engine over wl[0 .. nw) (nullptr: all local sources): batches of tiles alternate between the lane
// streams; everything is ordered behind `st` at the start and `st` waits for all of it at the end.  Sources that do not
// finish here are appended to P.ovfList / P.ovfCount.
int runLanes(smsgpu_ctx *ctx, const GraphDev &G, SearchParams P, const RawOut &R, const int32_t *wl, long long nw, cudaStream_t st) {
    const int itemSize = ctx->pairsMirror ? 2 : 1;
    const long long items = (nw + itemSize - 1) / itemSize;
    if (int rc = ensureLanes(ctx, items)) return rc;
    P.worklist = wl;
    P.nwork = nw;
    P.tier = 1;
    PairsKArgs K{G, P, R, ctx->lanesFcaps};
    CU(ctx->dLanesArgs.ensure(sizeof K));
    CU(cudaMemcpyAsync(ctx->dLanesArgs.p, &K, sizeof K, cudaMemcpyHostToDevice, st)); // pageable source: staged before return
    CU(cudaEventRecord(ctx->lanesT0, st));
    CU(cudaEventRecord(ctx->lanesForkEv, st));
    for (int k = 0; k < ctx->lanesSets; ++k) CU(cudaStreamWaitEvent(ctx->laneStream[k], ctx->lanesForkEv, 0));
    const long long B = ctx->lanesTilesPerSet;
    int batch = 0;
    for (long long first = 0; first < items; first += B, ++batch) {
        const int set = batch % ctx->lanesSets;
        cudaStream_t s = ctx->laneStream[set];
        if (++ctx->lanesStamp == 0) { // stamps exhausted (2^32 batches): forget every old entry
            CU(cudaDeviceSynchronize());
            CU(cudaMemsetAsync(ctx->lanesScratch, 0, ctx->lanesScratchBytes, s));
            CU(cudaStreamSynchronize(s));
            ctx->lanesStamp = 1;
        }
        LanesArgs A{};
        A.cfg = ctx->lanesCfg;
        A.rowinfo = ctx->dRowinfo.as<uint2>();
        A.tiles = static_cast<unsigned char *>(ctx->lanesScratch) + (size_t)set * (size_t)B * ctx->lanesTileStride;
        A.tileStride = ctx->lanesTileStride;
        A.first = first;
        A.count = std::min<long long>(B, items - first);
        A.itemSize = itemSize;
        A.stamp = ctx->lanesStamp;
        A.kargs = ctx->dLanesArgs.as<PairsKArgs>();
        A.warpScratch = ctx->lanesWarpScratch.as<unsigned char>() + (size_t)set * (size_t)ctx->lanesFinishCtas * 4 * ctx->lanesWarpStride;
        A.warpScratchStride = ctx->lanesWarpStride;
        A.fcaps = ctx->lanesFcaps;
        ocwLanesSearchKernel<<<(unsigned)((A.count + SMSGPU_LANES_BLOCK - 1) / SMSGPU_LANES_BLOCK), SMSGPU_LANES_BLOCK, 0, s>>>(G, P, A);
        CU(cudaGetLastError());
        ocwLanesFinishKernel<<<(unsigned)std::min<long long>((A.count + 3) / 4, ctx->lanesFinishCtas), 128, 0, s>>>(A);
        CU(cudaGetLastError());
        ctx->stats.search_launches += 2;
        ctx->stats.total_launches += 2;
    }
    for (int k = 0; k < ctx->lanesSets; ++k) {
        CU(cudaEventRecord(ctx->laneEv[k], ctx->laneStream[k]));
        CU(cudaStreamWaitEvent(st, ctx->laneEv[k], 0));
    }
    CU(cudaEventRecord(ctx->lanesT1, st));
    return 0;
}

} // namespace

extern "C" {

int smsgpu_abi_version(void) { return SMSGPU_ABI_VERSION; }

void smsgpu_default_params(smsgpu_params *p) {
    if (!p) return;
    std::memset(p, 0, sizeof *p);
    p->max_intermediates = 10;
    p->best_dist_init = 1.0 - 1e-6;
    p->dedup = 1;
}

const char *smsgpu_last_error(const smsgpu_ctx *ctx) { return ctx ? ctx->err.c_str() : g_createError.c_str(); }

int smsgpu_create(smsgpu_ctx **out, int64_t n, int64_t m, const int64_t *rowptr, const int32_t *col, const int32_t *eid,
                  int device) {
    smsgpu_ctx *ctx = nullptr; // errors before the context exists go to the thread-local slot
    if (!out) return fail(ctx, SMSGPU_ERR_ARG, "out is null");
    *out = nullptr;
    if (n <= 0 || m < 0 || !rowptr || (m > 0 && (!col || !eid))) return fail(ctx, SMSGPU_ERR_ARG, "null or empty graph");
    if (n >= (1ll << 30) || 2 * m >= (1ll << 30)) return fail(ctx, SMSGPU_ERR_ARG, "graph too large for 32-bit arc ids");
    // the reference rejects non-compact / un-indexed graphs (ocw_seperator.cpp:16-19); here: a malformed CSR
    if (rowptr[0] != 0 || rowptr[n] != 2 * m) return fail(ctx, SMSGPU_ERR_ARG, "rowptr[0] != 0 or rowptr[n] != 2m");
    {
        std::vector<int32_t> seen(n, -1);
        std::vector<uint8_t> eidCount(m, 0);
        for (int64_t u = 0; u < n; ++u) {
            if (rowptr[u + 1] < rowptr[u] || rowptr[u + 1] > 2 * m) return fail(ctx, SMSGPU_ERR_ARG, "rowptr not monotone or beyond 2m at %lld", (long long)u);
            for (int64_t a = rowptr[u]; a < rowptr[u + 1]; ++a) {
                const int32_t v = col[a];
                if (v < 0 || v >= n) return fail(ctx, SMSGPU_ERR_ARG, "neighbour out of range at arc %lld", (long long)a);
                if (v == u) return fail(ctx, SMSGPU_ERR_ARG, "self loop at vertex %lld", (long long)u);
                if (seen[v] == (int32_t)u) return fail(ctx, SMSGPU_ERR_ARG, "multi-edge %lld-%d", (long long)u, v);
                seen[v] = (int32_t)u;
                if (eid[a] < 0 || eid[a] >= m) return fail(ctx, SMSGPU_ERR_ARG, "edge id out of range at arc %lld", (long long)a);
                if (++eidCount[eid[a]] > 2) return fail(ctx, SMSGPU_ERR_ARG, "edge id %d used more than twice", eid[a]);
            }
        }
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        (void)cudaGetLastError();
        return fail(ctx, SMSGPU_ERR_CUDA, "no CUDA device (libsmsgpu has no CPU fallback)");
    }
    if (device < 0 || device >= ndev) return fail(ctx, SMSGPU_ERR_ARG, "device %d out of range (%d devices)", device, ndev);
    ctx = new smsgpu_ctx();
    ctx->device = device;
    ctx->n = n;
    ctx->m = m;
    ctx->hRowptr.assign(rowptr, rowptr + n + 1);
    ctx->hCol.assign(col, col + 2 * m);
    auto bail = [&](int code) {
        g_createError = ctx->err;
        smsgpu_destroy(ctx);
        return code;
    };
#define CUB(call)                                                                                          \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) {                                                                           \
            (void)cudaGetLastError();                                                                      \
            fail(ctx, SMSGPU_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return bail(e_ == cudaErrorMemoryAllocation ? SMSGPU_ERR_NOMEM : SMSGPU_ERR_CUDA);             \
        }                                                                                                  \
    } while (0)
    CUB(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUB(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        fail(ctx, SMSGPU_ERR_CUDA, "device %d is sm_%d%d; libsmsgpu is built for sm_100a only", device, prop.major, prop.minor);
        return bail(SMSGPU_ERR_CUDA);
    }
    ctx->smCount = prop.multiProcessorCount;
    CUB(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    for (auto &e : ctx->ev) CUB(cudaEventCreate(&e));
    for (auto &e : ctx->tierEv) CUB(cudaEventCreate(&e));
    for (auto &e : ctx->scEv) CUB(cudaEventCreate(&e));
    {
        int prLow = 0, prHigh = 0; // its CTAs take free SM slots before tier 1's
        CUB(cudaDeviceGetStreamPriorityRange(&prLow, &prHigh));
        CUB(cudaStreamCreateWithPriority(&ctx->stream2, cudaStreamNonBlocking, prHigh));
    }
    CUB(cudaEventCreateWithFlags(&ctx->joinEv, cudaEventDisableTiming));
    CUB(cudaEventCreateWithFlags(&ctx->heavyEv, cudaEventDisableTiming));
    CUB(ctx->hHeavy.ensure(16));
    CUB(ctx->dRowptr.ensure((size_t)(n + 1) * sizeof(int64_t)));
    CUB(ctx->dCol.ensure((size_t)std::max<int64_t>(2 * m, 4) * sizeof(int32_t)));
    CUB(ctx->dEid.ensure((size_t)std::max<int64_t>(2 * m, 4) * sizeof(int32_t)));
    CUB(ctx->dArcs.ensure((size_t)std::max<int64_t>(2 * m, 4) * sizeof(Arc)));
#ifdef SMSGPU_PAIRS_ARCS12
    CUB(ctx->dArcs12.ensure((size_t)std::max<int64_t>(2 * m, 4) * 12));
#endif
    CUB(cudaMemcpyAsync(ctx->dRowptr.p, rowptr, (size_t)(n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
    if (m > 0) {
        CUB(cudaMemcpyAsync(ctx->dCol.p, col, (size_t)2 * m * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
        CUB(cudaMemcpyAsync(ctx->dEid.p, eid, (size_t)2 * m * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    }
    {
        // Arc::row of every adjacency entry: the neighbour's first arc in the low rowShift bits, its degree above
        // (saturated: the all-ones field sends the reader to the row record)
        uint32_t shift = 1;
        while (shift < 31 && (1ll << shift) < std::max<int64_t>(2 * m, 2)) ++shift;
        ctx->rowShift = shift;
        const uint32_t degMax = 0xFFFFFFFFu >> shift;
        std::vector<uint32_t> rp((size_t)std::max<int64_t>(2 * m, 4), 0u);
        for (int64_t a = 0; a < 2 * m; ++a) {
            const int64_t v = col[a];
            const uint64_t deg = (uint64_t)(rowptr[v + 1] - rowptr[v]);
            rp[(size_t)a] = (uint32_t)rowptr[v] | ((uint32_t)std::min<uint64_t>(deg, degMax) << shift);
        }
        CUB(ctx->dRowpack.ensure(rp.size() * sizeof(uint32_t)));
        CUB(cudaMemcpy(ctx->dRowpack.p, rp.data(), rp.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    }
    CUB(ctx->dX.ensure((size_t)std::max<int64_t>(m, 1) * sizeof(double)));
    CUB(ctx->dBl.ensure((size_t)n));
    CUB(ctx->hX.ensure((size_t)std::max<int64_t>(m, 1) * sizeof(double)));
    CUB(ctx->hBl.ensure((size_t)n));
    CUB(ctx->counters.ensure(24 * sizeof(unsigned long long)));
    CUB(ctx->hCounters.ensure(24 * sizeof(unsigned long long)));
    CUB(cudaStreamSynchronize(ctx->stream));
    // tile width by average degree: one lane per arc of a typical row
    const double avgDeg = n ? 2.0 * (double)m / (double)n : 0.0;
    (void)avgDeg;
    ctx->tileWidth = 32;
    // pairs engine (tier 1): needs 32-bit arc indices; SMSGPU_PAIRS=0 selects the one-source-per-warp engine
    {
        const char *ev = std::getenv("SMSGPU_PAIRS");
        ctx->pairs = (2 * m < (int64_t)0xFFFFFFFFll) && !(ev && ev[0] == '0');
        if (const char *em = std::getenv("SMSGPU_MIRROR")) ctx->pairsMirror = !(em[0] == '0');
        if (ctx->pairs) {
            std::vector<uint2> ri((size_t)std::max<int64_t>(n, 1));
            for (int64_t u = 0; u < n; ++u) ri[u] = make_uint2((uint32_t)rowptr[u], (uint32_t)(rowptr[u + 1] - rowptr[u]));
            CUB(ctx->dRowinfo.ensure(ri.size() * sizeof(uint2)));
            CUB(cudaMemcpy(ctx->dRowinfo.p, ri.data(), (size_t)n * sizeof(uint2), cudaMemcpyHostToDevice));
            // bucket map: 2^u, u in [0, L] over [0, bestInit/2]; L ~ log2 of the ball growth over the radius.  Measured on
            // config 5 (degree 8): L = 6..8 leaves 0.2 % of the sources with a full bucket, L = 10 0.6 %, L = 14 9 %, L <= 4
            // more again (and slower extraction); the round is 4 % faster at 7 than at 10.  SMSGPU_PAIRS_L overrides, 0 = linear
            int L = (int)std::lround(0.85 * avgDeg);
            if (const char *el = std::getenv("SMSGPU_PAIRS_L")) L = std::atoi(el);
            L = std::max(0, std::min(L, 20));
            const double half = 0.5 * (1.0 - 1e-6); // bestInit / 2 of the default parameters; larger radii clamp to the last bucket
            const double NBd = (double)SMSGPU_PAIRS_BUCKETS;
            if (L > 0) {
                ctx->pairsScaleU = (float)((double)L / half);
                ctx->pairsScaleB = (float)(NBd / (std::ldexp(1.0, L) - 1.0));
            } else {
                ctx->pairsScaleU = 0.f;
                ctx->pairsScaleB = (float)(NBd / half);
            }
        }
    }
    // The arc records are the one array every search shares; everything else the searches touch is private working memory
    // that streams through L2 once or twice.  With SMSGPU_L2_PERSIST=1 the arc array gets an access-policy window with
    // the persisting property on both streams (as much of it as the device lets persist), so that the streaming
    // working-memory traffic cannot evict it.  Off by default: the carve-out is a device-wide limit.
    if (const char *ep = std::getenv("SMSGPU_L2_PERSIST")) {
        if (ep[0] == '1' && m > 0) {
            const size_t arcBytes = (size_t)2 * m * sizeof(Arc);
            const size_t persist = std::min<size_t>((size_t)prop.persistingL2CacheMaxSize, arcBytes);
            const size_t window = std::min<size_t>((size_t)prop.accessPolicyMaxWindowSize, arcBytes);
            if (persist > 0 && window > 0) {
                CUB(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, persist));
                cudaStreamAttrValue av{};
                av.accessPolicyWindow.base_ptr = ctx->dArcs.p;
                av.accessPolicyWindow.num_bytes = window;
                av.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)persist / (double)window);
                av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                CUB(cudaStreamSetAttribute(ctx->stream, cudaStreamAttributeAccessPolicyWindow, &av));
                CUB(cudaStreamSetAttribute(ctx->stream2, cudaStreamAttributeAccessPolicyWindow, &av));
                if (std::getenv("SMSGPU_DEBUG_OVF"))
                    std::fprintf(stderr, "[smsgpu] L2 persisting window: %zu of %zu arc bytes (L2 %d MB, max persisting %d MB, max window %d MB)\n",
                                 persist, arcBytes, prop.l2CacheSize >> 20, prop.persistingL2CacheMaxSize >> 20, prop.accessPolicyMaxWindowSize >> 20);
            }
        }
    }
    // L2 fetch granularity (a device-wide hint, default 64 B): the searches read isolated 32-byte sectors
    if (const char *eg = std::getenv("SMSGPU_L2_FETCH")) {
        const int g = std::atoi(eg);
        if (g == 32 || g == 64 || g == 128) {
            if (cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)g) != cudaSuccess) (void)cudaGetLastError();
            size_t got = 0;
            if (cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity) != cudaSuccess) (void)cudaGetLastError();
            if (std::getenv("SMSGPU_DEBUG_OVF")) std::fprintf(stderr, "[smsgpu] L2 fetch granularity: asked %d, now %zu\n", g, got);
        }
    }
    if (const char *el = std::getenv("SMSGPU_LANES")) ctx->lanesMode = el[0] == '1' ? 1 : (el[0] == '2' ? 2 : 0);
    if (const char *el = std::getenv("SMSGPU_LANES_MIN")) ctx->lanesMinSources = std::atoll(el);
    if (const char *el = std::getenv("SMSGPU_LANES_TILES")) ctx->lanesMaxTiles = std::max<long long>(128, std::atoll(el));
    if (const char *el = std::getenv("SMSGPU_LANES_TABLG")) ctx->lanesTabLg = std::atoi(el) > 0 ? std::max(6, std::min(20, std::atoi(el))) : 0;
    if (const char *el = std::getenv("SMSGPU_LANES_NB")) ctx->lanesBuckets = std::max(64, std::min(2048, std::atoi(el) & ~3));
    if (const char *el = std::getenv("SMSGPU_LANES_QCAP")) ctx->lanesQCap = std::max(64, std::min(8192, std::atoi(el)));
    if (const char *el = std::getenv("SMSGPU_LANES_SETS")) ctx->lanesSetsWanted = std::max(1, std::min(4, std::atoi(el)));
    if (const char *eh = std::getenv("SMSGPU_NO_HISTORY")) ctx->useHistory = !(eh[0] == '1');
    if (const char *eo = std::getenv("SMSGPU_NO_OVERLAP")) ctx->overlapHeavy = !(eo[0] == '1');
    configureTiers(ctx);
    ctx->nsrcGlobal = ctx->nsrcLocal = n; // default source list 0..n-1 (conshdlr_oscw.cpp:188-195)
    ctx->rank = 0;
    ctx->nranks = 1;
    *out = ctx;
    return SMSGPU_OK;
#undef CUB
}

void smsgpu_destroy(smsgpu_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    smsgpu_comm_destroy(ctx);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->stream2) cudaStreamSynchronize(ctx->stream2);
    DevBuf *dev[] = {&ctx->dArcs12, &ctx->dPairsArgs, &ctx->dRowinfo, &ctx->dRowpack, &ctx->dRowptr, &ctx->dCol, &ctx->dEid, &ctx->dArcs, &ctx->dX, &ctx->dBl, &ctx->dSources,
                     &ctx->tier[0].scratch, &ctx->tier[1].scratch, &ctx->tier[2].scratch, &ctx->tier[3].scratch, &ctx->rNode, &ctx->rEid,
                     &ctx->rStay, &ctx->rCutOff, &ctx->rCutLen, &ctx->rCutSrc, &ctx->rCutSeq, &ctx->rCutTier, &ctx->rValid,
                     &ctx->labSrc, &ctx->labNode, &ctx->labHops, &ctx->labDist, &ctx->srcTier, &ctx->srcArcs,
                     &ctx->srcSettled, &ctx->srcWork, &ctx->srcOrder, &ctx->orderBins, &ctx->counters, &ctx->ovfA, &ctx->ovfB, &ctx->ovfC, &ctx->wCnode, &ctx->wCeid, &ctx->wCstay,
                     &ctx->wHash, &ctx->wRank, &ctx->wSlot, &ctx->wKeep, &ctx->wNcross, &ctx->wViol, &ctx->wTabKey,
                     &ctx->wTabRank, &ctx->wTabIdx, &ctx->wCnt, &ctx->wCursor, &ctx->wBase, &ctx->wOrder, &ctx->wOutLen,
                     &ctx->wBlockSums, &ctx->fin.off, &ctx->fin.node, &ctx->fin.eid, &ctx->fin.stay, &ctx->fin.ncross,
                     &ctx->fin.src, &ctx->fin.seq, &ctx->fin.viol, &ctx->scFin.off, &ctx->scFin.node, &ctx->scFin.eid,
                     &ctx->scFin.stay, &ctx->scFin.ncross, &ctx->scFin.src, &ctx->scFin.viol, &ctx->scNodes[0], &ctx->scNodes[1], &ctx->scEids[0],
                     &ctx->scEids[1], &ctx->gCounts,
                     &ctx->pBlock, &ctx->mBlocks, &ctx->dArcSrc, &ctx->dColSorted, &ctx->eHoleBase, &ctx->scCode, &ctx->scXc, &ctx->vStampA, &ctx->vStampB, &ctx->vStampC, &ctx->vIota, &ctx->vEmpty, &ctx->vEmptyB0, &ctx->vActive, &ctx->vFlags, &ctx->mNode, &ctx->mEid, &ctx->mStay, &ctx->mOff, &ctx->mLen, &ctx->mSrc,
                     &ctx->mSeq, &ctx->mValid};
    for (auto *b : dev) b->release();
    PinBuf *pin[] = {&ctx->hX, &ctx->hBl, &ctx->hCounters, &ctx->hHeavy, &ctx->host.off, &ctx->host.node, &ctx->host.eid,
                     &ctx->host.stay, &ctx->host.ncross, &ctx->host.src, &ctx->host.viol, &ctx->hLabOff, &ctx->hLabNode,
                     &ctx->hLabDist, &ctx->hLabHops, &ctx->hSrcArcs, &ctx->hSrcSettled, &ctx->hSrcTier, &ctx->scTmp, &ctx->scListHost[0], &ctx->scListHost[1],
                     &ctx->hGCounts};
    for (auto *b : pin) b->release();
    for (int k = 0; k < 2; ++k) {
        PinBuf *sc[] = {&ctx->scHost[k].off, &ctx->scHost[k].node, &ctx->scHost[k].eid, &ctx->scHost[k].stay,
                        &ctx->scHost[k].ncross, &ctx->scHost[k].src, &ctx->scHost[k].viol};
        for (auto *b : sc) b->release();
    }
    for (auto &e : ctx->ev)
        if (e) cudaEventDestroy(e);
    for (auto &e : ctx->tierEv)
        if (e) cudaEventDestroy(e);
    for (auto &e : ctx->scEv)
        if (e) cudaEventDestroy(e);
    for (auto &ls : ctx->laneStream)
        if (ls) cudaStreamDestroy(ls);
    for (auto &le : ctx->laneEv)
        if (le) cudaEventDestroy(le);
    if (ctx->lanesForkEv) cudaEventDestroy(ctx->lanesForkEv);
    if (ctx->lanesT0) cudaEventDestroy(ctx->lanesT0);
    if (ctx->lanesT1) cudaEventDestroy(ctx->lanesT1);
    if (ctx->lanesScratch) cudaFree(ctx->lanesScratch);
    ctx->lanesWarpScratch.release();
    ctx->dLanesArgs.release();
    ctx->ovfD.release();
    if (ctx->joinEv) cudaEventDestroy(ctx->joinEv);
    if (ctx->heavyEv) cudaEventDestroy(ctx->heavyEv);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

// ---- per round input ----------------------------------------------------------------------------------------
int smsgpu_set_round_input(smsgpu_ctx *ctx, const double *x, const uint8_t *bl) {
    if (!ctx) return SMSGPU_ERR_ARG;
    if (!x || !bl) return fail(ctx, SMSGPU_ERR_ARG, "x_by_eid or blacklist is null");
    CU(cudaSetDevice(ctx->device));
    // pageable -> pinned staging -> device, all on the context's stream.  No staging copy for the context's own buffers
    // (smsgpu_input_buffers) nor for any other page-locked buffer of the caller (cudaHostAlloc / cudaHostRegister): the DMA
    // reads it where it lies -- the caller must leave it alone until the round has been launched (smsgpu_separate*).
    auto pinned = [](const void *p) {
        cudaPointerAttributes a{};
        if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
            (void)cudaGetLastError();
            return false;
        }
        return a.type == cudaMemoryTypeHost;
    };
    const void *sx = x, *sb = bl;
    if (x != ctx->hX.p && !pinned(x)) {
        std::memcpy(ctx->hX.p, x, (size_t)ctx->m * sizeof(double));
        sx = ctx->hX.p;
    }
    if (bl != ctx->hBl.p && !pinned(bl)) {
        std::memcpy(ctx->hBl.p, bl, (size_t)ctx->n);
        sb = ctx->hBl.p;
    }
    CU(cudaEventRecord(ctx->ev[6], ctx->stream));
    if (ctx->m) CU(cudaMemcpyAsync(ctx->dX.p, sx, (size_t)ctx->m * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->dBl.p, sb, (size_t)ctx->n, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaEventRecord(ctx->ev[7], ctx->stream));
    ctx->xPtr = ctx->dX.as<double>();
    ctx->blPtr = ctx->dBl.as<uint8_t>();
    ctx->haveInput = true;
    ctx->arcsDirty = true;
    return SMSGPU_OK;
}

int smsgpu_input_buffers(smsgpu_ctx *ctx, double **x, uint8_t **bl) {
    if (!ctx || !x || !bl) return SMSGPU_ERR_ARG;
    *x = ctx->hX.as<double>();
    *bl = ctx->hBl.as<uint8_t>();
    return SMSGPU_OK;
}

int smsgpu_set_round_input_device(smsgpu_ctx *ctx, const double *dx, const uint8_t *dbl) {
    if (!ctx) return SMSGPU_ERR_ARG;
    if (!dx || !dbl) return fail(ctx, SMSGPU_ERR_ARG, "device x or blacklist is null");
    ctx->xPtr = dx;
    ctx->blPtr = dbl;
    ctx->haveInput = true;
    ctx->arcsDirty = true;
    CU(cudaSetDevice(ctx->device));
    CU(cudaEventRecord(ctx->ev[6], ctx->stream));
    CU(cudaEventRecord(ctx->ev[7], ctx->stream));
    return SMSGPU_OK;
}

int smsgpu_set_sources(smsgpu_ctx *ctx, const int32_t *sources, int64_t nsrc, int rank, int nranks) {
    if (!ctx) return SMSGPU_ERR_ARG;
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(ctx, SMSGPU_ERR_ARG, "bad rank %d / %d", rank, nranks);
    if (sources && nsrc < 0) return fail(ctx, SMSGPU_ERR_ARG, "negative source count");
    CU(cudaSetDevice(ctx->device));
    // the same list again (a handler that sets its sources before every round): the history of the previous round stays
    bool same = ctx->rank == rank && ctx->nranks == nranks && ctx->nsrcGlobal == (sources ? nsrc : ctx->n) &&
                ctx->explicitSources == (sources != nullptr);
    ctx->rank = rank;
    ctx->nranks = nranks;
    ctx->nsrcGlobal = sources ? nsrc : ctx->n;
    // PAIRS of consecutive sources are dealt cyclically (the two sides of a vertex stay on one GPU, where their mirror-image
    // searches share a warp): local index j <-> global index ((j >> 1) * nranks + rank) * 2 + (j & 1)
    {
        const long long N = ctx->nsrcGlobal, pairsAll = (N + 1) / 2;
        const long long pairsMine = pairsAll > rank ? (pairsAll - rank + nranks - 1) / nranks : 0;
        long long cnt = 2 * pairsMine;
        if (pairsMine > 0 && (N & 1) && ((pairsMine - 1) * nranks + rank) == pairsAll - 1) cnt--; // the odd last source
        ctx->nsrcLocal = cnt;
    }
    ctx->explicitSources = sources != nullptr;
    if (sources) {
        std::vector<int32_t> mine((size_t)ctx->nsrcLocal);
        for (int64_t j = 0; j < ctx->nsrcLocal; ++j) {
            mine[j] = sources[((j >> 1) * nranks + rank) * 2 + (j & 1)];
            if (mine[j] < 0) return fail(ctx, SMSGPU_ERR_ARG, "negative source id");
        }
        same = same && mine == ctx->hSources;
        if (!same) {
            CU(ctx->dSources.ensure(std::max<size_t>(mine.size(), 1) * sizeof(int32_t)));
            CU(cudaMemcpyAsync(ctx->dSources.p, mine.data(), mine.size() * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            ctx->hSources.swap(mine);
        }
    } else {
        ctx->hSources.clear();
    }
    if (!same) { // the local index space changed
        ctx->vWindow = 0;
        ctx->haveOrder = false;
        ctx->skipTier0 = 0;
        ctx->heavyPairs = 0;
        ctx->heavyPending = false;
    }
    return SMSGPU_OK;
}

} // extern "C"

// =====================================================================================================================
// one separation round
// =====================================================================================================================
namespace {

int exclusiveScan(smsgpu_ctx *ctx, const int32_t *in, long long n, long long *out) {
    if (n == 0) {
        CU(cudaMemsetAsync(out, 0, sizeof(long long), ctx->stream));
        return 0;
    }
    const long long nb = (n + kScanTile - 1) / kScanTile;
    CU(ctx->wBlockSums.ensure((size_t)(nb + 1) * sizeof(long long)));
    scanReduceKernel<<<(unsigned)nb, kScanBlock, 0, ctx->stream>>>(in, n, ctx->wBlockSums.as<long long>());
    scanBlockSumsKernel<<<1, kScanBlock, 0, ctx->stream>>>(ctx->wBlockSums.as<long long>(), nb);
    scanFinalKernel<<<(unsigned)nb, kScanBlock, 0, ctx->stream>>>(in, n, ctx->wBlockSums.as<long long>(), out);
    CU(cudaGetLastError());
    ctx->stats.total_launches += 3;
    return 0;
}

__global__ void validKernel(const long long n, const int8_t *cutTier, const int32_t *cutSrc, const int8_t *srcTier,
                            uint8_t *valid) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) valid[i] = cutTier[i] == srcTier[cutSrc[i]];
}

int ensureRawArena(smsgpu_ctx *ctx, long long cutCap, long long nodeCap) {
    if (cutCap > ctx->cutCap) {
        CU(ctx->rCutOff.ensure((size_t)cutCap * 4));
        CU(ctx->rCutLen.ensure((size_t)cutCap * 4));
        CU(ctx->rCutSrc.ensure((size_t)cutCap * 4));
        CU(ctx->rCutSeq.ensure((size_t)cutCap * 4));
        CU(ctx->rCutTier.ensure((size_t)cutCap));
        CU(ctx->rValid.ensure((size_t)cutCap));
        ctx->cutCap = cutCap;
    }
    if (nodeCap > ctx->nodeCap) {
        CU(ctx->rNode.ensure((size_t)nodeCap * 4));
        CU(ctx->rEid.ensure((size_t)nodeCap * 4));
        CU(ctx->rStay.ensure((size_t)nodeCap));
        ctx->nodeCap = nodeCap;
    }
    return 0;
}

// K5 on an arbitrary cut view; result in ctx->fin.  srcMul/srcAdd map the view's source index to the global one.
int runDedup(smsgpu_ctx *ctx, const CutView &V, long long totalNodes, bool canonicalise, bool dedup, int srcMul,
             int srcAdd, long long nsrcGlobal, double minEfficacy = 0.0) {
    const long long R = V.n;
    cudaStream_t st = ctx->stream;
    CutListDev &F = ctx->fin;
    if (R == 0) {
        F.ncuts = F.nnodes = 0;
        CU(F.off.ensure(sizeof(long long)));
        CU(cudaMemsetAsync(F.off.p, 0, sizeof(long long), st));
        return 0;
    }
    CU(ctx->wCnode.ensure((size_t)totalNodes * 4));
    CU(ctx->wCeid.ensure((size_t)totalNodes * 4));
    CU(ctx->wCstay.ensure((size_t)totalNodes));
    CU(ctx->wHash.ensure((size_t)R * 8));
    CU(ctx->wRank.ensure((size_t)R * 8));
    CU(ctx->wSlot.ensure((size_t)R * 4));
    CU(ctx->wKeep.ensure((size_t)R));
    CU(ctx->wNcross.ensure((size_t)R * 4));
    CU(ctx->wViol.ensure((size_t)R * 8));
    DedupWork D{};
    D.cnode = ctx->wCnode.as<int32_t>();
    D.ceid = ctx->wCeid.as<int32_t>();
    D.cstay = ctx->wCstay.as<uint8_t>();
    D.hash = ctx->wHash.as<unsigned long long>();
    D.rank = ctx->wRank.as<unsigned long long>();
    D.slot = ctx->wSlot.as<int32_t>();
    D.keep = ctx->wKeep.as<uint8_t>();
    D.ncross = ctx->wNcross.as<int32_t>();
    D.viol = ctx->wViol.as<double>();
    D.collisions = ctx->counters.as<unsigned long long>() + 11;
    canonKernel<<<gridFor(R, 128), 128, 0, st>>>(V, D, ctx->xPtr, canonicalise ? 1 : 0, srcMul, srcAdd, minEfficacy);
    ctx->stats.total_launches += 1;
    if (dedup) {
        const uint64_t tabCap = 1ull << log2ceil((uint64_t)2 * R + 16);
        CU(ctx->wTabKey.ensure(tabCap * 8));
        CU(ctx->wTabRank.ensure(tabCap * 8));
        CU(ctx->wTabIdx.ensure(tabCap * 4));
        D.tabKey = ctx->wTabKey.as<unsigned long long>();
        D.tabRank = ctx->wTabRank.as<unsigned long long>();
        D.tabIdx = ctx->wTabIdx.as<int32_t>();
        D.tabMask = tabCap - 1;
        CU(cudaMemsetAsync(D.tabKey, 0, tabCap * 8, st));
        CU(cudaMemsetAsync(D.tabRank, 0xFF, tabCap * 8, st));
        dedupInsertKernel<<<gridFor(R, 256), 256, 0, st>>>(R, D);
        dedupMarkKernel<<<gridFor(R, 256), 256, 0, st>>>(R, D);
        dedupVerifyKernel<<<gridFor(R, 256), 256, 0, st>>>(V, D);
        ctx->stats.total_launches += 3;
    }
    // order by (global source index, emission index)
    CU(ctx->wCnt.ensure((size_t)(nsrcGlobal + 1) * 4));
    CU(ctx->wCursor.ensure((size_t)(nsrcGlobal + 1) * 4));
    CU(ctx->wBase.ensure((size_t)(nsrcGlobal + 2) * 8));
    CU(ctx->wOrder.ensure((size_t)R * 4));
    CU(ctx->wOutLen.ensure((size_t)R * 4));
    CU(cudaMemsetAsync(ctx->wCnt.p, 0, (size_t)(nsrcGlobal + 1) * 4, st));
    CU(cudaMemsetAsync(ctx->wCursor.p, 0, (size_t)(nsrcGlobal + 1) * 4, st));
    countPerSrcKernel<<<gridFor(R, 256), 256, 0, st>>>(R, D, ctx->wCnt.as<int32_t>());
    ctx->stats.total_launches += 1;
    if (int rc = exclusiveScan(ctx, ctx->wCnt.as<int32_t>(), nsrcGlobal, ctx->wBase.as<long long>())) return rc;
    scatterPerSrcKernel<<<gridFor(R, 256), 256, 0, st>>>(R, D, ctx->wBase.as<long long>(), ctx->wCursor.as<int32_t>(),
                                                        ctx->wOrder.as<int32_t>());
    sortSegmentsKernel<<<gridFor(nsrcGlobal, 256), 256, 0, st>>>(nsrcGlobal, D, ctx->wBase.as<long long>(),
                                                                ctx->wOrder.as<int32_t>(), ctx->wOutLen.as<int32_t>(), V.len);
    ctx->stats.total_launches += 2;
    CU(cudaGetLastError());
    // number of unique cuts -> host (sizes the final arrays)
    unsigned long long *hc = ctx->hCounters.as<unsigned long long>();
    CU(cudaMemcpyAsync(hc + 12, ctx->wBase.as<long long>() + nsrcGlobal, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(hc + 11, ctx->counters.as<unsigned long long>() + 11, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const long long U = (long long)hc[12];
    F.ncuts = U;
    CU(F.off.ensure((size_t)(U + 1) * 8));
    if (int rc = exclusiveScan(ctx, ctx->wOutLen.as<int32_t>(), U, F.off.as<long long>())) return rc;
    CU(cudaMemcpyAsync(hc + 13, F.off.as<long long>() + U, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    F.nnodes = (long long)hc[13];
    CU(F.node.ensure((size_t)std::max<long long>(F.nnodes, 1) * 4));
    CU(F.eid.ensure((size_t)std::max<long long>(F.nnodes, 1) * 4));
    CU(F.stay.ensure((size_t)std::max<long long>(F.nnodes, 1)));
    CU(F.ncross.ensure((size_t)std::max<long long>(U, 1) * 4));
    CU(F.src.ensure((size_t)std::max<long long>(U, 1) * 4));
    CU(F.seq.ensure((size_t)std::max<long long>(U, 1) * 4));
    CU(F.viol.ensure((size_t)std::max<long long>(U, 1) * 8));
    if (U > 0) {
        FinalCuts fc{F.off.as<long long>(), F.node.as<int32_t>(), F.eid.as<int32_t>(), F.stay.as<uint8_t>(),
                     F.ncross.as<int32_t>(), F.src.as<int32_t>(), F.seq.as<int32_t>(), F.viol.as<double>()};
        gatherFinalKernel<<<gridFor(U, 256), 256, 0, st>>>(U, V, D, ctx->wOrder.as<int32_t>(), fc);
        ctx->stats.total_launches += 1;
        CU(cudaGetLastError());
    }
    return 0;
}

int mergeAcrossRanks(smsgpu_ctx *ctx, bool dedup, int localRc); // multi-GPU part below
int registerSmallCycles(smsgpu_ctx *ctx, int k, int64_t T);
int separateLocal(smsgpu_ctx *ctx, const smsgpu_params &P);

} // namespace

extern "C" {

int smsgpu_separate_async(smsgpu_ctx *ctx, const smsgpu_params *params) {
    if (!ctx) return SMSGPU_ERR_ARG;
    if (!ctx->haveInput) return fail(ctx, SMSGPU_ERR_STATE, "smsgpu_separate before smsgpu_set_round_input");
    smsgpu_params P;
    if (params)
        P = *params;
    else
        smsgpu_default_params(&P);
    if (P.max_intermediates < 0 || P.max_intermediates > 16) return fail(ctx, SMSGPU_ERR_ARG, "max_intermediates must be 0..16");
    if (P.mode != SMSGPU_MODE_INDEPENDENT && P.mode != SMSGPU_MODE_VERBATIM) return fail(ctx, SMSGPU_ERR_ARG, "mode must be 0 (independent) or 1 (verbatim)");
    if (!(P.min_efficacy >= 0.0)) return fail(ctx, SMSGPU_ERR_ARG, "min_efficacy must be >= 0");
    // the bucket maps assume radii up to bestInit / 2 <= 0.5; a cycle of length >= kInfDist is never labelled anyway
    if (!(P.best_dist_init > 0.0) || !(P.best_dist_init <= 1.0 + 1e-3))
        return fail(ctx, SMSGPU_ERR_ARG, "best_dist_init must lie in (0, 1.001]");
    if (P.start_tier < 0 || P.start_tier >= kNumTiers) return fail(ctx, SMSGPU_ERR_ARG, "start_tier must be 0..3");
    CU(cudaSetDevice(ctx->device));
    ctx->lastParams = P;
    ctx->fetched = false;
    ctx->haveLabels = ctx->haveSrcStats = false;
    std::memset(&ctx->stats, 0, sizeof ctx->stats);
    // A rank whose local round fails still joins the exchange (with its error flag): its peers must not wait forever.
    const int rcLocal = separateLocal(ctx, P);
    if (ctx->ncclComm && ctx->commSize > 1) {
        const int rcMerge = mergeAcrossRanks(ctx, P.dedup != 0, rcLocal);
        if (rcLocal) return rcLocal;
        if (rcMerge) return rcMerge;
    } else if (rcLocal) {
        return rcLocal;
    }
    ctx->stats.unique_cuts = ctx->fin.ncuts;
    CU(cudaEventRecord(ctx->ev[3], ctx->stream));
    return SMSGPU_OK;
}

} // extern "C"

namespace {

// the local part of a round: K1 -> K2-K4 (tiers) -> K5 over this context's share of the sources; result in ctx->fin
int separateLocal(smsgpu_ctx *ctx, const smsgpu_params &P) {
    cudaStream_t st = ctx->stream;
    const long long nsrc = ctx->nsrcLocal;
    unsigned long long *dc = ctx->counters.as<unsigned long long>();
    unsigned long long *hc = ctx->hCounters.as<unsigned long long>();

    CU(cudaEventRecord(ctx->ev[0], st));
    // ---- K1 ----
    if (ctx->arcsDirty && ctx->m > 0) {
        dcWeightsKernel<<<std::min<unsigned>(gridFor((2 * ctx->m + 3) / 4, 256), (unsigned)ctx->smCount * 16), 256, 0, st>>>(
            2 * ctx->m, ctx->dCol.as<int32_t>(), ctx->dEid.as<int32_t>(), ctx->dRowpack.as<uint32_t>(), ctx->xPtr, ctx->blPtr,
            ctx->dArcs.as<Arc>(), ctx->dArcs12.as<uint32_t>());
        CU(cudaGetLastError());
        ctx->stats.total_launches += 1;
    }
    ctx->arcsDirty = false;
    CU(cudaEventRecord(ctx->ev[1], st));

    // ---- K2-K4: tiers ----
    CU(ctx->srcTier.ensure((size_t)std::max<long long>(nsrc, 1)));
    CU(ctx->ovfA.ensure((size_t)std::max<long long>(nsrc, 1) * 4));
    CU(ctx->ovfB.ensure((size_t)std::max<long long>(nsrc, 1) * 4));
    CU(ctx->srcWork.ensure((size_t)std::max<long long>(nsrc, 1) * 4));
    CU(ctx->srcOrder.ensure((size_t)std::max<long long>(nsrc, 1) * 4));
    CU(ctx->orderBins.ensure((4 * kOrderBins + 1) * sizeof(unsigned int)));
    if (P.export_src_stats) {
        CU(ctx->srcArcs.ensure((size_t)std::max<long long>(nsrc, 1) * 8));
        CU(ctx->srcSettled.ensure((size_t)std::max<long long>(nsrc, 1) * 8));
        CU(cudaMemsetAsync(ctx->srcArcs.p, 0, (size_t)std::max<long long>(nsrc, 1) * 8, st));
        CU(cudaMemsetAsync(ctx->srcSettled.p, 0, (size_t)std::max<long long>(nsrc, 1) * 8, st));
    }
    if (ctx->cutCap == 0) {
        const long long cc = 4 * nsrc + 1024;
        if (int rc = ensureRawArena(ctx, cc, 8 * cc)) return rc;
    }
    if (P.export_labels && ctx->labCap == 0) {
        ctx->labCap = 64 * nsrc + 1024;
        CU(ctx->labSrc.ensure((size_t)ctx->labCap * 4));
        CU(ctx->labNode.ensure((size_t)ctx->labCap * 4));
        CU(ctx->labHops.ensure((size_t)ctx->labCap * 4));
        CU(ctx->labDist.ensure((size_t)ctx->labCap * 8));
    }
    GraphDev G{ctx->n, ctx->m, ctx->dRowptr.as<int64_t>(), ctx->dArcs.as<Arc>(), ctx->dEid.as<int32_t>(), nullptr, nullptr, nullptr};
    long long rawCuts = 0, rawNodes = 0;
    // One sweep over the sources: every tier in turn until each source has finished somewhere; raw walks in the arena.
    // wl0 == nullptr: all local sources, issued by the history of the previous round (independent-source mode's one and
    // only sweep); else the given work list, no history (the sweeps of SMSGPU_MODE_VERBATIM).
    // mode V, windowed: the sweep runs the sources [lo, hi) only; stamps, per-source flags and the walks of the windows in front
    // of it are final and stay (stamps start from `base`, the arena counters from keepCuts / keepNodes)
    struct {
        bool on = false;
        long long lo = 0, hi = 0, keepCuts = 0, keepNodes = 0;
        const int32_t *base = nullptr;
    } vw;
    constexpr int kArenaGrew = 0x7A7A; // (internal) sweep(): the arena was re-allocated under a windowed mode-V round
    auto sweep = [&](const int32_t *wl0, long long nw0, const bool plain) -> int {
    for (int attempt = 0;; ++attempt) {
        if (attempt > 4) return fail(ctx, SMSGPU_ERR_INTERNAL, "raw arena kept overflowing");
        CU(cudaMemsetAsync(dc, 0, 24 * sizeof(unsigned long long), st));
        if (vw.on) {
            hc[22] = (unsigned long long)vw.keepCuts;
            hc[23] = (unsigned long long)vw.keepNodes;
            CU(cudaMemcpyAsync(dc, hc + 22, 16, cudaMemcpyHostToDevice, st));
            CU(cudaMemcpyAsync(G.blStampNext, vw.base, (size_t)ctx->n * 4, cudaMemcpyDeviceToDevice, st));
            CU(cudaMemsetAsync(G.srcEmpty + vw.lo, 0, (size_t)(vw.hi - vw.lo), st));
            CU(cudaMemsetAsync(ctx->srcTier.as<int8_t>() + vw.lo, 0xFF, (size_t)(vw.hi - vw.lo), st));
        } else {
            if (G.blStampNext) { // mode V, one window: this sweep's stamps start from the round's initial blacklist
                initStampsKernel<<<ctx->smCount * 4, 256, 0, st>>>(ctx->n, ctx->blPtr, G.blStampNext);
                CU(cudaGetLastError());
                CU(cudaMemsetAsync(G.srcEmpty, 0, (size_t)std::max<long long>(nsrc, 1), st));
                ctx->stats.total_launches += 1;
            }
            CU(cudaMemsetAsync(ctx->srcTier.p, 0xFF, (size_t)std::max<long long>(nsrc, 1), st));
            CU(cudaMemsetAsync(ctx->srcWork.p, 0, (size_t)std::max<long long>(nsrc, 1) * 4, st)); // skipped sources: no work
        }
        RawOut R{};
        R.node = ctx->rNode.as<int32_t>();
        R.eid = ctx->rEid.as<int32_t>();
        R.stay = ctx->rStay.as<uint8_t>();
        R.cutOff = ctx->rCutOff.as<int32_t>();
        R.cutLen = ctx->rCutLen.as<int32_t>();
        R.cutSrc = ctx->rCutSrc.as<int32_t>();
        R.cutSeq = ctx->rCutSeq.as<int32_t>();
        R.cutTier = ctx->rCutTier.as<int8_t>();
        R.counters = dc;
        R.cutCap = ctx->cutCap;
        R.nodeCap = std::min<long long>(ctx->nodeCap, 0x7FFFFFFFll);
        R.labSrc = ctx->labSrc.as<int32_t>();
        R.labNode = ctx->labNode.as<int32_t>();
        R.labHops = ctx->labHops.as<int32_t>();
        R.labDist = ctx->labDist.as<double>();
        R.labCap = P.export_labels ? ctx->labCap : 0;
        R.srcTier = ctx->srcTier.as<int8_t>();
        R.srcArcs = P.export_src_stats ? ctx->srcArcs.as<long long>() : nullptr;
        R.srcSettled = P.export_src_stats ? ctx->srcSettled.as<long long>() : nullptr;
        R.srcWork = ctx->srcWork.as<uint32_t>();

        SearchParams SP{};
        SP.sources = ctx->explicitSources ? ctx->dSources.as<int32_t>() : nullptr;
        SP.nsrc = nsrc;
        SP.srcBegin = ctx->rank;
        SP.srcStride = ctx->nranks;
        SP.blacklist = ctx->blPtr;
        SP.bestInit = P.best_dist_init;
        SP.maxInter = P.max_intermediates;
        SP.exportLabels = P.export_labels;
        SP.debugFlags = P.debug_flags;
        SP.queueCounter = dc + 8;
        SP.runSlot = 4;
        long long nwork = plain ? nsrc : nw0;
        const int32_t *worklist = plain ? nullptr : wl0;
        if (plain && ctx->useHistory && ctx->haveOrder) worklist = ctx->srcOrder.as<int32_t>(); // longest first (previous round)
        int firstTier = P.start_tier;
        if (ctx->useHistory && ctx->skipTier0 > 0 && firstTier == 0) firstTier = 1;
        int32_t *ovfBufs[2] = {ctx->ovfA.as<int32_t>(), ctx->ovfB.as<int32_t>()};
        long long done[kNumTiers] = {0, 0, 0, 0};
        long long prevRun = 0;
        bool tierRan[kNumTiers] = {false, false, false, false};
        bool lanesRan = false;
        // one search launch (no synchronisation): tier `ti` over `wl[0 .. nw)`, overflows appended to `ovf`
        auto launchTier = [&](int ti, const int32_t *wl, long long nw, int32_t *ovf, int ovfSlot, int queueSlot, int runSlot,
                              cudaStream_t s, int evIdx, int ctasCap) -> int {
            SearchParams Q = SP;
            Q.tier = ti;
            Q.worklist = wl;
            Q.nwork = nw;
            Q.ovfList = ovf;
            Q.ovfCount = dc + ovfSlot;
            Q.queueCounter = dc + queueSlot;
            Q.runSlot = runSlot;
            CU(cudaMemsetAsync(dc + queueSlot, 0, 8, s));
            CU(cudaMemsetAsync(dc + ovfSlot, 0, 8, s));
            if (ctx->tier[ti].caps.nbuckets) {
                TierCfg &tc = ctx->tier[ti];
                if (tc.stampNext + (uint64_t)nw + 2 >= (1ull << 32)) { // stamps exhausted: forget every old entry
                    const size_t stride = tileBytes(tc.caps) + (tc.hotInSmem ? 0 : hotBytes(tc.caps));
                    initTablesKernel<<<ctx->smCount * 8, 256, 0, s>>>(tc.scratch.as<unsigned char>(), stride, tc.ntiles, tc.caps.tabCap);
                    CU(cudaGetLastError());
                    tc.stampNext = 0;
                }
                Q.stampBase = (uint32_t)tc.stampNext;
                Q.hotInSmem = tc.hotInSmem ? 1 : 0;
                tc.stampNext += (uint64_t)nw;
            }
            CU(cudaEventRecord(ctx->tierEv[evIdx], s));
            CU(launchSearch(ctx, ti, G, Q, R, s, ctasCap));
            CU(cudaEventRecord(ctx->tierEv[evIdx + 1], s));
            ctx->stats.search_launches += 1;
            ctx->stats.total_launches += 1;
            return 0;
        };
        // The pairs that needed tier 2 last round (the prefix of the order) start in tier 2 at once, on a second stream,
        // with as many CTAs as they have sources for; tier 1 fills the rest of the GPU with everything else.  Run one
        // after the other, tier 2 is a tail of a few thousand long searches on an otherwise idle GPU.
        long long heavy = 0;
        if (plain && worklist && ctx->heavyPending) {
            CU(cudaEventSynchronize(ctx->heavyEv));
            ctx->heavyPairs = (long long)*ctx->hHeavy.as<unsigned int>();
            ctx->heavyPending = false;
        }
        if (plain && worklist && ctx->overlapHeavy && firstTier == 1 && ctx->heavyPairs > 0 && kNumTiers > 2 && ctx->tier[2].caps.nbuckets) {
            heavy = std::min<long long>(2 * ctx->heavyPairs, nsrc & ~1ll);
            if (int rc = ensureTierScratch(ctx, 1)) return rc;
            if (int rc = ensureTierScratch(ctx, 2)) return rc;
            CU(ctx->ovfC.ensure((size_t)heavy * 4));
            // Tier 1's CTAs are persistent and fill every SM: tier 2 has to be resident BEFORE them.  A cross-stream event
            // wait would let the tier-1 launch overtake it (measured: tier 2 then runs after tier 1), hence the host wait.
            CU(cudaStreamSynchronize(st));
            const int tilesPerCta = 128 / ctx->tileWidth;
            if (int rc = launchTier(2, worklist, heavy, ctx->ovfC.as<int32_t>(), 17, 16, 18, ctx->stream2, 8,
                                    (int)std::min<long long>((heavy + tilesPerCta - 1) / tilesPerCta, 1 << 20)))
                return rc;
            CU(cudaEventRecord(ctx->joinEv, ctx->stream2));
            worklist += heavy;
            nwork -= heavy;
        }
        for (int ti = firstTier; ti < kNumTiers && nwork > 0; ++ti) {
            if (int rc = ensureTierScratch(ctx, ti)) return rc;
            const int which = ti & 1;
            if (heavy > 0 && ti >= 2) CU(cudaStreamWaitEvent(st, ctx->joinEv, 0)); // tier 2's tiles are in use until then
            if (ti == 1 && plain && !G.blStampNext && ctx->tier[1].pairs &&
                (ctx->lanesMode == 1 || (ctx->lanesMode == 2 && nwork >= ctx->lanesMinSources))) {
                // the lanes engine (one source per thread) first; what it gives up -- a capacity hit, a second source that is
                // not the mirror image of the first -- is the pairs engine's work list
                CU(ctx->ovfD.ensure((size_t)std::max<long long>(nsrc, 1) * 4));
                CU(cudaMemsetAsync(dc + 21, 0, 8, st));
                SearchParams LQ = SP;
                LQ.ovfList = ctx->ovfD.as<int32_t>();
                LQ.ovfCount = dc + 21;
                if (int rc = runLanes(ctx, G, LQ, R, worklist, nwork, st)) return rc;
                CU(cudaMemcpyAsync(hc, dc, 24 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
                CU(cudaStreamSynchronize(st));
                lanesRan = true;
                if (std::getenv("SMSGPU_DEBUG_OVF"))
                    std::fprintf(stderr, "[smsgpu] lanes engine: %lld sources in, %llu left for the pairs engine\n", nwork, hc[21]);
                done[1] += (long long)hc[4] - prevRun;
                prevRun = (long long)hc[4];
                worklist = ctx->ovfD.as<int32_t>();
                nwork = (long long)hc[21];
                if (nwork == 0) continue;
            }
            if (int rc = launchTier(ti, worklist, nwork, ovfBufs[which], 9 + which, 8, 4, st, 2 * ti, 0)) return rc;
            tierRan[ti] = true;
            CU(cudaMemcpyAsync(hc, dc, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            if (std::getenv("SMSGPU_DEBUG_OVF"))
                std::fprintf(stderr, "[smsgpu] tier %d: %lld sources in, %llu overflowed (table %llu, bucket %llu at mean bucket %.0f, history %llu)\n", ti,
                             nwork, hc[9 + which], hc[12], hc[13], hc[13] ? (double)hc[15] / (double)hc[13] : 0.0, hc[14]);
            done[ti] += (long long)hc[4] - prevRun;
            prevRun = (long long)hc[4];
            nwork = (long long)hc[9 + which];
            worklist = ovfBufs[which];
            if (ti == 0 && nwork > 1 && kNumTiers > 1 && ctx->tier[1].pairs && ctx->pairsMirror) {
                // tier 0's overflow list is in completion order; the pairs engine wants the two sides of a vertex adjacent
                CU(ctx->ovfC.ensure((size_t)(nsrc + 1) * 4));
                CU(cudaMemsetAsync(dc + 19, 0, 16, st));
                pairCompactKernel<<<std::min<unsigned>(gridFor((nsrc + 1) / 2, 256), (unsigned)ctx->smCount * 8), 256, 0, st>>>(
                    nsrc, ctx->srcTier.as<int8_t>(), ctx->ovfC.as<int32_t>(), ovfBufs[which], dc + 19);
                CU(cudaGetLastError());
                ctx->stats.total_launches += 1;
                CU(cudaMemcpyAsync(hc + 19, dc + 19, 16, cudaMemcpyDeviceToHost, st));
                CU(cudaStreamSynchronize(st));
                const long long both = (long long)hc[19], singles = (long long)hc[20];
                if (singles > 0)
                    CU(cudaMemcpyAsync(ctx->ovfC.as<int32_t>() + 2 * both, ovfBufs[which], (size_t)singles * 4, cudaMemcpyDeviceToDevice, st));
                nwork = 2 * both + singles; // the overflowed sources plus the skipped ones (skipped again by the next tier)
                worklist = ctx->ovfC.as<int32_t>();
            }
        }
        long long heavyRun = 0;
        bool heavyRan = false;
        if (heavy > 0 && nwork == 0) {
            CU(cudaStreamWaitEvent(st, ctx->joinEv, 0));
            CU(cudaMemcpyAsync(hc, dc, 24 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            heavyRan = true;
            heavyRun = (long long)hc[18];
            done[2] += heavyRun;
            nwork = (long long)hc[17];
            if (std::getenv("SMSGPU_DEBUG_OVF"))
                std::fprintf(stderr, "[smsgpu] tier 2 beside tier 1: %lld sources in, %lld overflowed\n", heavy, nwork);
            if (nwork > 0) { // they exceeded tier 2 as well: the dense tier
                if (int rc = ensureTierScratch(ctx, 3)) return rc;
                if (int rc = launchTier(3, ctx->ovfC.as<int32_t>(), nwork, ovfBufs[0], 9, 8, 4, st, 6, 0)) return rc;
                tierRan[3] = true;
                CU(cudaMemcpyAsync(hc, dc, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
                CU(cudaStreamSynchronize(st));
                done[3] += (long long)hc[4] - prevRun;
                prevRun = (long long)hc[4];
                nwork = (long long)hc[9];
            }
        } else if (heavy > 0) {
            CU(cudaStreamSynchronize(ctx->stream2)); // failing below: leave no launch behind
        }
        if (nwork > 0)
            return fail(ctx, SMSGPU_ERR_INTERNAL, "%lld sources exceeded the capacities of the dense tier", nwork);
        rawCuts = (long long)hc[0];
        rawNodes = (long long)hc[1];
        const long long nlab = (long long)hc[5];
        bool redo = false;
        if (rawCuts > ctx->cutCap || rawNodes > std::min<long long>(ctx->nodeCap, 0x7FFFFFFFll)) {
            if (rawNodes > 0x7FFFFFFFll) return fail(ctx, SMSGPU_ERR_INTERNAL, "more than 2^31 walk nodes in one round");
            if (int rc = ensureRawArena(ctx, std::max(ctx->cutCap, rawCuts + rawCuts / 4), std::max(ctx->nodeCap, rawNodes + rawNodes / 4)))
                return rc;
            redo = true;
        }
        if (P.export_labels && nlab > ctx->labCap) {
            ctx->labCap = nlab + nlab / 4;
            CU(ctx->labSrc.ensure((size_t)ctx->labCap * 4));
            CU(ctx->labNode.ensure((size_t)ctx->labCap * 4));
            CU(ctx->labHops.ensure((size_t)ctx->labCap * 4));
            CU(ctx->labDist.ensure((size_t)ctx->labCap * 8));
            redo = true;
        }
        if (redo && vw.on) return kArenaGrew;
        if (redo) continue;
        if (plain && ctx->useHistory && nsrc > 0) {
            // tier 0 (shared memory) is skipped for the next 15 rounds when it completed fewer than 1/4 of its sources
            if (tierRan[0])
                ctx->skipTier0 = (done[0] * 4 < nsrc) ? 15 : 0;
            else if (ctx->skipTier0 > 0)
                ctx->skipTier0--;
            unsigned int *bins = ctx->orderBins.as<unsigned int>();
            CU(cudaMemsetAsync(bins, 0, (4 * kOrderBins + 1) * sizeof(unsigned int), st));
            const unsigned g = std::min<unsigned>(gridFor(nsrc, 256), (unsigned)ctx->smCount * 8);
            orderHistKernel<<<g, 256, 0, st>>>(nsrc, ctx->srcWork.as<uint32_t>(), ctx->srcTier.as<int8_t>(), bins);
            orderScanKernel<<<1, kOrderBins, 0, st>>>(bins);
            orderScatterKernel<<<g, 256, 0, st>>>(nsrc, ctx->srcWork.as<uint32_t>(), ctx->srcTier.as<int8_t>(), bins,
                                                 ctx->srcOrder.as<int32_t>());
            CU(cudaGetLastError());
            ctx->stats.total_launches += 3;
            // read when the next round needs it (heavyEv): no host wait here
            CU(cudaMemcpyAsync(ctx->hHeavy.p, bins + 4 * kOrderBins, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
            CU(cudaEventRecord(ctx->heavyEv, st));
            ctx->heavyPending = true;
            ctx->haveOrder = true;
        }
        ctx->stats.sources_run = (long long)hc[4] + heavyRun;
        ctx->stats.arcs_scanned += (long long)hc[2]; // summed over the sweeps of mode V
        ctx->stats.settled += (long long)hc[3];
        ctx->stats.mirrored_sources = (long long)hc[6];
        ctx->stats.mirrored_arcs = (long long)hc[7];
        ctx->stats.raw_walks = rawCuts;
        ctx->nLabels = nlab;
        for (int k = 0; k < kNumTiers; ++k) {
            ctx->stats.tier_sources[k] = done[k];
            float tms = 0;
            ctx->stats.ms_tier[k] = 0.0;
            if (tierRan[k] && cudaEventElapsedTime(&tms, ctx->tierEv[2 * k], ctx->tierEv[2 * k + 1]) == cudaSuccess) ctx->stats.ms_tier[k] = tms;
        }
        float lms = 0;
        if (lanesRan && cudaEventElapsedTime(&lms, ctx->lanesT0, ctx->lanesT1) == cudaSuccess) ctx->stats.ms_tier[1] += lms;
        float hms = 0; // the overlapped launch ran beside tier 1: the tier times no longer add up to the search time
        if (heavyRan && cudaEventElapsedTime(&hms, ctx->tierEv[8], ctx->tierEv[9]) == cudaSuccess) ctx->stats.ms_tier[2] += hms;
        break;
    }
    return 0;
    }; // sweep

    if (P.mode == SMSGPU_MODE_INDEPENDENT) {
        if (int rc = sweep(nullptr, nsrc, true)) return rc;
    } else {
        // ---- SMSGPU_MODE_VERBATIM: the sequential sweep as the fixed point of independent sweeps over time stamps --------
        // (include/smsgpu.h; CPU statement of the same scheme: oracle/ocw_oracle.hpp stampBegin...)
        if (ctx->nranks != 1) return fail(ctx, SMSGPU_ERR_ARG, "SMSGPU_MODE_VERBATIM runs on one GPU (the stamps are positions in the whole source list)");
        const size_t nb = (size_t)std::max<long long>(nsrc, 1);
        CU(ctx->vStampA.ensure((size_t)ctx->n * 4));
        CU(ctx->vStampB.ensure((size_t)ctx->n * 4));
        CU(ctx->vEmpty.ensure(nb));
        CU(ctx->vEmptyB0.ensure(nb));
        CU(ctx->vActive.ensure(nb * 4));
        CU(ctx->vFlags.ensure(4 * sizeof(unsigned int)));
        CU(ctx->vStampC.ensure((size_t)ctx->n * 4));
        CU(ctx->vIota.ensure(nb * 4));
        int32_t *cur = ctx->vStampA.as<int32_t>(), *next = ctx->vStampB.as<int32_t>(), *base = ctx->vStampC.as<int32_t>();
        unsigned int *flags = ctx->vFlags.as<unsigned int>(); // [0] flood changed, [1] stamps differ, [2] active count
        unsigned int hf[4] = {0, 0, 0, 0};
        const unsigned gv = (unsigned)ctx->smCount * 8;
        iotaKernel<<<gv, 256, 0, st>>>(nsrc, ctx->vIota.as<int32_t>());
        CU(cudaGetLastError());
        const bool savedMirror = ctx->pairsMirror;
        ctx->pairsMirror = false; // the two sides of a vertex sit at different times: each is searched
        const bool shortcutOn = !(std::getenv("SMSGPU_VERBATIM_FULL") && std::getenv("SMSGPU_VERBATIM_FULL")[0] == '1');
        const int32_t *srcList = ctx->explicitSources ? ctx->dSources.as<int32_t>() : nullptr;
        // WINDOWS.  A source only reads stamps of EARLIER sources, so the list can be brought to its fixed point window by
        // window, in order: the stamps of the finished windows are final (`base`), the sweeps of a window run its sources only.
        // The sweeps a window needs are the length of the dependency chains INSIDE it, so short windows need few -- the work
        // falls from (chain length of the whole list) x n to about (chain length of a window) x n.  But every sweep has a fixed
        // cost (~10 launches, 4 host waits), so windows only pay when the chains are long or the sweeps heavy (measured: torus
        // 100x100, LP iterates, 111 sweeps: 3.3 s -> 0.85 s per round with 10 windows; the same torus with uniform x*, 14 sweeps:
        // 3.2 -> 13.7 ms).  Rule: one window, unless the list is long (> 200 000 sources: n / 16) or the previous round of this
        // context needed >= 48 sweeps with one window (then n / 10, at least 1024).  SMSGPU_VERBATIM_WINDOW overrides, 0 = one.
        long long W = ctx->vWindow > 0 ? ctx->vWindow : (nsrc > 200000 ? ((nsrc / 16 + 1) & ~1ll) : nsrc);
        if (const char *ew = std::getenv("SMSGPU_VERBATIM_WINDOW")) W = std::atoll(ew);
        if (W <= 0 || W > nsrc) W = std::max<long long>(nsrc, 1);
        int rcV = 0;
        long long sweepsTotal = 0;
        auto cuv = [&](cudaError_t e) { return e == cudaSuccess ? 0 : fail(ctx, SMSGPU_ERR_CUDA, "mode V: %s", cudaGetErrorString(e)); };
        for (int restart = 0; restart < 4 && !rcV; ++restart) { // (restarted from the first window when the walk arena had to grow)
        bool again = false;
        initStampsKernel<<<ctx->smCount * 4, 256, 0, st>>>(ctx->n, ctx->blPtr, base);
        if ((rcV = cuv(cudaGetLastError()))) break;
        vw.keepCuts = vw.keepNodes = 0;
        for (long long lo = 0; lo < nsrc && !rcV && !again; lo += W) {
            const long long hi = std::min(nsrc, lo + W);
            vw.on = W < nsrc;
            vw.lo = lo;
            vw.hi = hi;
            vw.base = base;
            if ((rcV = cuv(cudaMemcpyAsync(cur, base, (size_t)ctx->n * 4, cudaMemcpyDeviceToDevice, st)))) break;
            bool shortcut = shortcutOn, verifying = false;
            const int32_t *active = nullptr;
            long long nactive = hi - lo;
            for (long long sweepNo = 1;; ++sweepNo) {
                if (sweepNo > (hi - lo) + 2) {
                    rcV = fail(ctx, SMSGPU_ERR_INTERNAL, "the stamps did not reach a fixed point in %lld sweeps", sweepNo - 1);
                    break;
                }
                const bool full = active == nullptr;
                G.blStamp = cur;
                G.blStampNext = next;
                G.srcEmpty = ctx->vEmpty.as<uint8_t>();
                rcV = sweep(full ? ctx->vIota.as<int32_t>() + lo : active, full ? hi - lo : nactive, false);
                if (rcV == kArenaGrew) { // the arena was re-allocated: the walks of the finished windows are gone
                    rcV = 0;
                    again = true;
                    break;
                }
                if (rcV) break;
                ++sweepsTotal;
                ctx->stats.verbatim_sweeps = sweepsTotal;
                if (!full) {
                    reseedEmptySourcesKernel<<<gv, 256, 0, st>>>(G, lo, hi, srcList, ctx->vEmptyB0.as<uint8_t>());
                    ctx->stats.total_launches += 1;
                }
                for (long long pass = 0;; ++pass) { // flood of the empty sources' stamps over integral edges
                    if ((rcV = cuv(cudaMemsetAsync(flags, 0, 8, st)))) break;
                    floodStampsKernel<<<gv, 256, 0, st>>>(G, flags);
                    ctx->stats.total_launches += 1;
                    if ((rcV = cuv(cudaMemcpyAsync(hf, flags, 4, cudaMemcpyDeviceToHost, st)))) break;
                    if ((rcV = cuv(cudaStreamSynchronize(st)))) break;
                    if (!hf[0]) break;
                    if (pass > ctx->n + 2) {
                        rcV = fail(ctx, SMSGPU_ERR_INTERNAL, "stamp flood did not settle");
                        break;
                    }
                }
                if (rcV) break;
                compareStampsKernel<<<gv, 256, 0, st>>>(ctx->n, cur, next, flags + 1);
                ctx->stats.total_launches += 1;
                if (sweepNo == 1 && shortcut) { // sources that found nothing against the stamps in front of the window never find anything
                    if ((rcV = cuv(cudaMemcpyAsync(ctx->vEmptyB0.as<uint8_t>() + lo, ctx->vEmpty.as<uint8_t>() + lo, (size_t)(hi - lo),
                                                   cudaMemcpyDeviceToDevice, st))))
                        break;
                    if ((rcV = cuv(cudaMemsetAsync(flags + 2, 0, 4, st)))) break;
                    activeSourcesKernel<<<gv, 256, 0, st>>>(lo, hi, ctx->vEmptyB0.as<uint8_t>(), ctx->vActive.as<int32_t>(), flags + 2);
                    ctx->stats.total_launches += 1;
                }
                if ((rcV = cuv(cudaMemcpyAsync(hf, flags, 12, cudaMemcpyDeviceToHost, st)))) break;
                if ((rcV = cuv(cudaStreamSynchronize(st)))) break;
                if (sweepNo == 1 && shortcut) ctx->vActiveCount = (long long)hf[2];
                std::swap(cur, next);
                if (std::getenv("SMSGPU_DEBUG_OVF"))
                    std::fprintf(stderr, "[smsgpu] mode V window [%lld, %lld) sweep %lld (%s, %lld sources): stamps %s\n", lo, hi, sweepNo,
                                 full ? "full" : "active set", full ? hi - lo : nactive, hf[1] ? "changed" : "reproduced");
                if (!hf[1]) {
                    if (full) break; // a full sweep of the window reproduced the stamps it read: its walks are the sequential sweep's
                    verifying = true; // the active set has converged: one full sweep has to confirm it
                    active = nullptr;
                    continue;
                }
                if (verifying) shortcut = false; // the confirmation failed (never observed): full sweeps from here on
                if (shortcut && !verifying) {
                    active = ctx->vActive.as<int32_t>();
                    nactive = ctx->vActiveCount;
                } else {
                    active = nullptr;
                }
            }
            if (rcV || again) break;
            // the window is final: its stamps join the base, its walks stay in the arena
            std::swap(base, cur);
            vw.keepCuts = rawCuts;
            vw.keepNodes = rawNodes;
        }
        if (!again) break;
        }
        vw.on = false;
        if (!rcV && W >= nsrc && sweepsTotal >= 48 && nsrc >= 2048) ctx->vWindow = std::max<long long>(1024, (nsrc / 10 + 1) & ~1ll);
        ctx->pairsMirror = savedMirror;
        if (rcV) return rcV;
    }
    CU(cudaEventRecord(ctx->ev[2], st));

    // ---- K5 ----
    if (rawCuts > 0) {
        validKernel<<<gridFor(rawCuts, 256), 256, 0, st>>>(rawCuts, ctx->rCutTier.as<int8_t>(), ctx->rCutSrc.as<int32_t>(),
                                                           ctx->srcTier.as<int8_t>(), ctx->rValid.as<uint8_t>());
        ctx->stats.total_launches += 1;
    }
    CutView V{ctx->rNode.as<int32_t>(), ctx->rEid.as<int32_t>(), ctx->rStay.as<uint8_t>(), ctx->rCutOff.as<int32_t>(),
              ctx->rCutLen.as<int32_t>(), ctx->rCutSrc.as<int32_t>(), ctx->rCutSeq.as<int32_t>(), ctx->rValid.as<uint8_t>(),
              rawCuts};
#ifdef SMSGPU_RUNS_CLOCKS
    {
        unsigned long long h[24];
        if (cudaMemcpyFromSymbol(h, g_runClk, sizeof h) == cudaSuccess) {
            std::fprintf(stderr, "[smsgpu] runs engine, CTA 0, cycles: offers + grouping %llu | apply %llu | record logic %llu | rollback %llu || pool minimum %llu | "
                                 "collect %llu | member sort %llu | chunk set-up %llu | rest of the tail %llu | table reset %llu || selection %llu | buildWalk %llu | "
                                 "simple cycles + chords %llu | walks %llu of mean length %.1f, max %llu || extractCycles x %llu: positions %llu | prefix sums %llu | "
                                 "chord test %llu | sort %llu | greedy + emission %llu\n",
                         h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7], h[8], h[9], h[10], h[11], h[12], h[14], h[14] ? (double)h[13] / (double)h[14] : 0.0, h[15],
                         h[21], h[16], h[17], h[18], h[19], h[20]);
            std::memset(h, 0, sizeof h);
            cudaMemcpyToSymbol(g_runClk, h, sizeof h);
        }
    }
#endif
    return runDedup(ctx, V, rawNodes, P.dedup != 0, P.dedup != 0, ctx->nranks, ctx->rank, ctx->nsrcGlobal, P.min_efficacy);
}

} // namespace

extern "C" {

int smsgpu_fetch_cuts(smsgpu_ctx *ctx) {
    if (!ctx) return SMSGPU_ERR_ARG;
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    CutListDev &F = ctx->fin;
    CutListHost &H = ctx->host;
    const long long U = F.ncuts, N = F.nnodes;
    CU(H.off.ensure((size_t)(U + 1) * 8));
    CU(H.node.ensure((size_t)std::max<long long>(N, 1) * 4));
    CU(H.eid.ensure((size_t)std::max<long long>(N, 1) * 4));
    CU(H.stay.ensure((size_t)std::max<long long>(N, 1)));
    CU(H.ncross.ensure((size_t)std::max<long long>(U, 1) * 4));
    CU(H.src.ensure((size_t)std::max<long long>(U, 1) * 4));
    CU(H.viol.ensure((size_t)std::max<long long>(U, 1) * 8));
    CU(cudaEventRecord(ctx->ev[4], st));
    CU(cudaMemcpyAsync(H.off.p, F.off.p, (size_t)(U + 1) * 8, cudaMemcpyDeviceToHost, st));
    if (U > 0) {
        CU(cudaMemcpyAsync(H.node.p, F.node.p, (size_t)N * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.eid.p, F.eid.p, (size_t)N * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.stay.p, F.stay.p, (size_t)N, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.ncross.p, F.ncross.p, (size_t)U * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.src.p, F.src.p, (size_t)U * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.viol.p, F.viol.p, (size_t)U * 8, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaEventRecord(ctx->ev[5], st));
    CU(cudaStreamSynchronize(st));
    H.ncuts = U;
    H.nnodes = N;
    ctx->fetched = true;
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
    ctx->stats.ms_weights = ms;
    cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]);
    ctx->stats.ms_search = ms;
    cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]);
    ctx->stats.ms_dedup = ms;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[3]);
    ctx->stats.ms_total_device = ms;
    cudaEventElapsedTime(&ms, ctx->ev[4], ctx->ev[5]);
    ctx->stats.ms_d2h = ms;
    if (cudaEventElapsedTime(&ms, ctx->ev[6], ctx->ev[7]) == cudaSuccess) ctx->stats.ms_h2d = ms;
    (void)cudaGetLastError();
    return SMSGPU_OK;
}

int smsgpu_separate(smsgpu_ctx *ctx, const smsgpu_params *params) {
    if (int rc = smsgpu_separate_async(ctx, params)) return rc;
    return smsgpu_fetch_cuts(ctx);
}

int smsgpu_get_cuts(smsgpu_ctx *ctx, smsgpu_cuts *out) {
    if (!ctx || !out) return SMSGPU_ERR_ARG;
    if (!ctx->fetched) return fail(ctx, SMSGPU_ERR_STATE, "no fetched result: call smsgpu_separate or smsgpu_fetch_cuts first");
    const CutListHost &H = ctx->host;
    out->ncuts = H.ncuts;
    out->nnodes = H.nnodes;
    out->off = H.off.as<int64_t>();
    out->node = H.node.as<int32_t>();
    out->eid = H.eid.as<int32_t>();
    out->is_stay = H.stay.as<uint8_t>();
    out->n_cross = H.ncross.as<int32_t>();
    out->src = H.src.as<int32_t>();
    out->violation = H.viol.as<double>();
    return SMSGPU_OK;
}

int smsgpu_get_stats(smsgpu_ctx *ctx, smsgpu_stats *out) {
    if (!ctx || !out) return SMSGPU_ERR_ARG;
    *out = ctx->stats;
    return SMSGPU_OK;
}

} // extern "C"

// =====================================================================================================================
// parity instrumentation
// =====================================================================================================================
extern "C" {

int smsgpu_get_labels(smsgpu_ctx *ctx, int64_t *nlabels, const int64_t **label_off, const int32_t **label_node,
                      const double **label_dist, const int32_t **label_hops) {
    if (!ctx || !nlabels || !label_off || !label_node || !label_dist || !label_hops) return SMSGPU_ERR_ARG;
    if (!ctx->lastParams.export_labels) return fail(ctx, SMSGPU_ERR_STATE, "last round ran without export_labels");
    CU(cudaSetDevice(ctx->device));
    const long long L = ctx->nLabels, nsrc = ctx->nsrcLocal;
    if (!ctx->haveLabels) {
        std::vector<int32_t> src((size_t)L), node((size_t)L), hops((size_t)L);
        std::vector<double> dist((size_t)L);
        if (L > 0) {
            CU(cudaMemcpy(src.data(), ctx->labSrc.p, (size_t)L * 4, cudaMemcpyDeviceToHost));
            CU(cudaMemcpy(node.data(), ctx->labNode.p, (size_t)L * 4, cudaMemcpyDeviceToHost));
            CU(cudaMemcpy(hops.data(), ctx->labHops.p, (size_t)L * 4, cudaMemcpyDeviceToHost));
            CU(cudaMemcpy(dist.data(), ctx->labDist.p, (size_t)L * 8, cudaMemcpyDeviceToHost));
        }
        std::vector<long long> order((size_t)L);
        for (long long i = 0; i < L; ++i) order[i] = i;
        std::sort(order.begin(), order.end(), [&](long long a, long long b) {
            return src[a] != src[b] ? src[a] < src[b] : node[a] < node[b];
        });
        CU(ctx->hLabOff.ensure((size_t)(nsrc + 1) * 8));
        CU(ctx->hLabNode.ensure((size_t)std::max<long long>(L, 1) * 4));
        CU(ctx->hLabDist.ensure((size_t)std::max<long long>(L, 1) * 8));
        CU(ctx->hLabHops.ensure((size_t)std::max<long long>(L, 1) * 4));
        int64_t *off = ctx->hLabOff.as<int64_t>();
        std::fill(off, off + nsrc + 1, 0);
        for (long long i = 0; i < L; ++i) {
            const long long o = order[i];
            off[src[o] + 1]++;
            ctx->hLabNode.as<int32_t>()[i] = node[o];
            ctx->hLabDist.as<double>()[i] = dist[o];
            ctx->hLabHops.as<int32_t>()[i] = hops[o];
        }
        for (long long s = 0; s < nsrc; ++s) off[s + 1] += off[s];
        ctx->haveLabels = true;
    }
    *nlabels = L;
    *label_off = ctx->hLabOff.as<int64_t>();
    *label_node = ctx->hLabNode.as<int32_t>();
    *label_dist = ctx->hLabDist.as<double>();
    *label_hops = ctx->hLabHops.as<int32_t>();
    return SMSGPU_OK;
}

int smsgpu_get_src_stats(smsgpu_ctx *ctx, int64_t *nsrcOut, const int64_t **arcs, const int64_t **settled,
                         const int32_t **tier) {
    if (!ctx || !nsrcOut || !arcs || !settled || !tier) return SMSGPU_ERR_ARG;
    if (!ctx->lastParams.export_src_stats) return fail(ctx, SMSGPU_ERR_STATE, "last round ran without export_src_stats");
    CU(cudaSetDevice(ctx->device));
    const long long nsrc = ctx->nsrcLocal;
    if (!ctx->haveSrcStats) {
        CU(ctx->hSrcArcs.ensure((size_t)std::max<long long>(nsrc, 1) * 8));
        CU(ctx->hSrcSettled.ensure((size_t)std::max<long long>(nsrc, 1) * 8));
        CU(ctx->hSrcTier.ensure((size_t)std::max<long long>(nsrc, 1) * 4));
        std::vector<int8_t> t8((size_t)nsrc);
        if (nsrc > 0) {
            CU(cudaMemcpy(ctx->hSrcArcs.p, ctx->srcArcs.p, (size_t)nsrc * 8, cudaMemcpyDeviceToHost));
            CU(cudaMemcpy(ctx->hSrcSettled.p, ctx->srcSettled.p, (size_t)nsrc * 8, cudaMemcpyDeviceToHost));
            CU(cudaMemcpy(t8.data(), ctx->srcTier.p, (size_t)nsrc, cudaMemcpyDeviceToHost));
        }
        for (long long s = 0; s < nsrc; ++s) ctx->hSrcTier.as<int32_t>()[s] = t8[s];
        ctx->haveSrcStats = true;
    }
    *nsrcOut = nsrc;
    *arcs = ctx->hSrcArcs.as<int64_t>();
    *settled = ctx->hSrcSettled.as<int64_t>();
    *tier = ctx->hSrcTier.as<int32_t>();
    return SMSGPU_OK;
}

// =====================================================================================================================
// K6: small-cycle scan
// =====================================================================================================================
int smsgpu_small_cycles_set(smsgpu_ctx *ctx, int k, int64_t T, const int32_t *cycles) {
    if (!ctx) return SMSGPU_ERR_ARG;
    if ((k != 3 && k != 4) || T < 0 || (T > 0 && !cycles)) return fail(ctx, SMSGPU_ERR_ARG, "k must be 3 or 4, cycles non-null");
    if (T * k >= (1ll << 31)) return fail(ctx, SMSGPU_ERR_ARG, "cycle list too long");
    CU(cudaSetDevice(ctx->device));
    const int s = k - 3;
    for (int64_t i = 0; i < T * k; ++i)
        if (cycles[i] < 0 || cycles[i] >= ctx->n) return fail(ctx, SMSGPU_ERR_ARG, "cycle vertex out of range");
    CU(ctx->scNodes[s].ensure((size_t)std::max<int64_t>(T * k, 4) * 4));
    if (T > 0) CU(cudaMemcpyAsync(ctx->scNodes[s].p, cycles, (size_t)T * k * 4, cudaMemcpyHostToDevice, ctx->stream));
    return registerSmallCycles(ctx, k, T);
}

} // extern "C"

namespace {
// the list of k-cycles sits in ctx->scNodes[k - 3] (device): resolve its edge ids once, size the scan's buffers
int registerSmallCycles(smsgpu_ctx *ctx, int k, int64_t T) {
    const int s = k - 3;
    ctx->scCount[s] = T;
    CU(ctx->scNodes[s].ensure((size_t)std::max<int64_t>(T * k, 4) * 4));
    CU(ctx->scEids[s].ensure((size_t)std::max<int64_t>(T * k, 4) * 4));
    if (T == 0) return SMSGPU_OK;
    unsigned long long *dc = ctx->counters.as<unsigned long long>();
    CU(cudaMemsetAsync(dc + 14, 0, 8, ctx->stream));
    int *missing = reinterpret_cast<int *>(dc + 14);
    if (k == 3)
        resolveCycleEidsKernel<3><<<gridFor(T, 256), 256, 0, ctx->stream>>>(T, ctx->scNodes[s].as<int32_t>(), ctx->dRowptr.as<long long>(),
                                                                           ctx->dCol.as<int32_t>(), ctx->dEid.as<int32_t>(),
                                                                           ctx->scEids[s].as<int32_t>(), missing);
    else
        resolveCycleEidsKernel<4><<<gridFor(T, 256), 256, 0, ctx->stream>>>(T, ctx->scNodes[s].as<int32_t>(), ctx->dRowptr.as<long long>(),
                                                                           ctx->dCol.as<int32_t>(), ctx->dEid.as<int32_t>(),
                                                                           ctx->scEids[s].as<int32_t>(), missing);
    CU(cudaGetLastError());
    int hm = 0;
    CU(cudaMemcpyAsync(&hm, missing, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (hm) {
        ctx->scCount[s] = 0;
        return fail(ctx, SMSGPU_ERR_ARG, "%d cycle edges are not edges of the graph", hm);
    }
    return SMSGPU_OK;
}
} // namespace

extern "C" {

// SmallChordlessCycles::computeSmallCCs (src/graph/small_ccs.cpp:18-43) on the device: triangles, and chordless
// four-cycles when want_holes, in the reference's list order; both lists are registered for the scan.
int smsgpu_small_cycles_enumerate(smsgpu_ctx *ctx, int want_holes, int64_t *ntriangles, int64_t *nholes) {
    if (!ctx) return SMSGPU_ERR_ARG;
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const long long narcs = 2 * ctx->m;
    if (ntriangles) *ntriangles = 0;
    if (nholes) *nholes = 0;
    if (narcs == 0) {
        if (int rc = registerSmallCycles(ctx, 3, 0)) return rc;
        return want_holes ? registerSmallCycles(ctx, 4, 0) : SMSGPU_OK;
    }
    if (!ctx->enumReady) { // owner vertex of every arc + per-row sorted adjacency (hasEdge = binary search), built once
        std::vector<int32_t> src((size_t)narcs), sorted(ctx->hCol);
        for (long long u = 0; u < ctx->n; ++u) {
            std::fill(src.begin() + ctx->hRowptr[u], src.begin() + ctx->hRowptr[u + 1], (int32_t)u);
            std::sort(sorted.begin() + ctx->hRowptr[u], sorted.begin() + ctx->hRowptr[u + 1]);
        }
        CU(ctx->dArcSrc.ensure((size_t)narcs * 4));
        CU(ctx->dColSorted.ensure((size_t)narcs * 4));
        CU(cudaMemcpy(ctx->dArcSrc.p, src.data(), (size_t)narcs * 4, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(ctx->dColSorted.p, sorted.data(), (size_t)narcs * 4, cudaMemcpyHostToDevice));
        ctx->enumReady = true;
    }
    CU(ctx->wCnt.ensure((size_t)(narcs + 1) * 4));
    CU(ctx->wCursor.ensure((size_t)(narcs + 1) * 4));
    CU(ctx->wBase.ensure((size_t)(narcs + 2) * 8));
    CU(ctx->eHoleBase.ensure((size_t)(narcs + 2) * 8));
    const unsigned grid = (unsigned)std::min<long long>((narcs + 7) / 8, (long long)ctx->smCount * 32);
    int32_t *triCnt = ctx->wCnt.as<int32_t>(), *holeCnt = ctx->wCursor.as<int32_t>();
    long long *triBase = ctx->wBase.as<long long>(), *holeBase = ctx->eHoleBase.as<long long>();
    enumerateSmallCyclesKernel<false><<<grid, 256, 0, st>>>(narcs, ctx->dArcSrc.as<int32_t>(), ctx->dRowptr.as<long long>(),
                                                            ctx->dCol.as<int32_t>(), ctx->dColSorted.as<int32_t>(), triCnt, holeCnt,
                                                            nullptr, nullptr, nullptr, nullptr, want_holes ? 1 : 0);
    CU(cudaGetLastError());
    if (int rc = exclusiveScan(ctx, triCnt, narcs, triBase)) return rc;
    if (int rc = exclusiveScan(ctx, holeCnt, narcs, holeBase)) return rc;
    long long tot[2] = {0, 0};
    CU(cudaMemcpyAsync(&tot[0], triBase + narcs, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&tot[1], holeBase + narcs, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (tot[0] * 3 >= (1ll << 31) || tot[1] * 4 >= (1ll << 31)) return fail(ctx, SMSGPU_ERR_ARG, "more than 2^31 cycle vertices (%lld triangles, %lld holes)", tot[0], tot[1]);
    CU(ctx->scNodes[0].ensure((size_t)std::max<long long>(tot[0] * 3, 4) * 4));
    CU(ctx->scNodes[1].ensure((size_t)std::max<long long>(tot[1] * 4, 4) * 4));
    if (tot[0] + tot[1] > 0) {
        enumerateSmallCyclesKernel<true><<<grid, 256, 0, st>>>(narcs, ctx->dArcSrc.as<int32_t>(), ctx->dRowptr.as<long long>(),
                                                               ctx->dCol.as<int32_t>(), ctx->dColSorted.as<int32_t>(), nullptr, nullptr,
                                                               triBase, holeBase, ctx->scNodes[0].as<int32_t>(),
                                                               ctx->scNodes[1].as<int32_t>(), want_holes ? 1 : 0);
        CU(cudaGetLastError());
    }
    if (int rc = registerSmallCycles(ctx, 3, tot[0])) return rc;
    if (want_holes)
        if (int rc = registerSmallCycles(ctx, 4, tot[1])) return rc;
    if (ntriangles) *ntriangles = tot[0];
    if (nholes) *nholes = want_holes ? tot[1] : 0;
    return SMSGPU_OK;
}

// host copy of the registered list of k-cycles (vertex ids, k per cycle), e.g. after smsgpu_small_cycles_enumerate
int smsgpu_small_cycles_get(smsgpu_ctx *ctx, int k, int64_t *T, const int32_t **cycles) {
    if (!ctx || !T || !cycles) return SMSGPU_ERR_ARG;
    if (k != 3 && k != 4) return fail(ctx, SMSGPU_ERR_ARG, "k must be 3 or 4");
    CU(cudaSetDevice(ctx->device));
    const int s = k - 3;
    const long long cnt = ctx->scCount[s];
    CU(ctx->scListHost[s].ensure((size_t)std::max<long long>(cnt * k, 1) * 4));
    if (cnt > 0) {
        CU(cudaMemcpyAsync(ctx->scListHost[s].p, ctx->scNodes[s].p, (size_t)cnt * k * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    *T = cnt;
    *cycles = ctx->scListHost[s].as<int32_t>();
    return SMSGPU_OK;
}

// device part of the scan: codes, ranks, the violated walks in ctx->scFin; returns with the number of violated cycles known
static int smallCyclesScanDevice(smsgpu_ctx *ctx, int k, long long *Uout) {
    if (!ctx) return SMSGPU_ERR_ARG;
    if (k != 3 && k != 4) return fail(ctx, SMSGPU_ERR_ARG, "k must be 3 or 4");
    if (!ctx->haveInput) return fail(ctx, SMSGPU_ERR_STATE, "small_cycles_scan before set_round_input");
    CU(cudaSetDevice(ctx->device));
    const int s = k - 3;
    const long long T = ctx->scCount[s];
    cudaStream_t st = ctx->stream;
    *Uout = 0;
    ctx->scMs[0] = ctx->scMs[1] = ctx->scMs[2] = 0.0;
    ctx->scFin.ncuts = ctx->scFin.nnodes = 0;
    if (T == 0) return SMSGPU_OK;
    const long long ntiles = (T + kSmallTile - 1) / kSmallTile;
    CU(ctx->scCode.ensure((size_t)T));
    CU(ctx->wCnt.ensure((size_t)(ntiles + 1) * 4));
    CU(ctx->wBase.ensure((size_t)(ntiles + 2) * 8));
    CU(ctx->scXc.ensure((size_t)std::max<long long>(ctx->m, 1) * 8));
    CU(cudaEventRecord(ctx->scEv[0], st));
    // clamp(x*) once per edge (m values) instead of once per gather (k T values)
    clampWeightsKernel<<<std::min<unsigned>(gridFor(ctx->m, 256), (unsigned)ctx->smCount * 8), 256, 0, st>>>(ctx->m, ctx->xPtr, ctx->scXc.as<double>());
    if (k == 3)
        smallCycleScanKernel<3, true><<<(unsigned)ntiles, kSmallBlock, 0, st>>>(T, ctx->scEids[s].as<int32_t>(), ctx->scXc.as<double>(),
                                                                              ctx->scCode.as<uint8_t>(), ctx->wCnt.as<int32_t>());
    else
        smallCycleScanKernel<4, true><<<(unsigned)ntiles, kSmallBlock, 0, st>>>(T, ctx->scEids[s].as<int32_t>(), ctx->scXc.as<double>(),
                                                                              ctx->scCode.as<uint8_t>(), ctx->wCnt.as<int32_t>());
    CU(cudaGetLastError());
    CU(cudaEventRecord(ctx->scEv[1], st));
    if (int rc = exclusiveScan(ctx, ctx->wCnt.as<int32_t>(), ntiles, ctx->wBase.as<long long>())) return rc;
    long long U = 0;
    CU(cudaMemcpyAsync(&U, ctx->wBase.as<long long>() + ntiles, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CutListDev &F = ctx->scFin;
    CU(F.off.ensure((size_t)(U + 1) * 8));
    CU(F.node.ensure((size_t)std::max<long long>(U * k, 1) * 4));
    CU(F.eid.ensure((size_t)std::max<long long>(U * k, 1) * 4));
    CU(F.stay.ensure((size_t)std::max<long long>(U * k, 1)));
    CU(F.ncross.ensure((size_t)std::max<long long>(U, 1) * 4));
    CU(F.src.ensure((size_t)std::max<long long>(U, 1) * 4));
    CU(F.viol.ensure((size_t)std::max<long long>(U, 1) * 8));
    if (U > 0) {
        if (k == 3)
            smallCycleGatherKernel<3><<<(unsigned)ntiles, kSmallBlock, 0, st>>>(
                T, ctx->scNodes[s].as<int32_t>(), ctx->scEids[s].as<int32_t>(), ctx->xPtr, ctx->scCode.as<uint8_t>(),
                ctx->wBase.as<long long>(), F.node.as<int32_t>(), F.eid.as<int32_t>(), F.stay.as<uint8_t>(), F.ncross.as<int32_t>(),
                F.src.as<int32_t>(), F.viol.as<double>(), F.off.as<long long>());
        else
            smallCycleGatherKernel<4><<<(unsigned)ntiles, kSmallBlock, 0, st>>>(
                T, ctx->scNodes[s].as<int32_t>(), ctx->scEids[s].as<int32_t>(), ctx->xPtr, ctx->scCode.as<uint8_t>(),
                ctx->wBase.as<long long>(), F.node.as<int32_t>(), F.eid.as<int32_t>(), F.stay.as<uint8_t>(), F.ncross.as<int32_t>(),
                F.src.as<int32_t>(), F.viol.as<double>(), F.off.as<long long>());
        smallCycleTailKernel<<<1, 1, 0, st>>>(F.off.as<long long>(), U, k);
        CU(cudaGetLastError());
    }
    CU(cudaEventRecord(ctx->scEv[2], st));
    F.ncuts = U;
    F.nnodes = U * k;
    *Uout = U;
    return SMSGPU_OK;
}
static void smallCyclesTimes(smsgpu_ctx *ctx, bool withD2h) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, ctx->scEv[0], ctx->scEv[1]) == cudaSuccess) ctx->scMs[0] = ms;
    if (cudaEventElapsedTime(&ms, ctx->scEv[1], ctx->scEv[2]) == cudaSuccess) ctx->scMs[1] = ms;
    if (withD2h && cudaEventElapsedTime(&ms, ctx->scEv[2], ctx->scEv[3]) == cudaSuccess) ctx->scMs[2] = ms;
    (void)cudaGetLastError();
}

int smsgpu_small_cycles_scan_device(smsgpu_ctx *ctx, int k) {
    long long U = 0;
    if (int rc = smallCyclesScanDevice(ctx, k, &U)) return rc;
    if (ctx->scCount[k - 3] > 0) {
        CU(cudaStreamSynchronize(ctx->stream));
        smallCyclesTimes(ctx, false);
    }
    return SMSGPU_OK;
}

int smsgpu_small_cycles_scan(smsgpu_ctx *ctx, int k) {
    long long U = 0;
    if (int rc = smallCyclesScanDevice(ctx, k, &U)) return rc;
    const int s = k - 3;
    cudaStream_t st = ctx->stream;
    CutListHost &H = ctx->scHost[s];
    CutListDev &F = ctx->scFin;
    H.ncuts = H.nnodes = 0;
    CU(H.off.ensure((size_t)(U + 1) * 8));
    H.off.as<int64_t>()[0] = 0;
    if (ctx->scCount[s] == 0) return SMSGPU_OK;
    if (U > 0) {
        CU(H.node.ensure((size_t)U * k * 4));
        CU(H.eid.ensure((size_t)U * k * 4));
        CU(H.stay.ensure((size_t)U * k));
        CU(H.ncross.ensure((size_t)U * 4));
        CU(H.src.ensure((size_t)U * 4));
        CU(H.viol.ensure((size_t)U * 8));
        CU(cudaMemcpyAsync(H.off.p, F.off.p, (size_t)(U + 1) * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.node.p, F.node.p, (size_t)U * k * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.eid.p, F.eid.p, (size_t)U * k * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.stay.p, F.stay.p, (size_t)U * k, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.ncross.p, F.ncross.p, (size_t)U * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.src.p, F.src.p, (size_t)U * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(H.viol.p, F.viol.p, (size_t)U * 8, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaEventRecord(ctx->scEv[3], st));
    CU(cudaStreamSynchronize(st));
    smallCyclesTimes(ctx, true);
    H.ncuts = U;
    H.nnodes = U * k;
    return SMSGPU_OK;
}

int smsgpu_small_cycles_timing(smsgpu_ctx *ctx, double *ms_scan, double *ms_gather, double *ms_d2h) {
    if (!ctx) return SMSGPU_ERR_ARG;
    if (ms_scan) *ms_scan = ctx->scMs[0];
    if (ms_gather) *ms_gather = ctx->scMs[1];
    if (ms_d2h) *ms_d2h = ctx->scMs[2];
    return SMSGPU_OK;
}

int smsgpu_get_small_cycle_cuts(smsgpu_ctx *ctx, int k, smsgpu_cuts *out) {
    if (!ctx || !out) return SMSGPU_ERR_ARG;
    if (k != 3 && k != 4) return fail(ctx, SMSGPU_ERR_ARG, "k must be 3 or 4");
    const CutListHost &H = ctx->scHost[k - 3];
    out->ncuts = H.ncuts;
    out->nnodes = H.nnodes;
    out->off = H.off.as<int64_t>();
    out->node = H.node.as<int32_t>();
    out->eid = H.eid.as<int32_t>();
    out->is_stay = H.stay.as<uint8_t>();
    out->n_cross = H.ncross.as<int32_t>();
    out->src = H.src.as<int32_t>();
    out->violation = H.viol.as<double>();
    return SMSGPU_OK;
}

} // extern "C"

// =====================================================================================================================
// multi-GPU: NCCL all-gather of the per-GPU cut lists + global dedup (C1).  NCCL is loaded lazily with dlopen so
// that the library has no link-time dependency on it (a torch process already carries libnccl.so.2).
// =====================================================================================================================
#include <dlfcn.h>
#include <nccl.h>

namespace {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};
NcclApi loadNccl() {
    NcclApi api;
    void *h = nullptr;
    for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
        h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) return api;
    api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))dlsym(h, "ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))dlsym(h, "ncclCommDestroy");
    api.AllGather = (decltype(api.AllGather))dlsym(h, "ncclAllGather");
    api.GetErrorString = (decltype(api.GetErrorString))dlsym(h, "ncclGetErrorString");
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.GetErrorString;
    api.handle = h; // published last, with every symbol resolved
    return api;
}
NcclApi &nccl() {
    static NcclApi api = loadNccl(); // function-local static: initialised once, thread-safe (one thread per GPU may race here)
    return api;
}

#define NC(call)                                                                                                  \
    do {                                                                                                          \
        ncclResult_t r_ = (call);                                                                                 \
        if (r_ != ncclSuccess) return fail(ctx, SMSGPU_ERR_NCCL, "%s: %s", #call, nccl().GetErrorString(r_));    \
    } while (0)

// The exchange is TWO collectives per round: an all-gather of a 4-word header per rank
// {cuts, nodes, error flag, -} and ONE all-gather of a packed byte block per rank
//   [len | src | seq | off : int32 x maxU each][node | eid : int32 x maxN each][stay : uint8 x maxN, padded to 16]
// with maxU / maxN the largest counts of the header.  A rank whose local round failed still joins the header
// exchange with its error flag set, so that every rank returns an error together instead of waiting for it forever.
struct PackLayout {
    long long maxU, maxN, oLen, oSrc, oSeq, oOff, oNode, oEid, oStay, bytes;
};
__host__ __device__ inline PackLayout packLayout(long long maxU, long long maxN) {
    PackLayout L;
    L.maxU = maxU;
    L.maxN = maxN;
    L.oLen = 0;
    L.oSrc = 4 * maxU;
    L.oSeq = 8 * maxU;
    L.oOff = 12 * maxU;
    L.oNode = 16 * maxU;
    L.oEid = L.oNode + 4 * maxN;
    L.oStay = L.oEid + 4 * maxN;
    L.bytes = (L.oStay + maxN + 15) & ~15ll;
    return L;
}
// local final list -> this rank's block
__global__ void packCutsKernel(const long long U, const long long N, const long long *off, const int32_t *src,
                               const int32_t *seq, const int32_t *node, const int32_t *eid, const uint8_t *stay,
                               const PackLayout L, unsigned char *blk) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int32_t *pLen = reinterpret_cast<int32_t *>(blk + L.oLen), *pSrc = reinterpret_cast<int32_t *>(blk + L.oSrc);
    int32_t *pSeq = reinterpret_cast<int32_t *>(blk + L.oSeq), *pOff = reinterpret_cast<int32_t *>(blk + L.oOff);
    int32_t *pNode = reinterpret_cast<int32_t *>(blk + L.oNode), *pEid = reinterpret_cast<int32_t *>(blk + L.oEid);
    uint8_t *pStay = blk + L.oStay;
    for (long long k = t0; k < U; k += stride) {
        pLen[k] = (int32_t)(off[k + 1] - off[k]);
        pOff[k] = (int32_t)off[k];
        pSrc[k] = src[k];
        pSeq[k] = seq[k];
    }
    for (long long j = t0; j < N; j += stride) {
        pNode[j] = node[j];
        pEid[j] = eid[j];
        pStay[j] = stay[j];
    }
}
// gathered blocks -> one cut view over all ranks (cut r * maxU + k = cut k of rank r; padding cuts are invalid)
__global__ void unpackMergedKernel(const int size, const PackLayout L, const long long *hdr, const unsigned char *all,
                                   int32_t *mLen, int32_t *mSrc, int32_t *mSeq, int32_t *mOff, uint8_t *mValid,
                                   int32_t *mNode, int32_t *mEid, uint8_t *mStay) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (long long i = t0; i < (long long)size * L.maxU; i += stride) {
        const long long r = i / L.maxU, k = i % L.maxU;
        const unsigned char *blk = all + r * L.bytes;
        const bool v = k < hdr[4 * r];
        mValid[i] = v;
        mLen[i] = v ? reinterpret_cast<const int32_t *>(blk + L.oLen)[k] : 0;
        mSrc[i] = v ? reinterpret_cast<const int32_t *>(blk + L.oSrc)[k] : 0;
        mSeq[i] = v ? reinterpret_cast<const int32_t *>(blk + L.oSeq)[k] : 0;
        mOff[i] = v ? (int32_t)(r * L.maxN + reinterpret_cast<const int32_t *>(blk + L.oOff)[k]) : 0;
    }
    for (long long i = t0; i < (long long)size * L.maxN; i += stride) {
        const long long r = i / L.maxN, j = i % L.maxN;
        const unsigned char *blk = all + r * L.bytes;
        const bool v = j < hdr[4 * r + 1];
        mNode[i] = v ? reinterpret_cast<const int32_t *>(blk + L.oNode)[j] : 0;
        mEid[i] = v ? reinterpret_cast<const int32_t *>(blk + L.oEid)[j] : 0;
        mStay[i] = v ? (blk + L.oStay)[j] : 0;
    }
}

int mergeAcrossRanks(smsgpu_ctx *ctx, bool dedup, int localRc) {
    NcclApi &A = nccl();
    cudaStream_t st = ctx->stream;
    const int size = ctx->commSize;
    if (ctx->nranks != size || ctx->rank != ctx->commRank)
        return fail(ctx, SMSGPU_ERR_STATE, "source sharding (%d/%d) does not match the communicator (%d/%d)", ctx->rank,
                    ctx->nranks, ctx->commRank, size);
    ncclComm_t comm = static_cast<ncclComm_t>(ctx->ncclComm);
    CutListDev &F = ctx->fin;
    // 1. header
    CU(ctx->gCounts.ensure((size_t)(4 * size + 4) * 8));
    CU(ctx->hGCounts.ensure((size_t)(4 * size + 4) * 8));
    long long *hg = ctx->hGCounts.as<long long>();
    long long *dg = ctx->gCounts.as<long long>();
    hg[4 * size] = localRc ? 0 : F.ncuts;
    hg[4 * size + 1] = localRc ? 0 : F.nnodes;
    hg[4 * size + 2] = localRc;
    hg[4 * size + 3] = 0;
    CU(cudaMemcpyAsync(dg + 4 * size, hg + 4 * size, 32, cudaMemcpyHostToDevice, st));
    NC(A.AllGather(dg + 4 * size, dg, 4, ncclInt64, comm, st));
    CU(cudaMemcpyAsync(hg, dg, (size_t)4 * size * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    long long maxU = 0, maxN = 0;
    for (int r = 0; r < size; ++r) {
        if (hg[4 * r + 2] != 0) {
            if (localRc) return localRc; // this rank's own error text stays
            return fail(ctx, SMSGPU_ERR_INTERNAL, "rank %d failed its local round (status %lld); no rank returns cuts", r, hg[4 * r + 2]);
        }
        maxU = std::max(maxU, hg[4 * r]);
        maxN = std::max(maxN, hg[4 * r + 1]);
    }
    if (maxU == 0) return 0;
    maxN = std::max<long long>(maxN, 1);
    if ((long long)size * maxN > 0x7FFFFFFFll) return fail(ctx, SMSGPU_ERR_INTERNAL, "merged cut list exceeds 2^31 nodes");
    // 2. pack + gather
    const PackLayout L = packLayout(maxU, maxN);
    CU(ctx->pBlock.ensure((size_t)L.bytes));
    CU(ctx->mBlocks.ensure((size_t)L.bytes * size));
    const long long tot = (long long)size * maxU, totN = (long long)size * maxN;
    CU(ctx->mLen.ensure((size_t)tot * 4));
    CU(ctx->mSrc.ensure((size_t)tot * 4));
    CU(ctx->mSeq.ensure((size_t)tot * 4));
    CU(ctx->mOff.ensure((size_t)tot * 4));
    CU(ctx->mValid.ensure((size_t)tot));
    CU(ctx->mNode.ensure((size_t)totN * 4));
    CU(ctx->mEid.ensure((size_t)totN * 4));
    CU(ctx->mStay.ensure((size_t)totN));
    packCutsKernel<<<ctx->smCount * 4, 256, 0, st>>>(F.ncuts, F.nnodes, F.off.as<long long>(), F.src.as<int32_t>(),
                                                     F.seq.as<int32_t>(), F.node.as<int32_t>(), F.eid.as<int32_t>(),
                                                     F.stay.as<uint8_t>(), L, ctx->pBlock.as<unsigned char>());
    CU(cudaGetLastError());
    NC(A.AllGather(ctx->pBlock.p, ctx->mBlocks.p, (size_t)L.bytes, ncclUint8, comm, st));
    unpackMergedKernel<<<ctx->smCount * 4, 256, 0, st>>>(size, L, dg, ctx->mBlocks.as<unsigned char>(), ctx->mLen.as<int32_t>(),
                                                         ctx->mSrc.as<int32_t>(), ctx->mSeq.as<int32_t>(), ctx->mOff.as<int32_t>(),
                                                         ctx->mValid.as<uint8_t>(), ctx->mNode.as<int32_t>(), ctx->mEid.as<int32_t>(),
                                                         ctx->mStay.as<uint8_t>());
    CU(cudaGetLastError());
    ctx->stats.total_launches += 2;
    // 3. global dedup: every rank computes the identical full list
    CutView V{ctx->mNode.as<int32_t>(), ctx->mEid.as<int32_t>(), ctx->mStay.as<uint8_t>(), ctx->mOff.as<int32_t>(),
              ctx->mLen.as<int32_t>(), ctx->mSrc.as<int32_t>(), ctx->mSeq.as<int32_t>(), ctx->mValid.as<uint8_t>(), tot};
    return runDedup(ctx, V, totN, false, dedup, 1, 0, ctx->nsrcGlobal);
}

} // namespace

extern "C" {

int smsgpu_nccl_unique_id(void *id128) {
    smsgpu_ctx *ctx = nullptr;
    if (!id128) return fail(ctx, SMSGPU_ERR_ARG, "id128 is null");
    NcclApi &A = nccl();
    if (!A.ok) return fail(ctx, SMSGPU_ERR_NCCL, "libnccl.so.2 could not be loaded");
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    NC(A.GetUniqueId(&id));
    std::memcpy(id128, &id, 128);
    return SMSGPU_OK;
}

int smsgpu_comm_init(smsgpu_ctx *ctx, const void *id128, int rank, int nranks) {
    if (!ctx) return SMSGPU_ERR_ARG;
    if (!id128 || nranks < 1 || rank < 0 || rank >= nranks) return fail(ctx, SMSGPU_ERR_ARG, "bad communicator arguments");
    NcclApi &A = nccl();
    if (!A.ok) return fail(ctx, SMSGPU_ERR_NCCL, "libnccl.so.2 could not be loaded");
    CU(cudaSetDevice(ctx->device));
    smsgpu_comm_destroy(ctx);
    ncclUniqueId id;
    std::memcpy(&id, id128, 128);
    ncclComm_t comm = nullptr;
    NC(A.CommInitRank(&comm, nranks, id, rank));
    ctx->ncclComm = comm;
    ctx->commRank = rank;
    ctx->commSize = nranks;
    return SMSGPU_OK;
}

int smsgpu_comm_destroy(smsgpu_ctx *ctx) {
    if (!ctx) return SMSGPU_ERR_ARG;
    if (ctx->ncclComm) {
        if (ctx->stream) cudaStreamSynchronize(ctx->stream);
        nccl().CommDestroy(static_cast<ncclComm_t>(ctx->ncclComm));
        ctx->ncclComm = nullptr;
    }
    ctx->commRank = 0;
    ctx->commSize = 1;
    return SMSGPU_OK;
}

} // extern "C"

// =====================================================================================================================
// K6 without any allocation: smsgpu_small_cycles_scan_noalloc (include/smsgpu.h).  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
// loops, harnesses that forbid it).  Same scan / rank / gather as smallCyclesScanDevice above, but stateless: inputs and
// outputs are the caller's device buffers, and the scratch of the scan (one code byte per cycle, one count and one base
// per tile of 1024 cycles) lives in ONE file-scope device array of fixed size, which every GPU that loads the library
// holds until exit.  Every call on a device shares that array: calls that overlap on the GPU race.
// =====================================================================================================================
#include <mutex>

namespace {
constexpr long long kNoallocMaxCycles = SMSGPU_NOALLOC_MAX_CYCLES;
constexpr long long kNoallocMaxTiles = (kNoallocMaxCycles + kSmallTile - 1) / kSmallTile;
static_assert(kNoallocMaxTiles <= 4096, "the tile scan below is one CTA of 1024 threads x 4 tiles");
constexpr size_t kNoallocCodeBytes = ((size_t)kNoallocMaxCycles + 255) & ~(size_t)255;
constexpr size_t kNoallocCountBytes = (((size_t)kNoallocMaxTiles + 1) * 4 + 255) & ~(size_t)255;
constexpr size_t kNoallocBaseBytes = (((size_t)kNoallocMaxTiles + 2) * 8 + 255) & ~(size_t)255;
constexpr size_t kNoallocScratchBytes = kNoallocCodeBytes + kNoallocCountBytes + kNoallocBaseBytes; // ~4.05 MiB
static_assert(kNoallocScratchBytes == SMSGPU_NOALLOC_SCRATCH_BYTES, "include/smsgpu.h states the size of the array");
} // namespace

// In host code this name is a host-side stand-in, NOT a device address: the address the kernels get comes from
// cudaGetSymbolAddress (noallocScratchOf), once per device.
static __device__ __align__(256) unsigned char g_noallocScratch[kNoallocScratchBytes];

namespace {
// exclusive scan of the per-tile counts by one CTA (ntiles <= 4096), the total, and the closing offset of the walk list
__global__ void __launch_bounds__(1024) noallocTileScanKernel(const int ntiles, const int32_t *__restrict__ cnt,
                                                              long long *__restrict__ base, const long long cap, const int K,
                                                              long long *__restrict__ oOff, long long *__restrict__ dCount) {
    __shared__ long long warpSum[32];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    int v[4];
    long long mine = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int t = 4 * (int)threadIdx.x + i;
        v[i] = t < ntiles ? cnt[t] : 0;
        mine += v[i];
    }
    long long incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const long long a = __shfl_up_sync(0xFFFFFFFFu, incl, off);
        if (lane >= off) incl += a;
    }
    if (lane == 31) warpSum[wrp] = incl;
    __syncthreads();
    long long before = 0, total = 0;
    for (int k = 0; k < 32; ++k) {
        if (k < wrp) before += warpSum[k];
        total += warpSum[k];
    }
    long long run = before + incl - mine;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int t = 4 * (int)threadIdx.x + i;
        if (t < ntiles) base[t] = run;
        run += v[i];
    }
    if (threadIdx.x == 0) {
        base[ntiles] = total;
        *dCount = total;
        const long long kept = total < cap ? total : cap;
        oOff[kept] = (long long)K * kept;
    }
}

// smallCycleGatherKernel (small_cycles.cuh) with a bound on the output: violated cycles of list rank >= cap are dropped
template <int K>
__global__ void __launch_bounds__(kSmallBlock) noallocGatherKernel(const long long T, const int32_t *__restrict__ nodes,
                                                                   const int32_t *__restrict__ eids, const double *__restrict__ x,
                                                                   const uint8_t *__restrict__ code,
                                                                   const long long *__restrict__ tileBase, const long long cap,
                                                                   int32_t *oNode, int32_t *oEid, uint8_t *oStay, int32_t *oNcross,
                                                                   int32_t *oSrc, double *oViol, long long *oOff) {
    __shared__ int warpSum[kSmallBlock / 32];
    const long long base = (long long)blockIdx.x * kSmallTile;
    const long long out0 = tileBase[blockIdx.x];
    if (out0 >= cap || tileBase[blockIdx.x + 1] == out0) return; // CTA uniform: nothing of this tile is kept
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    int carry = 0; // violated cycles of the earlier items of this tile (every thread keeps its own copy)
    for (int it = 0; it < kSmallItems; ++it) {
        const long long t = base + it * kSmallBlock + threadIdx.x;
        const unsigned c = t < T ? code[t] : 0u;
        const int v = c ? 1 : 0;
        int incl = v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int a = __shfl_up_sync(0xFFFFFFFFu, incl, off);
            if (lane >= off) incl += a;
        }
        if (lane == 31) warpSum[wrp] = incl;
        __syncthreads();
        int before = carry, tot = 0;
#pragma unroll
        for (int k = 0; k < kSmallBlock / 32; ++k) {
            if (k < wrp) before += warpSum[k];
            tot += warpSum[k];
        }
        const long long k = out0 + before + incl - 1;
        if (v && k < cap) {
            const unsigned sm = c & 0x7Fu;
            int nc = 0;
            double w[K];
#pragma unroll
            for (int i = 0; i < K; ++i) {
                const int32_t e = eids[K * t + i];
                w[i] = clamp01(__ldg(x + e));
                oNode[K * k + i] = nodes[K * t + i];
                oEid[K * k + i] = e;
                const int s = (sm >> i) & 1u;
                oStay[K * k + i] = (uint8_t)s;
                nc += 1 - s;
            }
            double len = 0; // same summation order as the scan's and as smallCycleGatherKernel's
#pragma unroll
            for (int i = 0; i < K; ++i) len += ((sm >> i) & 1u) ? w[i] : 1 - w[i];
            oNcross[k] = nc;
            oSrc[k] = (int32_t)t;
            oViol[k] = 1.0 - len;
            oOff[k] = (long long)K * k;
        }
        carry += tot;
        __syncthreads(); // warpSum is rewritten by the next item
    }
}

constexpr int kNoallocMaxDevices = 64;
std::mutex g_noallocMutex;
unsigned char *g_noallocPtr[kNoallocMaxDevices] = {};

// device address of g_noallocScratch on `device` (which must be the current device), resolved once per device
cudaError_t noallocScratchOf(const int device, unsigned char **out) {
    std::lock_guard<std::mutex> lock(g_noallocMutex);
    if (!g_noallocPtr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noallocScratch);
        if (e != cudaSuccess) return e;
        g_noallocPtr[device] = static_cast<unsigned char *>(p);
    }
    *out = g_noallocPtr[device];
    return cudaSuccess;
}

int noallocFail(char *err, const size_t errLen, const int code, const char *fmt, ...) {
    if (err && errLen > 0) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, errLen, fmt, ap);
        va_end(ap);
    }
    return code;
}
} // namespace

extern "C" int smsgpu_small_cycles_scan_noalloc(int device, void *stream, int k, int64_t T, const int32_t *d_cycles,
                                                const int32_t *d_cycle_eids, const double *d_x, int64_t out_capacity,
                                                int32_t *d_node, int32_t *d_eid, uint8_t *d_is_stay, int32_t *d_n_cross,
                                                int32_t *d_src, double *d_violation, int64_t *d_off, int64_t *d_count,
                                                int64_t *h_count, char *err, size_t err_len) {
    if (err && err_len > 0) err[0] = 0;
    if (k != 3 && k != 4) return noallocFail(err, err_len, SMSGPU_ERR_ARG, "k must be 3 or 4 (got %d)", k);
    if (T < 0 || out_capacity < 0) return noallocFail(err, err_len, SMSGPU_ERR_ARG, "T and out_capacity must be >= 0");
    if (T > kNoallocMaxCycles)
        return noallocFail(err, err_len, SMSGPU_ERR_ARG,
                           "T = %lld cycles exceeds SMSGPU_NOALLOC_MAX_CYCLES = %lld (the static scratch array holds no more; "
                           "split the list or use smsgpu_small_cycles_scan)",
                           (long long)T, kNoallocMaxCycles);
    if (device < 0 || device >= kNoallocMaxDevices) return noallocFail(err, err_len, SMSGPU_ERR_ARG, "device %d out of range", device);
    if (T > 0 && (!d_cycles || !d_cycle_eids || !d_x)) return noallocFail(err, err_len, SMSGPU_ERR_ARG, "null input pointer");
    if (!d_off || !d_count) return noallocFail(err, err_len, SMSGPU_ERR_ARG, "d_off and d_count must not be null");
    if (out_capacity > 0 && (!d_node || !d_eid || !d_is_stay || !d_n_cross || !d_src || !d_violation))
        return noallocFail(err, err_len, SMSGPU_ERR_ARG, "null output pointer with out_capacity > 0");
#define NA_CU(call)                                                                                              \
    do {                                                                                                         \
        const cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) {                                                                                 \
            (void)cudaGetLastError();                                                                            \
            return noallocFail(err, err_len, SMSGPU_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call,         \
                               cudaGetErrorString(e_));                                                          \
        }                                                                                                        \
    } while (0)
    NA_CU(cudaSetDevice(device));
    unsigned char *scratch = nullptr;
    NA_CU(noallocScratchOf(device, &scratch));
    uint8_t *code = scratch;
    int32_t *tileCount = reinterpret_cast<int32_t *>(scratch + kNoallocCodeBytes);
    long long *tileBase = reinterpret_cast<long long *>(scratch + kNoallocCodeBytes + kNoallocCountBytes);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long ntiles = (T + kSmallTile - 1) / kSmallTile;
    // Nothing in the scratch has to start at zero: the scan writes code[0, T) and tileCount[0, ntiles), the tile scan
    // writes tileBase[0, ntiles], and only those ranges are read afterwards -- what an earlier call left behind is never seen.
    if (ntiles > 0) {
        if (k == 3)
            smallCycleScanKernel<3, false><<<(unsigned)ntiles, kSmallBlock, 0, st>>>(T, d_cycle_eids, d_x, code, tileCount);
        else
            smallCycleScanKernel<4, false><<<(unsigned)ntiles, kSmallBlock, 0, st>>>(T, d_cycle_eids, d_x, code, tileCount);
        NA_CU(cudaGetLastError());
    }
    noallocTileScanKernel<<<1, 1024, 0, st>>>((int)ntiles, tileCount, tileBase, out_capacity, k, reinterpret_cast<long long *>(d_off),
                                             reinterpret_cast<long long *>(d_count));
    NA_CU(cudaGetLastError());
    if (ntiles > 0 && out_capacity > 0) {
        if (k == 3)
            noallocGatherKernel<3><<<(unsigned)ntiles, kSmallBlock, 0, st>>>(T, d_cycles, d_cycle_eids, d_x, code, tileBase, out_capacity, d_node,
                                                                           d_eid, d_is_stay, d_n_cross, d_src, d_violation,
                                                                           reinterpret_cast<long long *>(d_off));
        else
            noallocGatherKernel<4><<<(unsigned)ntiles, kSmallBlock, 0, st>>>(T, d_cycles, d_cycle_eids, d_x, code, tileBase, out_capacity, d_node,
                                                                           d_eid, d_is_stay, d_n_cross, d_src, d_violation,
                                                                           reinterpret_cast<long long *>(d_off));
        NA_CU(cudaGetLastError());
    }
    if (h_count) {
        NA_CU(cudaMemcpyAsync(h_count, d_count, 8, cudaMemcpyDeviceToHost, st));
        NA_CU(cudaStreamSynchronize(st));
    }
#undef NA_CU
    return SMSGPU_OK;
}
