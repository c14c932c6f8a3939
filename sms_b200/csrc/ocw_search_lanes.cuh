// Kernels of the lanes engine (one source per THREAD; the sequential core and the reasons are in ocw_lanes_core.cuh).
//
// A launch pair works through one BATCH of work items, one tile of working memory per item:
//   ocwLanesSearchKernel   thread i searches item `first + i`: a PAIR (2p, 2p+1) of consecutive entries of the source
//                          list -- the two sides (v,0), (v,1) of one vertex in the handler's list, whose searches are mirror
//                          images of each other unless an extraction broke a (dist, hops) tie by id (twin symmetry,
//                          ocw_search_pairs.cuh) -- or a single entry when mirroring is off.  It leaves labels, queue
//                          entries (with the predecessors), the history of intermediates and a LanesRec in its tile.
//   ocwLanesFinishKernel   a warp per tile: selection of intermediates, walks, chord repair, label export, counters -- the
//                          32-lane code shared with the other engines (finishSourceMem).  Sources that hit a capacity, and
//                          second sources that are not mirror images, go to the overflow list: the pairs engine runs them.
// Batches alternate between streams with their own tile sets, so the finishing of one overlaps the search of the next.
#pragma once
#include "ocw_lanes_core.cuh"
#include "ocw_search_pairs.cuh"

namespace smsgpu {

struct LanesArgs {
    LanesCfg cfg;
    const uint2 *rowinfo;
    unsigned char *tiles;  // tile set of this launch
    size_t tileStride;
    long long first, count; // items [first, first + count) of this batch
    int itemSize;          // 2: pairs of list entries (twins mirrored), 1: single entries
    uint32_t stamp;        // stamp of this batch (unique per tile set: every tile runs one search per batch)
    // finishing
    const PairsKArgs *kargs;
    unsigned char *warpScratch; // walk scratch of the finishing warps
    size_t warpScratchStride;
    TileCaps fcaps;        // capacities as the finishing code sees them
};

#ifndef SMSGPU_LANES_BLOCK
#define SMSGPU_LANES_BLOCK 64
#endif
#ifndef SMSGPU_LANES_MINB // resident CTAs per SM the search kernel is compiled for (caps its registers)
#define SMSGPU_LANES_MINB 1
#endif

__global__ void __launch_bounds__(SMSGPU_LANES_BLOCK, SMSGPU_LANES_MINB) ocwLanesSearchKernel(const GraphDev G, const SearchParams P, const LanesArgs A) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= A.count) return;
    const LanesTile T = lanesCarve(A.tiles + (size_t)t * A.tileStride, A.cfg);
    const long long total = P.worklist ? P.nwork : P.nsrc;
    const long long w0 = (A.first + t) * A.itemSize;
    LanesRec rec;
    rec.srcIdx = -1;
    rec.twinIdx = -1;
    rec.flags = 0;
    rec.src = 0;
    rec.best = 0.0;
    rec.histN = rec.touchedN = rec.arcs = rec.settled = rec.stamp = rec.qN = 0;
    rec.pad[0] = rec.pad[1] = rec.pad[2] = rec.pad[3] = 0;
    // the entries of the item that are to be searched (same rules as the pairs engine's fetch)
    int32_t idx[2] = {-1, -1};
    uint32_t sid[2] = {0, 0};
    for (int k = 0; k < A.itemSize; ++k) {
        const long long w = w0 + k;
        if (w >= total) break;
        const int32_t i = P.worklist ? P.worklist[w] : (int32_t)w;
        const uint32_t s = P.sources ? (uint32_t)P.sources[i] : (uint32_t)(((((long long)(i >> 1)) * P.srcStride + P.srcBegin) << 1) | (i & 1));
        // `if (blacklist_[currentSource_]) return {}` -- indexed with the DOUBLED id (ocw_seperator.cpp:334)
        const bool skip = ((long long)s < G.n && P.blacklist[s] != 0) || (long long)(s >> 1) >= G.n;
        if (skip) continue;
        idx[k] = i;
        sid[k] = s;
    }
    int f = idx[0] >= 0 ? 0 : (idx[1] >= 0 ? 1 : -1);
    if (f >= 0) {
        const uint32_t src = sid[f];
        lanesSearch(G.arcs, A.rowinfo, A.cfg, T, src, A.rowinfo[src >> 1], A.stamp, P.bestInit, rec);
        rec.srcIdx = idx[f];
        if (f == 0 && idx[1] >= 0) {
            rec.twinIdx = idx[1];
            if (sid[1] == (src ^ 1u)) rec.flags |= 4u;
        }
    }
    uint4 *rp = reinterpret_cast<uint4 *>(T.rec);
    const uint4 *sp = reinterpret_cast<const uint4 *>(&rec);
    rp[0] = sp[0];
    rp[1] = sp[1];
    rp[2] = sp[2];
    rp[3] = sp[3];
}

__global__ void __launch_bounds__(128) ocwLanesFinishKernel(const LanesArgs A) {
    const int warpInCta = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + warpInCta;
    const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
    const PairsKArgs *K = A.kargs;
    for (long long t = warp; t < A.count; t += nwarps) {
        const LanesTile T = lanesCarve(A.tiles + (size_t)t * A.tileStride, A.cfg);
        const LanesRec rec = *T.rec;
        if (rec.srcIdx < 0) continue; // every entry of the item was skipped
        TileMem M;
        {
            TileCaps sc = A.fcaps; // the scratch arrays only: table and history live in the tile
            sc.tabCap = 0;
            sc.touchedCap = 0;
            sc.histCap = 0;
            carveTileMem(M, A.warpScratch + (size_t)warp * A.warpScratchStride, sc);
        }
        M.caps = A.fcaps;
        M.tab = T.tab;
        M.tabMask = A.fcaps.tabCap - 1;
        M.hist = T.hist;
        M.pred = nullptr;
        M.touched = nullptr;
        M.hopMask = kLanesHopMask;
        M.stamp = rec.stamp;
        M.lq = T.q;
        M.lqShift = kLanesHopBits;
        M.lqMask = kLanesQMask;
        const int overflow = (int)(rec.flags & 1u);
        const bool isTwin = (rec.flags & 4u) != 0;
        const int mirrorOk = (isTwin && !(rec.flags & 2u)) ? 1 : 0;
        const int ok = finishSourceMem(K, M, rec.best, rec.histN, rec.touchedN, rec.arcs, rec.settled, overflow, rec.src, rec.srcIdx,
                                       rec.twinIdx, mirrorOk);
        __syncwarp();
        // the second source of the item, when this search does not stand for it: the next engine searches it
        if (rec.twinIdx >= 0 && ok != 2 && lane == 0) {
            const unsigned long long k = atomicAdd(K->P.ovfCount, 1ull);
            K->P.ovfList[k] = rec.twinIdx;
        }
    }
}

} // namespace smsgpu
