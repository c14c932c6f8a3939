// Kernels of the separation round that touch the graph:
//   dcWeightsKernel  (K1)      per-round arc records: clamp(x*[eid]) + blacklist bit, CSR order
//   ocwSearchKernel  (K2-K4)   persistent tiles pulling sources from an atomic cursor
//   initTablesKernel           one-time clearing of the HBM label tables of tiers 1 / 2
#pragma once
#include "ocw_walks.cuh"
#include "ocw_search_pairs.cuh"
#include "ocw_search_lanes.cuh"
#include "ocw_search_runs.cuh"

namespace smsgpu {

// ---- K1 ------------------------------------------------------------------------------------------------------
// Replaces the m calls of OcwSeparator::updateWeights (ocw_seperator.cpp:25-31; clamp = std::min(std::max(0., w), 1.))
// and folds the round's blacklist (ocw_seperator.hpp:25-28) into the arc.  One thread per 4 consecutive arcs:
// 128-bit loads of col / eid, four 128-bit stores.
__global__ void __launch_bounds__(256) dcWeightsKernel(const long long narcs, const int32_t *__restrict__ col,
                                                       const int32_t *__restrict__ eid, const uint32_t *__restrict__ rowpack,
                                                       const double *__restrict__ x, const uint8_t *__restrict__ bl,
                                                       Arc *__restrict__ arcs, uint32_t *__restrict__ arcs12) {
    const long long nquads = narcs >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nquads; q += stride) {
        const int4 c = reinterpret_cast<const int4 *>(col)[q];
        const int4 e = reinterpret_cast<const int4 *>(eid)[q];
        const uint4 r = reinterpret_cast<const uint4 *>(rowpack)[q];
        const int cs[4] = {c.x, c.y, c.z, c.w};
        const int es[4] = {e.x, e.y, e.z, e.w};
        const uint32_t rs[4] = {r.x, r.y, r.z, r.w};
        double ws[4];
        int bs[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            ws[k] = __ldg(x + es[k]);
            bs[k] = __ldg(bl + cs[k]);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            Arc a;
            a.nbr = cs[k] | (bs[k] ? (int32_t)0x80000000 : 0);
            a.row = rs[k];
            a.w = clamp01(ws[k]);
            arcs[4 * q + k] = a;
            if (arcs12) { // 12-byte copy {nbr, w} for the pairs engine (SMSGPU_ARCS12: 96 instead of 128 MB at config 5)
                const unsigned long long wb = (unsigned long long)__double_as_longlong(a.w);
                arcs12[3 * (4 * q + k)] = (uint32_t)a.nbr;
                arcs12[3 * (4 * q + k) + 1] = (uint32_t)wb;
                arcs12[3 * (4 * q + k) + 2] = (uint32_t)(wb >> 32);
            }
        }
    }
    // tail
    const long long t = 4 * nquads + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < narcs) {
        Arc a;
        const int c = col[t];
        a.nbr = c | (bl[c] ? (int32_t)0x80000000 : 0);
        a.row = rowpack[t];
        a.w = clamp01(x[eid[t]]);
        arcs[t] = a;
        if (arcs12) {
            const unsigned long long wb = (unsigned long long)__double_as_longlong(a.w);
            arcs12[3 * t] = (uint32_t)a.nbr;
            arcs12[3 * t + 1] = (uint32_t)wb;
            arcs12[3 * t + 2] = (uint32_t)(wb >> 32);
        }
    }
}

__global__ void initTablesKernel(unsigned char *scratch, const size_t tileStride, const long long ntiles,
                                 const unsigned tabCap) {
    const long long total = ntiles * (long long)tabCap;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long t = i / tabCap, s = i % tabCap;
        storeEntry(reinterpret_cast<Entry *>(scratch + t * tileStride) + s, emptyEntry());
    }
}

// ---- longest-first order of the next round's sources -------------------------------------------------------------------
// A source is ~1250 DEPENDENT steps on one (half-)warp; a 10 000-step source started near the end of the launch keeps
// the kernel alive for tens of milliseconds with the GPU almost idle.  Ball sizes change slowly from one LP solve to
// the next, so the sources of a round are issued in descending order of the work they did in the previous round.
// The unit of the order is a PAIR of consecutive local source indices (2k, 2k+1): in the handler's source list these
// are the two sides (v,0), (v,1) of one vertex, whose searches are mirror images of each other -- issued together they
// run on the two halves of one warp and read the same arc rows at the same time (measured: 70 ms of a 920 ms round
// between pairs kept together and pairs scattered).  1024 bins of 16 settled nodes; the order inside a bin is
// arbitrary (results do not depend on the issue order).
constexpr int kOrderBins = 1024;
constexpr int kOrderShift = 4; // 16 settled nodes per bin
// Pairs with a member that needed tier 2 or 3 this round ("heavy") sort in front of all others, in bins
// [0, kOrderBins); the rest in [kOrderBins, 2 kOrderBins).  The next round starts the heavy prefix in tier 2 at once, on
// a second stream beside tier 1, instead of letting it overflow out of tier 1 first and run as a tail.
__device__ __forceinline__ int orderBinOfPair(const uint32_t *__restrict__ work, const int8_t *__restrict__ tier, const long long k,
                                              const long long nsrc) {
    uint32_t w = work[2 * k];
    int t = tier[2 * k];
    if (2 * k + 1 < nsrc) {
        w = max(w, work[2 * k + 1]);
        t = max(t, (int)tier[2 * k + 1]);
    }
    const int b = kOrderBins - 1 - (int)min(w >> kOrderShift, (uint32_t)(kOrderBins - 1)); // bin 0 = heaviest
    return t >= 2 ? b : kOrderBins + b;
}
__global__ void orderHistKernel(const long long nsrc, const uint32_t *__restrict__ work, const int8_t *__restrict__ tier,
                                unsigned int *__restrict__ bins) {
    __shared__ unsigned int h[2 * kOrderBins];
    for (int b = threadIdx.x; b < 2 * kOrderBins; b += blockDim.x) h[b] = 0;
    __syncthreads();
    const long long npairs = nsrc >> 1; // an odd last source goes last (orderScatterKernel)
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npairs; k += (long long)gridDim.x * blockDim.x)
        atomicAdd(&h[orderBinOfPair(work, tier, k, nsrc)], 1u);
    __syncthreads();
    for (int b = threadIdx.x; b < 2 * kOrderBins; b += blockDim.x)
        if (h[b]) atomicAdd(&bins[b], h[b]);
}
// bins[0 .. 2 kOrderBins) counts -> bins[2 kOrderBins .. 4 kOrderBins) running cursors; bins[4 kOrderBins] = heavy pairs.
// One block of kOrderBins threads, two bins per thread.
__global__ void __launch_bounds__(kOrderBins) orderScanKernel(unsigned int *bins) {
    static_assert(kOrderBins == 1024, "one block of 1024 threads scans two bins per thread");
    __shared__ unsigned int warpSums[32];
    const int t = threadIdx.x;
    const unsigned int a = bins[2 * t], b = bins[2 * t + 1];
    unsigned int incl = a + b;
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if ((t & 31) >= o) incl += v;
    }
    if ((t & 31) == 31) warpSums[t >> 5] = incl;
    __syncthreads();
    if (t < 32) {
        const unsigned int w = warpSums[t];
        unsigned int wi = w;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int v = __shfl_up_sync(0xFFFFFFFFu, wi, o);
            if (t >= o) wi += v;
        }
        warpSums[t] = wi - w;
    }
    __syncthreads();
    const unsigned int excl = incl - (a + b) + warpSums[t >> 5];
    bins[2 * kOrderBins + 2 * t] = excl;
    bins[2 * kOrderBins + 2 * t + 1] = excl + a;
    if (2 * t == kOrderBins) bins[4 * kOrderBins] = excl;
}
__global__ void orderScatterKernel(const long long nsrc, const uint32_t *__restrict__ work, const int8_t *__restrict__ tier,
                                   unsigned int *__restrict__ bins, int32_t *__restrict__ order) {
    const long long npairs = nsrc >> 1;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npairs; k += (long long)gridDim.x * blockDim.x) {
        const unsigned int pos = atomicAdd(&bins[2 * kOrderBins + orderBinOfPair(work, tier, k, nsrc)], 1u);
        order[2 * (long long)pos] = (int32_t)(2 * k);
        order[2 * (long long)pos + 1] = (int32_t)(2 * k + 1);
    }
    if ((nsrc & 1) && blockIdx.x == 0 && threadIdx.x == 0) order[nsrc - 1] = (int32_t)(nsrc - 1);
}

// The overflow list of a one-source-per-warp tier is in completion order: the two sides of a vertex end up far apart
// and the pairs engine behind it could not mirror them.  This rebuilds the list from srcTier (< 0: not finished yet,
// or skipped -- the next tier skips those again): pairs with both members unfinished first, adjacent and aligned,
// then the single members (`singles`, appended by the host once the counts are known).
__global__ void pairCompactKernel(const long long nsrc, const int8_t *__restrict__ tier, int32_t *__restrict__ pairsOut,
                                  int32_t *__restrict__ singles, unsigned long long *__restrict__ counts) {
    const long long npairs = (nsrc + 1) >> 1;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npairs; k += (long long)gridDim.x * blockDim.x) {
        const bool u0 = tier[2 * k] < 0;
        const bool u1 = 2 * k + 1 < nsrc && tier[2 * k + 1] < 0;
        if (u0 && u1) {
            const unsigned long long pos = atomicAdd(&counts[0], 1ull);
            pairsOut[2 * pos] = (int32_t)(2 * k);
            pairsOut[2 * pos + 1] = (int32_t)(2 * k + 1);
        } else if (u0 || u1) {
            singles[atomicAdd(&counts[1], 1ull)] = (int32_t)(2 * k + (u1 ? 1 : 0));
        }
    }
}

// ---- SMSGPU_MODE_VERBATIM: the time-stamped blacklist (GraphDev::blStamp) -------------------------------------------
// stamps of a fresh sweep: -1 where the round's initial blacklist is set, never elsewhere
__global__ void initStampsKernel(const long long n, const uint8_t *__restrict__ bl, int32_t *stamp) {
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < n; v += (long long)gridDim.x * blockDim.x)
        stamp[v] = bl[v] ? -1 : kNeverStamp;
}
// The rest of blacklistIntReachable (ocw_seperator.cpp:363-382) for every empty source of the sweep at once: the stamp t
// of a source that found nothing floods from the vertices that carry it over integral edges (weight exactly 0.0 / 1.0)
// to vertices that are not blacklisted before t (previous sweep's stamps) and do not carry a smaller stamp yet.  One
// relaxation pass per launch; the host repeats it until nothing changes (min-label propagation).
__global__ void floodStampsKernel(const GraphDev G, unsigned int *changed) {
    bool any = false;
    for (long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x; u < G.n; u += (long long)gridDim.x * blockDim.x) {
        const int32_t t = G.blStampNext[u];
        if (t < 0 || t == kNeverStamp || !G.srcEmpty[t]) continue; // only the stamps of empty sources flood
        for (long long a = G.rowptr[u]; a < G.rowptr[u + 1]; ++a) {
            const Arc arc = G.arcs[a];
            const uint32_t v = (uint32_t)arc.nbr & 0x7FFFFFFFu;
            if (!(arc.w == 0.0 || arc.w == 1.0) || stampedBefore(G, v, t)) continue;
            if (atomicMin(&G.blStampNext[v], t) > t) any = true;
        }
    }
    if (any) *changed = 1u;
}
// Sources known to find nothing whatever the blacklist (they found nothing against the round's INITIAL blacklist, and a
// larger blacklist only removes walks) are not searched again in later sweeps; their writes are repeated here: nothing
// when the source is blacklisted at its time (ocw_seperator.cpp:334), else the seed of blacklistIntReachable.
__global__ void reseedEmptySourcesKernel(const GraphDev G, const long long lo, const long long hi, const int32_t *__restrict__ sources,
                                         const uint8_t *__restrict__ emptyB0) {
    for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (long long)gridDim.x * blockDim.x) {
        if (!emptyB0[i]) continue;
        const int32_t now = (int32_t)i;
        const uint32_t src = sources ? (uint32_t)sources[i] : (uint32_t)i;
        if ((long long)src < G.n && stampedBefore(G, src, now)) { // skipped at its time: no search, no writes
            G.srcEmpty[i] = 0;
            continue;
        }
        G.srcEmpty[i] = 1;
        const uint32_t u = src >> 1;
        atomicMin(&G.blStampNext[u], now);
        for (long long a = G.rowptr[u]; a < G.rowptr[u + 1]; ++a) {
            const Arc arc = G.arcs[a];
            const uint32_t v = (uint32_t)arc.nbr & 0x7FFFFFFFu;
            if ((arc.w == 0.0 || arc.w == 1.0) && !stampedBefore(G, v, now)) atomicMin(&G.blStampNext[v], now);
        }
    }
}
// work list of the later sweeps: every source that is not known to be empty
__global__ void activeSourcesKernel(const long long lo, const long long hi, const uint8_t *__restrict__ emptyB0, int32_t *list,
                                    unsigned int *count) {
    for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (long long)gridDim.x * blockDim.x)
        if (!emptyB0[i]) list[atomicAdd(count, 1u)] = (int32_t)i;
}
__global__ void iotaKernel(const long long n, int32_t *out) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) out[i] = (int32_t)i;
}
__global__ void compareStampsKernel(const long long n, const int32_t *__restrict__ a, const int32_t *__restrict__ b, unsigned int *differ) {
    bool d = false;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < n; v += (long long)gridDim.x * blockDim.x) d |= a[v] != b[v];
    if (d) *differ = 1u;
}

// ---- K2-K4 ---------------------------------------------------------------------------------------------------
#ifndef SMSGPU_FAST_MIN_CTAS
#define SMSGPU_FAST_MIN_CTAS 5
#endif
template <int TW, bool SMEM, bool FAST>
__global__ void __launch_bounds__(128, FAST ? SMSGPU_FAST_MIN_CTAS : 1) ocwSearchKernel(const GraphDev G, const SearchParams P, const RawOut R,
                                                        const TileCaps caps, unsigned char *scratch) {
    extern __shared__ __align__(32) unsigned char smem[];
    const Tile<TW> T;
    const int tilesPerCta = blockDim.x / TW;
    const int tileInCta = threadIdx.x / TW;
    const size_t tb = tileBytes(caps), hb = hotBytes(caps);
    TileMem M;
    if (FAST) {
        // working memory in HBM; the hot words (bucket counters, presence bitmap) in shared memory or behind the tile
        unsigned char *tile = scratch + ((size_t)blockIdx.x * tilesPerCta + tileInCta) * (tb + (P.hotInSmem ? 0 : hb));
        carveTileMem(M, tile, caps, P.hotInSmem ? smem + (size_t)tileInCta * hb : tile + tb);
    } else {
        carveTileMem(M, SMEM ? smem + (size_t)tileInCta * tb
                             : scratch + ((size_t)blockIdx.x * tilesPerCta + tileInCta) * tb,
                     caps);
    }
    if (SMEM) {
        for (uint32_t i = T.lane; i < caps.tabCap; i += TW) storeEntry(&M.tab[i], emptyEntry());
        T.sync();
    }
    if ((P.debugFlags & 1) && T.shift != 0) return;
    const long long total = P.worklist ? P.nwork : P.nsrc;
    long long accArcs = 0, accSettled = 0, accSources = 0, accSteps = 0, accLoads = 0;
    uint32_t sel[16];
    const int maxInter = min(P.maxInter, 16);

    while (true) {
        long long wi = 0;
        if (T.lane == 0) wi = (long long)atomicAdd(P.queueCounter, 1ull);
        wi = T.shfl(wi, 0);
        if (wi >= total) break;
        const int32_t srcIdx = P.worklist ? P.worklist[wi] : (int32_t)wi;
        const uint32_t src = P.sources ? (uint32_t)P.sources[srcIdx] : (uint32_t)(((((long long)(srcIdx >> 1)) * P.srcStride + P.srcBegin) << 1) | (srcIdx & 1));
        // `if (blacklist_[currentSource_]) return {}` -- indexed with the DOUBLED id (ocw_seperator.cpp:334, quirk A.7.1)
        if ((long long)src < G.n && (G.blStamp ? stampedBefore(G, src, srcIdx) : P.blacklist[src] != 0)) continue;
        if ((long long)(src >> 1) >= G.n) continue; // not a node of the double cover

        SrcState S;
        if (FAST)
            searchSourceFast<TW>(T, G, M, S, src, P.bestInit, P.stampBase + (uint32_t)wi + 1u, srcIdx);
        else
            searchSource<TW>(T, G, M, S, src, P.bestInit, srcIdx);
        if (!S.overflow && S.histN > 0) {
            int nsel = selectIntermediates<TW>(T, M, S, maxInter, sel);
            if (!S.overflow) { // (uniform: selectIntermediates decides on uniform values)
                Emitter E;
                E.R = &R;
                E.srcIdx = srcIdx;
                E.seq = 0;
                E.tier = P.tier;
#pragma unroll
                for (int r = 0; r < 16; ++r) sel[r] = T.shfl(sel[r], 0); // lane 0 holds the selection
                walksOfSource(G, M, S, E, sel, nsel, src, srcIdx);
            }
            S.overflow = T.shfl(S.overflow, 0);
        }
        if (!S.overflow && G.blStampNext) verbatimSourceDone<TW>(T, G, src, srcIdx, S.histN == 0);
        if (!S.overflow) {
            if (P.exportLabels) {
                // every label the search kept: the touched list (heap engine) / the occupied slots (fast engine)
                const uint32_t nscan = FAST ? caps.tabCap : S.touchedN;
                for (uint32_t i = T.lane; i < nscan; i += TW) {
                    if (FAST && !((M.bm[i >> 5] >> (i & 31)) & 1u)) continue;
                    const Entry e = loadEntry(&M.tab[FAST ? i : M.touched[i]]);
#pragma unroll
                    for (int sd = 0; sd < 2; ++sd) {
                        const double d = sd ? e.d1 : e.d0;
                        if (d >= kUnlabelled) continue;
                        const long long li = (long long)atomicAdd(&R.counters[5], 1ull);
                        if (li < R.labCap) {
                            R.labSrc[li] = srcIdx;
                            R.labNode[li] = (int32_t)(2 * e.key + sd);
                            R.labDist[li] = d;
                            R.labHops[li] = (int32_t)(sd ? e.h1 : e.h0);
                        }
                    }
                }
            }
            if (T.lane == 0) {
                R.srcTier[srcIdx] = (int8_t)P.tier;
                if (R.srcWork) R.srcWork[srcIdx] = (uint32_t)min(S.settled, 0xFFFFFFFFll);
                if (R.srcArcs) {
                    R.srcArcs[srcIdx] = S.arcs;
                    R.srcSettled[srcIdx] = S.settled;
                }
            }
            accArcs += S.arcs;
            accSettled += S.settled;
            if (FAST) {
                accSteps += S.steps;
                accLoads += S.loads;
            }
            accSources++;
        } else if (T.lane == 0) {
            const unsigned long long k = atomicAdd(P.ovfCount, 1ull);
            P.ovfList[k] = srcIdx;
        }
        // leave the table clean for the next source (the fast engine resets its bitmap when a source starts)
        T.sync();
        if (!FAST) {
            for (uint32_t i = T.lane; i < S.touchedN; i += TW) storeEntry(&M.tab[M.touched[i]], emptyEntry());
            T.sync();
        }
    }
    if (T.lane == 0 && accSources) {
        atomicAdd(&R.counters[2], (unsigned long long)accArcs);
        atomicAdd(&R.counters[3], (unsigned long long)accSettled);
        atomicAdd(&R.counters[P.runSlot], (unsigned long long)accSources);
        if (FAST) {
            atomicAdd(&R.counters[6], (unsigned long long)accSteps);
            atomicAdd(&R.counters[7], (unsigned long long)accLoads);
        }
    }
}

} // namespace smsgpu
