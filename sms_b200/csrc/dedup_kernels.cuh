// K5: canonical form, hashing, duplicate removal and deterministic ordering of the violated cycles.
// The reference has no such step -- it hands every walk to SCIP and lets the cut pool drop duplicates
// (the same cycle is found from many sources, both sides, rotated and reflected: SURVEY.md A.7.2).
//
// Canonical form of a closed walk: among all rotations that start at a minimum vertex, in both directions,
// the lexicographically smallest sequence of (vertex, stay flag of the edge leaving it).
// Representative of a class of equal cycles: the one with the smallest (source index, emission index);
// the final list is ordered by that pair, so the output does not depend on scheduling.
#pragma once
#include "ocw_types.cuh"

namespace smsgpu {

// Input view: a list of closed walks in arena form.
struct CutView {
    const int32_t *node, *eid;
    const uint8_t *stay;
    const int32_t *off, *len, *src, *seq; // src: global source index
    const uint8_t *valid;                 // 0: ignore
    long long n;
};

struct DedupWork {
    // per input cut
    int32_t *cnode, *ceid; // canonical copies, same offsets as the input arena
    uint8_t *cstay;
    unsigned long long *hash;
    unsigned long long *rank;
    int32_t *slot;
    uint8_t *keep;
    int32_t *ncross;
    double *viol;
    // hash table
    unsigned long long *tabKey, *tabRank;
    int32_t *tabIdx;
    unsigned long long tabMask;
    unsigned long long *collisions;
};

__device__ __forceinline__ unsigned long long mix64(unsigned long long h, unsigned long long v) {
    h ^= v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
    h *= 0xFF51AFD7ED558CCDull;
    h ^= h >> 32;
    return h;
}

// One thread per cut: canonical rotation / direction, canonical copy, hash, #cross, violation.
// xByEid is the round's raw LP point; it is clamped here exactly as K1 clamps it.
// Cuts below the efficacy threshold are dropped here (keep = 0), before dedup and ordering.
__global__ void __launch_bounds__(128) canonKernel(const CutView V, const DedupWork D, const double *__restrict__ xByEid,
                                                   const int canonicalise, const int srcMul, const int srcAdd, const double minEfficacy) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V.n) return;
    if (!V.valid[i]) {
        D.keep[i] = 0;
        return;
    }
    const int L = V.len[i];
    const long long o = V.off[i];
    const int32_t *nd = V.node + o, *ed = V.eid + o;
    const uint8_t *st = V.stay + o;
    int bp = 0, bdir = 1;
    if (canonicalise) {
        int32_t mn = nd[0];
        for (int k = 1; k < L; ++k) mn = min(mn, nd[k]);
        bool have = false;
        for (int p = 0; p < L; ++p) {
            if (nd[p] != mn) continue;
            for (int dir = 1; dir >= -1; dir -= 2) {
                if (!have) {
                    bp = p;
                    bdir = dir;
                    have = true;
                    continue;
                }
                // compare candidate (p, dir) with best (bp, bdir)
                int cmp = 0;
                for (int k = 0; k < L && cmp == 0; ++k) {
                    int ia = p + dir * k, ib = bp + bdir * k;
                    ia = ((ia % L) + L) % L;
                    ib = ((ib % L) + L) % L;
                    const int ea = dir > 0 ? ia : (ia + L - 1) % L;
                    const int eb = bdir > 0 ? ib : (ib + L - 1) % L;
                    const int ka = nd[ia] * 2 + st[ea], kb = nd[ib] * 2 + st[eb];
                    cmp = (ka < kb) ? -1 : (ka > kb ? 1 : 0);
                }
                if (cmp < 0) {
                    bp = p;
                    bdir = dir;
                }
            }
        }
    }
    unsigned long long h = 0x243F6A8885A308D3ull ^ (unsigned long long)L;
    int nc = 0;
    double length = 0.;
    for (int k = 0; k < L; ++k) {
        int ia = bp + bdir * k;
        ia = ((ia % L) + L) % L;
        const int ea = bdir > 0 ? ia : (ia + L - 1) % L;
        const int32_t v = nd[ia], e = ed[ea];
        const int s = st[ea];
        D.cnode[o + k] = v;
        D.ceid[o + k] = e;
        D.cstay[o + k] = (uint8_t)s;
        h = mix64(h, ((unsigned long long)(uint32_t)v << 1) | (unsigned)s);
        nc += 1 - s;
        const double w = clamp01(xByEid[e]);
        length += s ? w : 1 - w;
    }
    D.hash[i] = h | 1ull; // 0 = empty slot
    D.ncross[i] = nc;
    D.viol[i] = 1.0 - length;
    // local -> global source index under the pair-cyclic deal of smsgpu_set_sources (identity for srcMul 1, srcAdd 0)
    const unsigned long long gsrc = ((((unsigned long long)V.src[i] >> 1) * srcMul + srcAdd) << 1) | ((unsigned)V.src[i] & 1u);
    D.rank[i] = (gsrc << 20) | (unsigned)V.seq[i];
    // efficacy pre-filter (smsgpu_params::min_efficacy): SCIP's efficacy of the row is (activity - rhs) / ||a||_2 =
    // (1 - length) / sqrt(len) (conshdlr_oscw.cpp:221,232 -> SCIPisCutEfficacious); relative slack 1e-9, SCIP decides the rest
    D.keep[i] = !(minEfficacy > 0.0) || (1.0 - length) > minEfficacy * (1.0 - 1e-9) * sqrt((double)L);
}

__global__ void __launch_bounds__(256) dedupInsertKernel(const long long n, const DedupWork D) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !D.keep[i]) return;
    const unsigned long long h = D.hash[i];
    unsigned long long s = (h * 0x9E3779B97F4A7C15ull) >> 20 & D.tabMask;
    while (true) {
        unsigned long long k = D.tabKey[s];
        if (k == 0) k = atomicCAS(&D.tabKey[s], 0ull, h);
        if (k == 0 || k == h) break;
        s = (s + 1) & D.tabMask;
    }
    atomicMin(&D.tabRank[s], D.rank[i]);
    D.slot[i] = (int32_t)s;
}
__global__ void __launch_bounds__(256) dedupMarkKernel(const long long n, const DedupWork D) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !D.keep[i]) return;
    const int32_t s = D.slot[i];
    if (D.tabRank[s] == D.rank[i]) D.tabIdx[s] = (int32_t)i;
}
// every non-representative compares itself with its representative: a 64-bit hash collision between two
// different cycles keeps both (and is counted) instead of silently dropping one
__global__ void __launch_bounds__(256) dedupVerifyKernel(const CutView V, const DedupWork D) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V.n || !D.keep[i]) return;
    const int32_t s = D.slot[i];
    const int32_t r = D.tabIdx[s];
    if (r == (int32_t)i) return;
    bool same = V.len[i] == V.len[r];
    if (same) {
        const long long oi = V.off[i], orr = V.off[r];
        for (int k = 0; k < V.len[i] && same; ++k)
            same = D.cnode[oi + k] == D.cnode[orr + k] && D.cstay[oi + k] == D.cstay[orr + k];
    }
    if (same)
        D.keep[i] = 0;
    else
        atomicAdd(D.collisions, 1ull);
}

// ---- deterministic order: by (global source index, emission index) ----------------------------------------------
__global__ void __launch_bounds__(256) countPerSrcKernel(const long long n, const DedupWork D, int32_t *cnt) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !D.keep[i]) return;
    atomicAdd(&cnt[D.rank[i] >> 20], 1);
}
__global__ void __launch_bounds__(256) scatterPerSrcKernel(const long long n, const DedupWork D, const long long *base,
                                                           int32_t *cursor, int32_t *order) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !D.keep[i]) return;
    const unsigned long long g = D.rank[i] >> 20;
    order[base[g] + atomicAdd(&cursor[g], 1)] = (int32_t)i;
}
__global__ void __launch_bounds__(256) sortSegmentsKernel(const long long nsrc, const DedupWork D, const long long *base,
                                                          int32_t *order, int32_t *outLen, const int32_t *inLen) {
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nsrc) return;
    const long long b = base[g], e = base[g + 1];
    for (long long a = b + 1; a < e; ++a) {
        const int32_t v = order[a];
        const unsigned long long rv = D.rank[v];
        long long c = a - 1;
        while (c >= b && D.rank[order[c]] > rv) {
            order[c + 1] = order[c];
            --c;
        }
        order[c + 1] = v;
    }
    for (long long a = b; a < e; ++a) outLen[a] = inLen[order[a]];
}

struct FinalCuts {
    long long *off; // U + 1
    int32_t *node, *eid;
    uint8_t *stay;
    int32_t *ncross, *src, *seq;
    double *viol;
};
// one thread per final cut (cycles are a handful of edges)
__global__ void __launch_bounds__(256) gatherFinalKernel(const long long U, const CutView V, const DedupWork D,
                                                         const int32_t *order, const FinalCuts F) {
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= U) return;
    const int32_t i = order[k];
    const long long o = V.off[i], fo = F.off[k];
    const int L = V.len[i];
    for (int q = 0; q < L; ++q) {
        F.node[fo + q] = D.cnode[o + q];
        F.eid[fo + q] = D.ceid[o + q];
        F.stay[fo + q] = D.cstay[o + q];
    }
    F.ncross[k] = D.ncross[i];
    F.viol[k] = D.viol[i];
    F.src[k] = (int32_t)(D.rank[i] >> 20);
    F.seq[k] = (int32_t)(D.rank[i] & 0xFFFFFu);
}

// ---- exclusive scan int32 -> int64 (3 phases, 1024 elements per block) -----------------------------------------
constexpr int kScanBlock = 256, kScanPer = 4, kScanTile = kScanBlock * kScanPer;
__device__ __forceinline__ long long blockScanExclusive(long long v, long long *total) {
    __shared__ long long warpSums[kScanBlock / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    long long inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const long long t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warpSums[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        long long w = lane < kScanBlock / 32 ? warpSums[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xFFFFFFFFu, w, o);
            if (lane >= o) w += t;
        }
        if (lane < kScanBlock / 32) warpSums[lane] = w;
    }
    __syncthreads();
    const long long prefix = wid ? warpSums[wid - 1] : 0;
    *total = warpSums[kScanBlock / 32 - 1];
    __syncthreads();
    return prefix + inc - v;
}
__global__ void __launch_bounds__(kScanBlock) scanReduceKernel(const int32_t *in, const long long n, long long *blockSums) {
    const long long b = (long long)blockIdx.x * kScanTile;
    long long v = 0;
    for (int k = 0; k < kScanPer; ++k) {
        const long long i = b + (long long)threadIdx.x * kScanPer + k;
        if (i < n) v += in[i];
    }
    long long total;
    blockScanExclusive(v, &total);
    if (threadIdx.x == 0) blockSums[blockIdx.x] = total;
}
__global__ void __launch_bounds__(kScanBlock) scanBlockSumsKernel(long long *blockSums, const long long nb) {
    long long carry = 0; // single block, walks the block sums in chunks
    for (long long c = 0; c < nb; c += kScanBlock) {
        const long long i = c + threadIdx.x;
        const long long v = i < nb ? blockSums[i] : 0;
        long long total;
        const long long ex = blockScanExclusive(v, &total);
        if (i < nb) blockSums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) blockSums[nb] = carry;
}
__global__ void __launch_bounds__(kScanBlock) scanFinalKernel(const int32_t *in, const long long n, const long long *blockSums,
                                                             long long *out) {
    const long long b = (long long)blockIdx.x * kScanTile;
    long long vals[kScanPer];
    long long v = 0;
    for (int k = 0; k < kScanPer; ++k) {
        const long long i = b + (long long)threadIdx.x * kScanPer + k;
        vals[k] = i < n ? in[i] : 0;
        v += vals[k];
    }
    long long total;
    long long ex = blockSums[blockIdx.x] + blockScanExclusive(v, &total);
    for (int k = 0; k < kScanPer; ++k) {
        const long long i = b + (long long)threadIdx.x * kScanPer + k;
        if (i < n) out[i] = ex;
        ex += vals[k];
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == kScanBlock - 1) out[n] = ex;
}

} // namespace smsgpu
