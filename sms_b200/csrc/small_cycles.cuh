// K6: the triangle-inequality / 4-hole scan.
//   smallCycleScanKernel   <- TriangleSeparator::getViolatedTriangles (include/sms/conshdlr/triangles.hpp:34-53),
//                             HoleSeparator::getViolatedFourHoles     (include/sms/conshdlr/holes.hpp:38-54),
//                             both = violatedOddSelectionClosedWalk   (src/graph/odd_closed_walk.cpp:8-65) on a fixed cycle
//   resolveCycleEidsKernel <- the per-call lpWeightedGraph_.weight(u, v) adjacency scans, done ONCE when the
//                             list is registered (addTriangle / addFourHole): the scan then gathers x* by edge id.
// One thread per four consecutive cycles; K edge ids (12 or 16 B) per cycle in, one code byte out; violated cycles are
// compacted in list order by the second pass.
#pragma once
#include "ocw_types.cuh"

namespace smsgpu {

template <int K>
__global__ void __launch_bounds__(256) resolveCycleEidsKernel(const long long T, const int32_t *__restrict__ nodes,
                                                              const long long *__restrict__ rowptr,
                                                              const int32_t *__restrict__ col,
                                                              const int32_t *__restrict__ eid, int32_t *__restrict__ out,
                                                              int *__restrict__ missing) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
#pragma unroll
    for (int i = 0; i < K; ++i) {
        const int32_t u = nodes[K * t + i], v = nodes[K * t + (i + 1) % K];
        int32_t id = -1;
        for (long long a = rowptr[u]; a < rowptr[u + 1]; ++a)
            if (col[a] == v) {
                id = eid[a];
                break;
            }
        if (id < 0) atomicAdd(missing, 1);
        out[K * t + i] = id;
    }
}

// violatedOddSelectionClosedWalk (src/graph/odd_closed_walk.cpp:8-65) on the K clamped weights of one cycle.
// Returns 0 when the cycle yields no violated inequality, else 0x80 | stay mask (bit i: edge i is a stay edge).
template <int K>
__device__ __forceinline__ unsigned oddSelectionCode(const double (&w)[K], double &length) {
    double curLength = 0;
    unsigned cross = 0;
    double lightestCross = 1, heaviestStay = 0;
    int lightestIdx = -1, heaviestIdx = -1, numCross = 0;
    bool none = false;
#pragma unroll
    for (int i = 0; i < K; ++i) {
        if (none) break;
        if (w[i] > 0.5) {
            curLength += 1 - w[i];
            cross |= 1u << i;
            numCross += 1;
            if (w[i] < lightestCross) {
                lightestCross = w[i];
                lightestIdx = i;
            }
        } else {
            curLength += w[i];
            if (w[i] > heaviestStay) {
                heaviestStay = w[i];
                heaviestIdx = i;
            }
        }
        if (curLength >= 1) none = true;
    }
    if (!none && (numCross % 2 == 0)) {
        if (lightestIdx >= 0 && (curLength + 2 * lightestCross - 1 < 1))
            cross &= ~(1u << lightestIdx);
        else if (heaviestIdx >= 0 && (curLength - 2 * heaviestStay + 1 < 1))
            cross |= 1u << heaviestIdx;
        else
            none = true;
    }
    if (none) return 0u;
    double len = 0;
#pragma unroll
    for (int i = 0; i < K; ++i) len += ((cross >> i) & 1u) ? 1 - w[i] : w[i];
    length = len;
    return 0x80u | (~cross & ((1u << K) - 1u));
}

// The scan is two passes over the list, kSmallTile cycles per CTA:
//   pass 1  K edge ids in (12 / 16 B per cycle, the only mandatory traffic), K gathers of x* (L2 resident), ONE code byte
//           out per cycle and one count per CTA;
//   (exclusive scan of the per-CTA counts: T / 1024 words)
//   pass 2  re-reads the code bytes, ranks the violated cycles inside the CTA (list order) and writes their walks.
constexpr int kSmallBlock = 256, kSmallItems = 4, kSmallTile = kSmallBlock * kSmallItems;

// The same decision without data-dependent branches (the scan's form: 32 cycles per warp take 32 different paths through the
// reference's loop).  Identical arithmetic in identical order: the reference's early `return {}` at the first prefix with
// curLength >= 1 equals one test of every prefix (nothing computed after it is used), and its running minimum / maximum with
// strict comparisons become selects.
template <int K>
__device__ __forceinline__ unsigned oddSelectionCodeFlat(const double (&w)[K]) {
    double cur = 0, lightestCross = 1, heaviestStay = 0;
    unsigned cross = 0;
    int li = -1, hi = -1;
    bool none = false;
#pragma unroll
    for (int i = 0; i < K; ++i) {
        const bool c = w[i] > 0.5;
        cur += c ? 1 - w[i] : w[i];
        cross |= (c ? 1u : 0u) << i;
        const bool lt = c && w[i] < lightestCross, gt = !c && w[i] > heaviestStay;
        lightestCross = lt ? w[i] : lightestCross;
        li = lt ? i : li;
        heaviestStay = gt ? w[i] : heaviestStay;
        hi = gt ? i : hi;
        none |= cur >= 1;
    }
    const bool even = !(__popc(cross) & 1);
    const bool fixL = li >= 0 && (cur + 2 * lightestCross - 1 < 1);
    const bool fixH = hi >= 0 && (cur - 2 * heaviestStay + 1 < 1);
    const unsigned crossL = cross & ~(1u << (li & 31)), crossH = cross | (1u << (hi & 31));
    cross = even ? (fixL ? crossL : (fixH ? crossH : cross)) : cross;
    none |= even && !fixL && !fixH;
    return none ? 0u : (0x80u | (~cross & ((1u << K) - 1u)));
}

// K = 3: the running minimum / maximum as a tournament over the three pairwise comparisons -- predicates and 32-bit selects
// only; the two values are picked once at the end.
__device__ __forceinline__ unsigned oddSelectionCodeFlat3(const double w0, const double w1, const double w2) {
    const bool c0 = w0 > 0.5, c1 = w1 > 0.5, c2 = w2 > 0.5;
    double cur = c0 ? 1 - w0 : w0;
    bool none = cur >= 1;
    cur += c1 ? 1 - w1 : w1;
    none |= cur >= 1;
    cur += c2 ? 1 - w2 : w2;
    none |= cur >= 1;
    const unsigned cross = (c0 ? 1u : 0u) | (c1 ? 2u : 0u) | (c2 ? 4u : 0u);
    // lightest cross edge: first index attaining the minimum among cross edges below the initial 1 (strict <)
    const bool l0 = c0 && w0 < 1, l1 = c1 && w1 < 1, l2 = c2 && w2 < 1;
    int li = l0 ? 0 : -1;
    li = (l1 && (li < 0 || w1 < w0)) ? 1 : li;
    li = (l2 && (li < 0 || (li == 1 ? w2 < w1 : w2 < w0))) ? 2 : li;
    // heaviest stay edge: first index attaining the maximum among stay edges above the initial 0 (strict >)
    const bool h0 = !c0 && w0 > 0, h1 = !c1 && w1 > 0, h2 = !c2 && w2 > 0;
    int hi = h0 ? 0 : -1;
    hi = (h1 && (hi < 0 || w1 > w0)) ? 1 : hi;
    hi = (h2 && (hi < 0 || (hi == 1 ? w2 > w1 : w2 > w0))) ? 2 : hi;
    const double lightestCross = li == 0 ? w0 : (li == 1 ? w1 : w2), heaviestStay = hi == 0 ? w0 : (hi == 1 ? w1 : w2);
    const bool even = !(__popc(cross) & 1);
    const bool fixL = li >= 0 && (cur + 2 * lightestCross - 1 < 1);
    const bool fixH = hi >= 0 && (cur - 2 * heaviestStay + 1 < 1);
    const unsigned crossL = cross & ~(1u << (li & 31)), crossH = cross | (1u << (hi & 31));
    const unsigned fin = even ? (fixL ? crossL : (fixH ? crossH : cross)) : cross;
    none |= even && !fixL && !fixH;
    return none ? 0u : (0x80u | (~fin & 7u));
}
template <int K>
__device__ __forceinline__ unsigned oddSelectionCodeScan(const double (&w)[K]) {
    if (K == 3) return oddSelectionCodeFlat3(w[0], w[1], w[2]);
    return oddSelectionCodeFlat<K>(w);
}
__global__ void __launch_bounds__(256) clampWeightsKernel(const long long m, const double *__restrict__ x, double *__restrict__ xc) {
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < m; e += (long long)gridDim.x * blockDim.x) xc[e] = clamp01(x[e]);
}

// One thread scans kSmallItems = 4 CONSECUTIVE cycles: their K * 4 edge ids arrive as K 128-bit loads, the four code bytes
// leave as one 32-bit store (round 2: 264 -> ~100 instructions per cycle; the kernel is bound by instruction issue, not by
// DRAM).  Lists whose edge ids / code bytes are not 16 / 4-byte aligned, and the last partial quad, take the scalar path.
// CLAMPED: x holds clamp01(x*) already (clampWeightsKernel; the context's scan), else the clamp is applied per gather.
template <int K, bool CLAMPED>
__global__ void __launch_bounds__(kSmallBlock) smallCycleScanKernel(const long long T, const int32_t *__restrict__ eids,
                                                                    const double *__restrict__ x, uint8_t *__restrict__ code,
                                                                    int32_t *__restrict__ tileCount) {
    const long long t0 = (long long)blockIdx.x * kSmallTile + (long long)threadIdx.x * kSmallItems;
    int mine = 0;
    const bool aligned = ((reinterpret_cast<uintptr_t>(eids) & 15u) == 0) && ((reinterpret_cast<uintptr_t>(code) & 3u) == 0);
    if (aligned && t0 + kSmallItems <= T) {
        int32_t e[K * kSmallItems];
        const int4 *src = reinterpret_cast<const int4 *>(eids + K * t0);
#pragma unroll
        for (int q = 0; q < K; ++q) {
            const int4 v = src[q];
            e[4 * q] = v.x;
            e[4 * q + 1] = v.y;
            e[4 * q + 2] = v.z;
            e[4 * q + 3] = v.w;
        }
        double wv[K * kSmallItems];
#pragma unroll
        for (int i = 0; i < K * kSmallItems; ++i) wv[i] = CLAMPED ? __ldg(x + e[i]) : clamp01(__ldg(x + e[i]));
        unsigned packed = 0;
#pragma unroll
        for (int it = 0; it < kSmallItems; ++it) {
            double w[K];
#pragma unroll
            for (int i = 0; i < K; ++i) w[i] = wv[K * it + i];
            const unsigned c = oddSelectionCodeScan<K>(w);
            packed |= c << (8 * it);
            mine += c ? 1 : 0;
        }
        *reinterpret_cast<uint32_t *>(code + t0) = packed;
    } else {
#pragma unroll
        for (int it = 0; it < kSmallItems; ++it) {
            const long long t = t0 + it;
            if (t >= T) continue;
            double w[K];
#pragma unroll
            for (int i = 0; i < K; ++i) w[i] = CLAMPED ? __ldg(x + eids[K * t + i]) : clamp01(__ldg(x + eids[K * t + i]));
            const unsigned c = oddSelectionCodeScan<K>(w);
            code[t] = (uint8_t)c;
            mine += c ? 1 : 0;
        }
    }
    // CTA total
    __shared__ int warpSum[kSmallBlock / 32];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, off);
    if ((threadIdx.x & 31) == 0) warpSum[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        int tot = 0;
#pragma unroll
        for (int k = 0; k < kSmallBlock / 32; ++k) tot += warpSum[k];
        tileCount[blockIdx.x] = tot;
    }
}

template <int K>
__global__ void __launch_bounds__(kSmallBlock) smallCycleGatherKernel(const long long T, const int32_t *__restrict__ nodes,
                                                                      const int32_t *__restrict__ eids,
                                                                      const double *__restrict__ x, const uint8_t *__restrict__ code,
                                                                      const long long *__restrict__ tileBase, int32_t *oNode,
                                                                      int32_t *oEid, uint8_t *oStay, int32_t *oNcross, int32_t *oSrc,
                                                                      double *oViol, long long *oOff) {
    __shared__ int warpSum[kSmallBlock / 32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const long long base = (long long)blockIdx.x * kSmallTile;
    const long long out0 = tileBase[blockIdx.x];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
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
        int before = carry;
        for (int k = 0; k < wrp; ++k) before += warpSum[k];
        if (v) {
            const long long k = out0 + before + incl - 1;
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
            double len = 0; // same summation order as the scan's
#pragma unroll
            for (int i = 0; i < K; ++i) len += ((sm >> i) & 1u) ? w[i] : 1 - w[i];
            oNcross[k] = nc;
            oSrc[k] = (int32_t)t;
            oViol[k] = 1.0 - len;
            oOff[k] = (long long)K * k;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0;
#pragma unroll
            for (int k = 0; k < kSmallBlock / 32; ++k) tot += warpSum[k];
            carry += tot;
        }
        __syncthreads();
    }
}
// oOff[U] = K * U
__global__ void smallCycleTailKernel(long long *oOff, const long long U, const int K) { oOff[U] = (long long)K * U; }

// ---- enumeration: replaces SmallChordlessCycles::computeSmallCCs (src/graph/small_ccs.cpp:18-43) -----------------------
//   for u ascending, for x in adj(u) with x > u, for y in adj(u) with y > x (both in adjacency order):
//     hasEdge(x, y)  -> triangle {x, u, y}
//     else for z in adj(x) (adjacency order) with z > u and !hasEdge(u, z): hasEdge(z, y) -> four-hole {x, u, y, z}
// The reference's list order is therefore lexicographic in (arc index of (u, x), position of y, position of z), and
// the arc index is the unit of work: one WARP per arc, lanes over the positions of y, 32 at a time, in order.  Two
// passes with the same traversal: COUNT (per arc, then an exclusive scan gives every arc its slice of the two lists)
// and FILL.  hasEdge is a binary search in a per-row SORTED copy of the adjacency (the reference scans the row).
__device__ __forceinline__ bool sortedHasEdge(const long long *__restrict__ rowptr, const int32_t *__restrict__ colSorted,
                                              const int32_t u, const int32_t v) {
    long long lo = rowptr[u], hi = rowptr[u + 1];
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        const int32_t c = colSorted[mid];
        if (c == v) return true;
        if (c < v)
            lo = mid + 1;
        else
            hi = mid;
    }
    return false;
}

template <bool FILL>
__global__ void __launch_bounds__(256) enumerateSmallCyclesKernel(const long long narcs, const int32_t *__restrict__ arcSrc,
                                                                  const long long *__restrict__ rowptr,
                                                                  const int32_t *__restrict__ col,
                                                                  const int32_t *__restrict__ colSorted,
                                                                  int32_t *__restrict__ triCount, int32_t *__restrict__ holeCount,
                                                                  const long long *__restrict__ triBase,
                                                                  const long long *__restrict__ holeBase,
                                                                  int32_t *__restrict__ tri, int32_t *__restrict__ holes,
                                                                  const int wantHoles) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long ai = warp; ai < narcs; ai += nwarps) {
        const int32_t u = arcSrc[ai], x = col[ai];
        if (x <= u) {
            if (!FILL && lane == 0) {
                triCount[ai] = 0;
                holeCount[ai] = 0;
            }
            continue;
        }
        long long tpos = FILL ? triBase[ai] : 0, hpos = FILL ? holeBase[ai] : 0;
        int tsum = 0, hsum = 0;
        const long long ub = rowptr[u], ue = rowptr[u + 1], xb = rowptr[x], xe = rowptr[x + 1];
        for (long long b0 = ub; b0 < ue; b0 += 32) {
            const long long b = b0 + lane;
            const int32_t y = b < ue ? col[b] : -1;
            int t = 0, h = 0;
            bool isTri = false;
            if (y > x) {
                isTri = sortedHasEdge(rowptr, colSorted, x, y);
                if (isTri) {
                    t = 1;
                } else if (wantHoles) {
                    for (long long c = xb; c < xe; ++c) {
                        const int32_t z = col[c];
                        if (z <= u || sortedHasEdge(rowptr, colSorted, u, z)) continue;
                        if (sortedHasEdge(rowptr, colSorted, z, y)) ++h;
                    }
                }
            }
            // exclusive prefix over the lanes: this lane's slice inside the chunk
            int tp = t, hp = h;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int a = __shfl_up_sync(0xFFFFFFFFu, tp, off), c = __shfl_up_sync(0xFFFFFFFFu, hp, off);
                if (lane >= off) {
                    tp += a;
                    hp += c;
                }
            }
            const int ttot = __shfl_sync(0xFFFFFFFFu, tp, 31), htot = __shfl_sync(0xFFFFFFFFu, hp, 31);
            if (FILL) {
                if (isTri) {
                    const long long o = 3 * (tpos + tp - t);
                    tri[o] = x;
                    tri[o + 1] = u;
                    tri[o + 2] = y;
                } else if (h > 0) {
                    long long o = 4 * (hpos + hp - h);
                    for (long long c = xb; c < xe; ++c) {
                        const int32_t z = col[c];
                        if (z <= u || sortedHasEdge(rowptr, colSorted, u, z)) continue;
                        if (sortedHasEdge(rowptr, colSorted, z, y)) {
                            holes[o] = x;
                            holes[o + 1] = u;
                            holes[o + 2] = y;
                            holes[o + 3] = z;
                            o += 4;
                        }
                    }
                }
                tpos += ttot;
                hpos += htot;
            } else {
                tsum += ttot;
                hsum += htot;
            }
        }
        if (!FILL && lane == 0) {
            triCount[ai] = tsum;
            holeCount[ai] = hsum;
        }
    }
}

} // namespace smsgpu
