// K6: the triangle-inequality / 4-hole scan.
//   smallCycleScanKernel   <- TriangleSeparator::getViolatedTriangles (include/sms/conshdlr/triangles.hpp:34-53),
//                             HoleSeparator::getViolatedFourHoles     (include/sms/conshdlr/holes.hpp:38-54),
//                             both = violatedOddSelectionClosedWalk   (src/graph/odd_closed_walk.cpp:8-65) on a fixed cycle
//   resolveCycleEidsKernel <- the per-call lpWeightedGraph_.weight(u, v) adjacency scans, done ONCE when the
//                             list is registered (addTriangle / addFourHole): the scan then gathers x* by edge id.
// One thread per cycle; K edge ids (12 or 16 B) in, one flag word out; violated cycles are compacted in list order.
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

// flag[t] = 1 if violated; stayMask[t] bit i = edge i (nodes[i] -> nodes[i+1 mod K]) is a stay edge
template <int K>
__global__ void __launch_bounds__(256) smallCycleScanKernel(const long long T, const int32_t *__restrict__ eids,
                                                            const double *__restrict__ x, int32_t *__restrict__ flag,
                                                            uint8_t *__restrict__ stayMask, double *__restrict__ viol) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    double w[K];
#pragma unroll
    for (int i = 0; i < K; ++i) w[i] = clamp01(__ldg(x + eids[K * t + i]));
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
    flag[t] = none ? 0 : 1;
    if (!none) {
        stayMask[t] = (uint8_t)(~cross & ((1u << K) - 1u));
        double len = 0;
#pragma unroll
        for (int i = 0; i < K; ++i) len += ((cross >> i) & 1u) ? 1 - w[i] : w[i];
        viol[t] = 1.0 - len;
    }
}

template <int K>
__global__ void __launch_bounds__(256) smallCycleGatherKernel(const long long T, const int32_t *__restrict__ nodes,
                                                              const int32_t *__restrict__ eids,
                                                              const int32_t *__restrict__ flag,
                                                              const uint8_t *__restrict__ stayMask,
                                                              const double *__restrict__ viol,
                                                              const long long *__restrict__ pos, int32_t *oNode,
                                                              int32_t *oEid, uint8_t *oStay, int32_t *oNcross, int32_t *oSrc,
                                                              double *oViol, long long *oOff) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T || !flag[t]) return;
    const long long k = pos[t];
    const unsigned sm = stayMask[t];
    int nc = 0;
#pragma unroll
    for (int i = 0; i < K; ++i) {
        oNode[K * k + i] = nodes[K * t + i];
        oEid[K * k + i] = eids[K * t + i];
        const int s = (sm >> i) & 1u;
        oStay[K * k + i] = (uint8_t)s;
        nc += 1 - s;
    }
    oNcross[k] = nc;
    oSrc[k] = (int32_t)t;
    oViol[k] = viol[t];
    oOff[k] = (long long)K * k;
    if (k == pos[T] - 1) oOff[k + 1] = (long long)K * (k + 1);
}

} // namespace smsgpu
