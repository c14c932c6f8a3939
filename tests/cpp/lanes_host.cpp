// Test infrastructure: the sequential core of the lanes engine (sms_b200/csrc/ocw_lanes_core.cuh) compiled for the HOST
// and run on the CPU, source by source, so that its labels / counters / history can be compared with the oracle without
// a GPU (tests/test_lanes_core.py).  The arc records and the configuration are built the way smsgpu.cu builds them.
#include "ocw_lanes_core.cuh"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

using namespace smsgpu;

static uint32_t log2ceil(uint64_t v) {
    uint32_t l = 0;
    while ((1ull << l) < v) ++l;
    return l;
}

extern "C" int lanes_host_sweep(int64_t n, int64_t m, const int64_t *rowptr, const int32_t *col, const int32_t *eid, const double *x,
                                const uint8_t *bl, const int32_t *sources, int64_t nsrc, double bestInit, int32_t tabLgCap,
                                int32_t nbuckets, int32_t parkCap, int64_t *arcsOut, int64_t *settledOut, int64_t *histNOut,
                                int32_t *flagsOut, double *bestOut, int64_t *labOff, int64_t labCap, int32_t *labNode,
                                double *labDist, int32_t *labHops, int32_t *labPred) {
    // arc records (K1) with the neighbour's row packed
    uint32_t shift = 1; // smsgpu_create's rule
    while (shift < 31 && (1ll << shift) < std::max<int64_t>(2 * m, 2)) ++shift;
    std::vector<Arc> arcs((size_t)2 * m);
    std::vector<uint2> rowinfo((size_t)n);
    for (int64_t u = 0; u < n; ++u) rowinfo[u] = make_uint2((uint32_t)rowptr[u], (uint32_t)(rowptr[u + 1] - rowptr[u]));
    const uint32_t degMax = 0xFFFFFFFFu >> shift;
    for (int64_t a = 0; a < 2 * m; ++a) {
        const int32_t v = col[a];
        Arc r;
        r.nbr = v | (bl[v] ? (int32_t)0x80000000 : 0);
        const uint32_t deg = rowinfo[v].y;
        r.row = rowinfo[v].x | (std::min(deg, degMax) << shift);
        r.w = clamp01(x[eid[a]]);
        arcs[a] = r;
    }
    LanesCfg C{};
    const uint32_t touchedMax = (uint32_t)n + 32;
    const uint32_t touched = std::min<uint32_t>(1u << tabLgCap, touchedMax);
    const uint32_t tabLg = std::max<uint32_t>(log2ceil((uint64_t)2 * touched), 5);
    C.tabCap = 1u << tabLg;
    C.tabShift = 32 - tabLg;
    C.touchedCap = std::min<uint32_t>(C.tabCap / 4 * 3, touchedMax);
    C.histCap = 1u << 12;
    C.nbuckets = (uint32_t)nbuckets;
    C.qCap = (uint32_t)parkCap; // the queue-entry capacity takes the place of the old parking capacity
    C.rowShift = shift;
    {
        const double half = 0.5 * bestInit, L = 7.0, NBd = (double)nbuckets;
        C.scaleU = (float)(L / half);
        C.scaleB = (float)(NBd / (std::ldexp(1.0, 7) - 1.0));
    }
    const size_t tb = lanesTileBytes(C);
    std::vector<unsigned char> tile(tb + 64, 0);
    unsigned char *base = tile.data() + ((32 - ((uintptr_t)tile.data() & 31)) & 31);
    const LanesTile T = lanesCarve(base, C);
    int64_t nlab = 0;
    for (int64_t k = 0; k < nsrc; ++k) {
        labOff[k] = nlab;
        arcsOut[k] = settledOut[k] = histNOut[k] = 0;
        flagsOut[k] = -1;
        bestOut[k] = 0;
        const uint32_t s = (uint32_t)sources[k];
        if (((int64_t)s < n && bl[s]) || (int64_t)(s >> 1) >= n) continue; // quirk A.7.1: indexed with the doubled id
        LanesRec rec{};
        const uint32_t stamp = (uint32_t)k + 1u; // the table is never cleared
        lanesSearch(arcs.data(), rowinfo.data(), C, T, s, rowinfo[s >> 1], stamp, bestInit, rec);
        arcsOut[k] = rec.arcs;
        settledOut[k] = rec.settled;
        histNOut[k] = rec.histN;
        flagsOut[k] = (int32_t)rec.flags;
        bestOut[k] = rec.best;
        if (rec.flags & 1u) continue;
        for (uint32_t i = 0; i < C.tabCap; ++i) {
            const Entry &e = T.tab[i];
            if (e.flags != stamp) continue;
            for (int sd = 0; sd < 2; ++sd) {
                const double d = sd ? e.d1 : e.d0;
                if (d >= kUnlabelled) continue;
                if (nlab >= labCap) return -1;
                const uint32_t hw = sd ? e.h1 : e.h0;
                labNode[nlab] = (int32_t)(2 * e.key + sd);
                labDist[nlab] = d;
                labHops[nlab] = (int32_t)(hw & kLanesHopMask);
                labPred[nlab] = (int32_t)T.q[2 * (size_t)((hw >> kLanesHopBits) & kLanesQMask) + 1].z;
                ++nlab;
            }
        }
    }
    labOff[nsrc] = nlab;
    return 0;
}
