// "Runs" engine of K2 (tier 3, the last tier): one source per CTA, and the CTA settles a whole RUN of queue keys at once.
//
// Why.  An LP point of a cutting loop is 70 - 100 % exact 0 / 1 values, so in the double cover most arcs cost exactly 0 and
// a search is a breadth-first flood of zero-distance components: thousands of queue keys share the same (dist, hops)
// and differ only in the id.  The bucket engines of tiers 1 / 2 give such a source up at once (no bucket holds a
// thousand keys of equal distance); the one-key-at-a-time heap engine that used to be the last tier then settled them
// at ~2.5 us per node on a single warp (torus 100x100, root LP iterates: the round was the critical path of one
// 20 000-node search).  Measured on those inputs (oracle statistics): a search of 20 000 settled nodes
// has ~100 distinct (dist, hops) values, a random 20k-node graph with LP-like x* settles 22 000 nodes in 8.
//
// What.  runDijsktra of the reference (src/graph/ocw_seperator.cpp:37-112) in the canonical order (dist, hops, doubled id),
// event for event, like the other engines -- but advanced run by run.  A RUN is the set of all queue keys with the
// minimal (dist, hops) = (d, h), in id order.  Facts that make a run processable in parallel:
//   * every offer a run member makes has hops h + 1, so no relaxation of the run can put a key INTO the run or improve the
//     label of another member: the run is known completely when it starts;
//   * what one member's relaxations do to a neighbour depends only on that neighbour's label (both sides of a vertex sit
//     in one entry), so offers to DIFFERENT vertices commute; offers to the same vertex are applied in the sequential
//     order (member id, adjacency position, stay before cross) by one thread;
//   * the one thing that is a history -- bestDist / bestHopDist / the list of intermediates (:71-82) -- is a function of the
//     events (label improvements whose D = dist[t] + dist[t^1] can matter) in sequential order: (bestDist, bestHopDist) in front
//     of an event is the lexicographic minimum of (D, H) over the earlier ones, so the whole record logic is a parallel scan;
//   * the loop condition of :88-90 is tested before every extraction: the target's position in the run is known up front
//     (the run is cut there); `2 d > bestDist` can become true in the middle of a run only through the replay, in which
//     case the chunk of members being processed is rolled back (undo log) and redone up to that member -- it ends the search.
// A run is processed in chunks of at most kRunOfferCap arcs, in three parallel phases (runChunk): the offers, one thread per
// arc, dropped at once unless they can improve the target's label; the surviving offers of a target vertex applied in order by
// one thread; the record logic as a lexicographic prefix minimum over the events.
//
// The queue is an unsorted pool of keys in HBM with tombstones (a per-node position array lets a superseded key be killed
// directly), scanned by the whole CTA once per run for the minimal (dist, hops) and once to collect the members.
// Labels live in a dense table indexed by vertex (O(n) per CTA, reset through the touched list), laid out like the other
// engines' tiles, so that selection of intermediates, walk extraction and chord repair are the shared code of ocw_walks.cuh.
#pragma once
#include "ocw_search_fast.cuh"
#include "ocw_walks.cuh"

namespace smsgpu {

constexpr int kRunThreads = 256;
constexpr uint32_t kRunOfferCap = 1024; // arcs per chunk (sort buffers in shared memory)

#ifdef SMSGPU_RUNS_CLOCKS // instrumentation build: cycles per phase of a chunk, accumulated by thread 0 of CTA 0
#define RCLK(i_)                                                                     \
    do {                                                                             \
        if (threadIdx.x == 0 && blockIdx.x == 0) {                                   \
            const long long now_ = clock64();                                        \
            atomicAdd(&g_runClk[i_], (unsigned long long)(now_ - rclkLast));         \
            rclkLast = now_;                                                         \
        }                                                                            \
    } while (0)
#define RCLK_DECL long long rclkLast = clock64()
#define RCLK_RESET rclkLast = clock64()
#else
#define RCLK(i_) ((void)0)
#define RCLK_DECL ((void)0)
#define RCLK_RESET ((void)0)
#endif


// per-CTA scratch behind the tile (global memory)
struct RunScratch {
    uint32_t *run;     // [runCap] members of the current run (doubled ids, sorted)
    uint32_t *qpos;    // [2n] pool position of a node's live key, kNoQpos
    uint32_t *bits;    // [(2n + 31) / 32] bitmap over the doubled ids (sorting long runs); all zero between uses
    uint32_t *undoVtx; // [cap] per target vertex of the chunk: undo log (label, predecessors, queue positions before the chunk)
    Entry *undoEntry;
    uint2 *undoPred, *undoQpos;
    uint32_t runCap;
};
__host__ __device__ inline size_t runScratchBytes(const long long n) {
    size_t rc = 2;
    while (rc < (size_t)(2 * n + 2)) rc <<= 1; // power of two: the member list is sorted with a bitonic network
    size_t b = rc * 4 + (size_t)(2 * n + 2) * 4 + (size_t)((2 * n + 31) / 32 + 1) * 4;
    b += (size_t)kRunOfferCap * (4 + 32 + 8 + 8);
    return (b + 255) & ~(size_t)255;
}
__device__ inline void carveRunScratch(RunScratch &X, unsigned char *p, const long long n) {
    uint32_t rc = 2;
    while (rc < (uint32_t)(2 * n + 2)) rc <<= 1;
    X.runCap = rc;
    // 16-byte members first
    X.undoEntry = reinterpret_cast<Entry *>(p);
    p += (size_t)kRunOfferCap * 32;
    X.undoPred = reinterpret_cast<uint2 *>(p);
    p += (size_t)kRunOfferCap * 8;
    X.undoQpos = reinterpret_cast<uint2 *>(p);
    p += (size_t)kRunOfferCap * 8;
    X.run = reinterpret_cast<uint32_t *>(p);
    p += (size_t)rc * 4;
    X.qpos = reinterpret_cast<uint32_t *>(p);
    p += (size_t)(2 * n + 2) * 4;
    X.bits = reinterpret_cast<uint32_t *>(p);
    p += (size_t)((2 * n + 31) / 32 + 1) * 4;
    X.undoVtx = reinterpret_cast<uint32_t *>(p);
}

// state of the search shared by the CTA
struct RunShared {
    double best, runD;
    unsigned long long minHi, minLo;
    uint32_t bestHop, histN, poolN, touchedN, runLen, runH, nev, nseg;
    long long arcs, settled;
    int overflow, done, cutoff;
    uint32_t src;
    int32_t srcIdx;
    long long wi;
    int skip;
};

// block-wide bitonic sort of n (a power of two) 32-bit keys in global memory, ascending (member lists too long for shared memory)
__device__ inline void bitonicSortGlobal32(uint32_t *k, const uint32_t n) {
    for (uint32_t size = 2; size <= n; size <<= 1) {
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (uint32_t t = threadIdx.x; t < (n >> 1); t += blockDim.x) {
                const uint32_t lo = 2 * t - (t & (stride - 1));
                const uint32_t hi = lo + stride;
                const bool up = (lo & size) == 0;
                const uint32_t a = k[lo], b = k[hi];
                if ((a > b) == up) {
                    k[lo] = b;
                    k[hi] = a;
                }
            }
        }
    }
    __syncthreads();
}
__device__ __forceinline__ uint32_t nextPow2(uint32_t v) {
    uint32_t p = 2;
    while (p < v) p <<= 1;
    return p;
}

// Shared memory of a CTA (dynamic): everything a chunk needs between its phases.
constexpr uint32_t kRunHashSlots = 2 * kRunOfferCap;
constexpr uint32_t kRunHashShift = 32 - 11; // log2(kRunHashSlots) = 11
static_assert(kRunHashSlots == 2048, "kRunHashShift assumes 2048 slots");
constexpr uint16_t kRunNil = 0xFFFFu;
struct RunSmem {
    double sW[kRunOfferCap];              // surviving offers: stay weight of the arc
    double evD[2 * kRunOfferCap];         // events, slot 2 j + (cross ? 1 : 0) of arc j: D (8.0 = no event); j order = sequential order
    uint32_t evH[2 * kRunOfferCap];       //   H
    uint32_t evNode[2 * kRunOfferCap];    //   doubled id of the improved node
    uint32_t offs[kRunOfferCap + 1];      // exclusive prefix of the degrees of the chunk's members
    uint32_t mCur[kRunOfferCap];          // per member: doubled id
    uint32_t mRow[kRunOfferCap];          //             first arc of its row
    uint32_t sNbr[kRunOfferCap];          // surviving offers: target vertex
    uint32_t hKey[kRunHashSlots];         // offers grouped by target vertex: open addressing, kEmptyKey = free
    uint16_t hHead[kRunHashSlots];        //   head of the vertex' list of surviving offers
    uint16_t sNext[kRunOfferCap];         //   next offer of the same vertex
    uint16_t sRank[kRunOfferCap];         // surviving offers: rank of the member that makes it
    uint16_t sJ[kRunOfferCap];            //                   arc index in the chunk (the sequence number)
    uint16_t gSlot[kRunOfferCap];         // hash slots in use (one per target vertex with a surviving offer)
    uint16_t aRank[kRunOfferCap];         // per arc of the chunk: rank of its member
};

struct LexMin { // (D, H) lexicographic minimum: bestDist / bestHopDist after a sequence of events (:73-76)
    double d;
    uint32_t h;
};
__device__ __forceinline__ LexMin lexmin(const LexMin a, const LexMin b) { return (b.d < a.d || (b.d == a.d && b.h < a.h)) ? b : a; }

// One chunk of the current run: the members with ranks [0, nm) described by S.mCur / S.mRow / S.offs -- or, when arcHi > arcLo,
// the arcs [arcLo, arcHi) of the single member of rank 0 (a row longer than the chunk capacity).
//   (1) one thread per arc: the offer, tested against the target's CURRENT label.  Labels only improve, so an offer that cannot
//       improve the label now never will: it is dropped here (most arcs of a flood point at nodes that are labelled already).
//       Survivors are hung into a shared-memory hash list of their target vertex.
//   (2) one thread per target vertex: updatePaths (:59-84) for its surviving offers in sequential order (arc index in the chunk =
//       (member id, adjacency position) order); label, predecessor, queue key; events into the slots of their arcs.
//   (3) the record logic (:71-82): bestDist / bestHopDist in front of an event is the lexicographic prefix minimum of (D, H) over
//       the event slots -- a parallel scan; an event is recorded iff D < bestDist or D == bestDist and H < 2 bestHopDist; the recorded
//       ones are appended to the list of intermediates in slot order.  The same scan tells where 2 d > bestDist becomes true.
// sh.cutoff: rank of the member behind which the search ends when later members of the chunk have to be undone, else -1.
__device__ inline void runChunk(const GraphDev &G, TileMem &M, RunScratch &X, RunShared &sh, RunSmem &S, const uint32_t nm,
                                const uint32_t arcLo, const uint32_t arcHi, const uint32_t noffers, const int32_t now,
                                const uint32_t snapPool, const bool lastOfRun) {
    const double d = sh.runD;
    const uint32_t hNew = sh.runH + 1;
    const double best0 = sh.best; // conservative (>= the running value) for pruning / queue admission, as in the other engines
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    __shared__ uint32_t nSurv, nGroups, firstDrop, chunkArcs;
    __shared__ double wD[kRunThreads / 32];
    __shared__ uint32_t wH[kRunThreads / 32], wC[kRunThreads / 32];
    RCLK_DECL;
    for (uint32_t i = threadIdx.x; i < kRunHashSlots; i += blockDim.x) {
        S.hKey[i] = kEmptyKey;
        S.hHead[i] = kRunNil;
    }
    if (threadIdx.x == 0) {
        nSurv = 0;
        chunkArcs = 0;
        nGroups = 0;
        firstDrop = 0xFFFFFFFFu;
        sh.nseg = 0;
        sh.cutoff = -1;
    }
    __syncthreads();
    // ---- (1) offers ---------------------------------------------------------------------------------------------------------
    // kRunOfferCap / kRunThreads arcs per thread, the loads of all of them in flight together: arc records first, then the labels
    int myArcs = 0;
    {
        constexpr int IA = kRunOfferCap / kRunThreads;
        uint32_t rr[IA], nb[IA];
        double ww[IA];
        bool act[IA];
#pragma unroll
        for (int i = 0; i < IA; ++i) {
            const uint32_t j = threadIdx.x + i * kRunThreads;
            act[i] = false;
            rr[i] = 0;
            nb[i] = 0;
            ww[i] = 0.0;
            if (j >= noffers) continue;
            uint32_t r = 0, a = arcLo + j;
            if (arcHi <= arcLo) { // member of arc j
                r = S.aRank[j];
                a = j - S.offs[r];
            }
            S.evD[2 * j] = 8.0; // no event (larger than any D)
            S.evD[2 * j + 1] = 8.0;
            const Arc arc = G.arcs[S.mRow[r] + a];
            rr[i] = r;
            nb[i] = (uint32_t)arc.nbr & 0x7FFFFFFFu;
            ww[i] = arc.w;
            act[i] = arc.nbr >= 0; // blacklisted neighbours are skipped (:98-99)
        }
        Entry ee[IA];
#pragma unroll
        for (int i = 0; i < IA; ++i) {
            if (G.blStamp && act[i]) act[i] = !stampedBefore(G, nb[i], now);
            if (act[i]) ee[i] = loadEntry(&M.tab[nb[i]]);
        }
#pragma unroll
        for (int i = 0; i < IA; ++i) {
            if (!act[i]) continue;
            myArcs++;
            const uint32_t j = threadIdx.x + i * kRunThreads;
            const Entry &e = ee[i];
            const uint32_t nbr = nb[i];
            const uint32_t side = S.mCur[rr[i]] & 1u;
            const double w = ww[i];
            const double dS = d + w, dX = d + (1.0 - w);
            const double dA = side ? e.d1 : e.d0, dB = side ? e.d0 : e.d1;
            const uint32_t hA = side ? e.h1 : e.h0, hB = side ? e.h0 : e.h1;
            // can the offer be accepted at all?  A side without a label accepts iff the label could still matter (label pruning, see
            // ocw_search.cuh): d_offer + d <= best or d_offer + dist[twin] <= best, where the twin's label can only fall to >= d here.
            const bool sS = dA >= kUnlabelled ? (dS < kInfDist && dS + fmin(dB, d) <= best0) : ((dA == dS) ? (hNew < hA) : (dS < dA));
            const bool sX = dB >= kUnlabelled ? (dX < kInfDist && dX + fmin(dA, d) <= best0) : ((dB == dX) ? (hNew < hB) : (dX < dB));
            if (!(sS || sX)) continue;
            const uint32_t sv = atomicAdd(&nSurv, 1u);
            S.sNbr[sv] = nbr;
            S.sW[sv] = w;
            S.sRank[sv] = (uint16_t)rr[i];
            S.sJ[sv] = (uint16_t)j;
            uint32_t hs = (nbr * 0x9E3779B1u) >> kRunHashShift;
            while (true) {
                const uint32_t old = atomicCAS(&S.hKey[hs], kEmptyKey, nbr);
                if (old == kEmptyKey) S.gSlot[atomicAdd(&nGroups, 1u)] = (uint16_t)hs; // first offer to this vertex in the chunk
                if (old == kEmptyKey || old == nbr) break;
                hs = (hs + 1) & (kRunHashSlots - 1);
            }
            // push onto the vertex' list (16-bit exchange through the containing word)
            uint32_t *word = reinterpret_cast<uint32_t *>(S.hHead) + (hs >> 1);
            const uint32_t shft = (hs & 1u) * 16u;
            uint32_t seen = *word, want;
            do {
                want = (seen & ~(0xFFFFu << shft)) | (sv << shft);
                const uint32_t got = atomicCAS(word, seen, want);
                if (got == seen) break;
                seen = got;
            } while (true);
            S.sNext[sv] = (uint16_t)((seen >> shft) & 0xFFFFu);
        }
    }
    // arcs scanned: one 32-bit shared-memory atomic per warp (a 64-bit one is a compare-and-swap loop: it was 16 % of the
    // kernel's stall samples, profiles/r02_runs_tier3_random20k_lplike_4k.md)
    myArcs = __reduce_add_sync(0xFFFFFFFFu, myArcs);
    if (lane == 0 && myArcs) atomicAdd(&chunkArcs, (uint32_t)myArcs);
    __syncthreads();
    if (threadIdx.x == 0) sh.arcs += chunkArcs;
    RCLK(0);
    // ---- (2) one thread per target vertex -----------------------------------------------------------------------------------
    const uint32_t ngroups = nGroups;
    for (uint32_t g = threadIdx.x; g < ngroups; g += blockDim.x) {
        const uint32_t hs = S.gSlot[g];
        const uint32_t v = S.hKey[hs];
        Entry e = loadEntry(&M.tab[v]);
        const Entry old = e;
        const uint2 oldPred = make_uint2(M.pred[2 * v], M.pred[2 * v + 1]);
        const bool fresh = e.key == kEmptyKey;
        uint32_t pr0 = oldPred.x, pr1 = oldPred.y;
        bool ch0 = false, ch1 = false;
        int last = -1;
        while (true) { // the vertex' offers in ascending arc index
            uint32_t bs = kRunNil, bj = 0xFFFFFFFFu;
            for (uint32_t q = S.hHead[hs]; q != kRunNil; q = S.sNext[q]) {
                const uint32_t jq = S.sJ[q];
                if ((int)jq > last && jq < bj) {
                    bj = jq;
                    bs = q;
                }
            }
            if (bs == kRunNil) break;
            last = (int)bj;
            const uint32_t cur = S.mCur[S.sRank[bs]];
            const uint32_t side = cur & 1u;
            const double w = S.sW[bs];
            const double dS = d + w, dX = d + (1.0 - w);
            const bool creatS = dS + d <= best0, creatX = dX + d <= best0; // label pruning, see ocw_search.cuh
            double dA = side ? e.d1 : e.d0, dB = side ? e.d0 : e.d1;
            uint32_t hA = side ? e.h1 : e.h0, hB = side ? e.h0 : e.h1;
            bool accS, accX;
            if (dA >= kUnlabelled)
                accS = dS < kInfDist && (creatS || dS + dB <= best0);
            else
                accS = (dA == dS) ? (hNew < hA) : (dS < dA);
            if (accS) {
                dA = dS;
                hA = hNew;
                const double D = dA + dB;
                if (D <= best0) {
                    S.evD[2 * bj] = D;
                    S.evH[2 * bj] = hA + hB;
                    S.evNode[2 * bj] = 2 * v + side;
                }
            }
            if (dB >= kUnlabelled)
                accX = dX < kInfDist && (creatX || dX + dA <= best0);
            else
                accX = (dB == dX) ? (hNew < hB) : (dX < dB);
            if (accX) {
                dB = dX;
                hB = hNew;
                const double D = dB + dA;
                if (D <= best0) {
                    S.evD[2 * bj + 1] = D;
                    S.evH[2 * bj + 1] = hB + hA;
                    S.evNode[2 * bj + 1] = 2 * v + (side ^ 1u);
                }
            }
            if (accS || accX) {
                if (side) {
                    e.d1 = dA;
                    e.h1 = hA;
                    e.d0 = dB;
                    e.h0 = hB;
                } else {
                    e.d0 = dA;
                    e.h0 = hA;
                    e.d1 = dB;
                    e.h1 = hB;
                }
                if (accS) {
                    (side ? pr1 : pr0) = cur;
                    (side ? ch1 : ch0) = true;
                }
                if (accX) {
                    (side ? pr0 : pr1) = cur;
                    (side ? ch0 : ch1) = true;
                }
            }
        }
        if (!(ch0 || ch1)) continue;
        // undo log of this vertex (a chunk is rolled back when the search ends in its middle)
        const uint32_t sidx = atomicAdd(&sh.nseg, 1u);
        X.undoVtx[sidx] = v;
        X.undoEntry[sidx] = old;
        X.undoPred[sidx] = oldPred;
        const uint2 oldQ = fresh ? make_uint2(kNoQpos, kNoQpos) : make_uint2(X.qpos[2 * v], X.qpos[2 * v + 1]);
        X.undoQpos[sidx] = oldQ;
        e.key = v;
        if (fresh) {
            e.flags = 0;
            const uint32_t ti = atomicAdd(&sh.touchedN, 1u);
            if (ti < M.caps.touchedCap)
                M.touched[ti] = v;
            else
                sh.overflow = 1;
        }
        storeEntry(&M.tab[v], e);
        M.pred[2 * v] = pr0;
        M.pred[2 * v + 1] = pr1;
        // queue: the superseded key dies, the new one is appended if the node can still be extracted
        uint32_t nq0 = oldQ.x, nq1 = oldQ.y;
#pragma unroll
        for (int sd = 0; sd < 2; ++sd) {
            if (!(sd ? ch1 : ch0)) continue;
            const uint32_t oq = sd ? oldQ.y : oldQ.x;
            if (oq != kNoQpos) M.heap[oq].hi = kDeadKey;
            const double dn = sd ? e.d1 : e.d0;
            uint32_t nq = kNoQpos;
            if (2 * dn <= best0) {
                nq = atomicAdd(&sh.poolN, 1u);
                if (nq < M.caps.heapCap) {
                    HeapEnt ke;
                    ke.hi = (unsigned long long)__double_as_longlong(dn);
                    ke.lo = ((unsigned long long)hNew << 32) | (2 * v + sd);
                    M.heap[nq] = ke;
                } else {
                    sh.overflow = 1;
                    nq = kNoQpos;
                }
            }
            (sd ? nq1 : nq0) = nq;
        }
        X.qpos[2 * v] = nq0;
        X.qpos[2 * v + 1] = nq1;
    }
    __syncthreads();
    RCLK(1);
    // ---- (3) the record logic over the event slots, in slot order ----------------------------------------------------------
    constexpr int IT = 2 * kRunOfferCap / kRunThreads; // slots per thread
    const uint32_t nslots = 2 * noffers;
    LexMin run = {sh.best, sh.bestHop}; // (bestDist, bestHopDist) in front of the tile
    uint32_t histBase = sh.histN;
    uint32_t pushMask = 0, pushPos = 0;
    {
        double dd[IT];
        uint32_t hh[IT];
        LexMin agg = {8.0, 0xFFFFFFFFu};
#pragma unroll
        for (int i = 0; i < IT; ++i) {
            const uint32_t q = threadIdx.x * IT + i;
            dd[i] = q < nslots ? S.evD[q] : 8.0;
            hh[i] = dd[i] < 8.0 ? S.evH[q] : 0xFFFFFFFFu;
            agg = lexmin(agg, LexMin{dd[i], hh[i]});
        }
        // exclusive prefix (lexicographic minimum) over the threads of the CTA
        LexMin incl = agg;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            LexMin o;
            o.d = __shfl_up_sync(0xFFFFFFFFu, incl.d, off);
            o.h = __shfl_up_sync(0xFFFFFFFFu, incl.h, off);
            if (lane >= off) incl = lexmin(o, incl);
        }
        LexMin excl;
        excl.d = __shfl_up_sync(0xFFFFFFFFu, incl.d, 1);
        excl.h = __shfl_up_sync(0xFFFFFFFFu, incl.h, 1);
        if (lane == 0) excl = LexMin{8.0, 0xFFFFFFFFu};
        if (lane == 31) {
            wD[wrp] = incl.d;
            wH[wrp] = incl.h;
        }
        __syncthreads();
        LexMin before = lexmin(run, excl);
        for (int w = 0; w < wrp; ++w) before = lexmin(before, LexMin{wD[w], wH[w]});
        int nc = 0;
#pragma unroll
        for (int i = 0; i < IT; ++i) {
            if (dd[i] < 8.0) {
                // `if (D < bestDist || (D == bestDist && H < bestHopDist)) {...; push} else if (D == bestDist && H < 2 bestHopDist) push`
                const bool push = dd[i] < before.d || (dd[i] == before.d && hh[i] < 2 * before.h);
                if (push) {
                    pushMask |= 1u << i;
                    nc++;
                }
                before = lexmin(before, LexMin{dd[i], hh[i]});
                if (2 * d > before.d) atomicMin(&firstDrop, threadIdx.x * IT + i); // the loop condition (:90) fails behind this member
            }
        }
        int cincl = nc;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int o = __shfl_up_sync(0xFFFFFFFFu, cincl, off);
            if (lane >= off) cincl += o;
        }
        if (lane == 31) wC[wrp] = (uint32_t)cincl;
        __syncthreads();
        pushPos = (uint32_t)(cincl - nc);
        for (int w = 0; w < wrp; ++w) pushPos += wC[w];
        if (threadIdx.x == 0) {
            LexMin tot = run;
            uint32_t np = 0;
            for (int w = 0; w < kRunThreads / 32; ++w) {
                tot = lexmin(tot, LexMin{wD[w], wH[w]});
                np += wC[w];
            }
            int cutoff = -1;
            bool drop = false;
            if (firstDrop != 0xFFFFFFFFu) {
                drop = true;
                if (arcHi <= arcLo) { // rank of the member whose event made 2 d > bestDist true
                    const uint32_t j = firstDrop >> 1;
                    uint32_t lo = 0, hi = nm;
                    while (hi - lo > 1) {
                        const uint32_t mid = (lo + hi) >> 1;
                        if (S.offs[mid] <= j)
                            lo = mid;
                        else
                            hi = mid;
                    }
                    if (lo + 1 < nm) cutoff = (int)lo; // later members of this chunk were never extracted: roll back, redo up to here
                }
            }
            if (cutoff < 0) {
                if (histBase + np > M.caps.histCap) {
                    sh.overflow = 1;
                } else {
                    sh.best = tot.d;
                    sh.bestHop = tot.h;
                    sh.histN = histBase + np;
                    if (drop && !lastOfRun) sh.done = 1; // the next member of the run is never extracted
                }
            }
            sh.cutoff = cutoff;
        }
        __syncthreads();
        if (sh.cutoff < 0 && !sh.overflow) { // append the recorded events to the list of intermediates, in slot order
#pragma unroll
            for (int i = 0; i < IT; ++i)
                if (pushMask & (1u << i)) M.hist[histBase + pushPos++] = S.evNode[threadIdx.x * IT + i];
        }
    }
    __syncthreads();
    RCLK(2);
    if (sh.cutoff >= 0) { // roll the chunk back; the caller redoes the members up to rank `cutoff` and ends the search
        const uint32_t nseg = sh.nseg;
        for (uint32_t k = threadIdx.x; k < nseg; k += blockDim.x) {
            const uint32_t v = X.undoVtx[k];
            const Entry old = X.undoEntry[k];
            storeEntry(&M.tab[v], old);
            M.pred[2 * v] = X.undoPred[k].x;
            M.pred[2 * v + 1] = X.undoPred[k].y;
            const uint2 oq = X.undoQpos[k];
            X.qpos[2 * v] = oq.x;
            X.qpos[2 * v + 1] = oq.y;
            if (oq.x != kNoQpos) { // revive the keys this chunk killed
                HeapEnt ke;
                ke.hi = (unsigned long long)__double_as_longlong(old.d0);
                ke.lo = ((unsigned long long)old.h0 << 32) | (2 * v);
                M.heap[oq.x] = ke;
            }
            if (oq.y != kNoQpos) {
                HeapEnt ke;
                ke.hi = (unsigned long long)__double_as_longlong(old.d1);
                ke.lo = ((unsigned long long)old.h1 << 32) | (2 * v + 1);
                M.heap[oq.y] = ke;
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) sh.poolN = snapPool;
        __syncthreads();
    }
    RCLK(3);
}

#ifndef SMSGPU_RUNS_MIN_CTAS
#define SMSGPU_RUNS_MIN_CTAS 2
#endif
__global__ void __launch_bounds__(kRunThreads, SMSGPU_RUNS_MIN_CTAS) ocwSearchRunsKernel(const GraphDev G, const SearchParams P, const RawOut R, const TileCaps caps,
                                                                   unsigned char *scratch) {
    extern __shared__ __align__(16) unsigned char runSmemRaw[];
    RunSmem &S = *reinterpret_cast<RunSmem *>(runSmemRaw);
    __shared__ RunShared sh;
    __shared__ unsigned long long redHi[kRunThreads / 32], redLo[kRunThreads / 32];
    const size_t tb = tileBytes(caps), xb = runScratchBytes(G.n);
    unsigned char *tile = scratch + (size_t)blockIdx.x * (tb + xb);
    TileMem M;
    carveTileMem(M, tile, caps);
    RunScratch X;
    carveRunScratch(X, tile + tb, G.n);
    const Tile<32> T;
    const long long total = P.worklist ? P.nwork : P.nsrc;
    const int maxInter = min(P.maxInter, 16);
    uint32_t sel[16];
    long long accArcs = 0, accSettled = 0, accSources = 0;
    for (uint32_t w = threadIdx.x; w < (uint32_t)((2 * G.n + 31) >> 5) + 1; w += blockDim.x) X.bits[w] = 0u;

    while (true) {
        __syncthreads();
        if (threadIdx.x == 0) {
            const long long wi = (long long)atomicAdd(P.queueCounter, 1ull);
            sh.wi = wi;
            sh.skip = 0;
            if (wi < total) {
                const int32_t srcIdx = P.worklist ? P.worklist[wi] : (int32_t)wi;
                const uint32_t src = P.sources ? (uint32_t)P.sources[srcIdx]
                                               : (uint32_t)(((((long long)(srcIdx >> 1)) * P.srcStride + P.srcBegin) << 1) | (srcIdx & 1));
                sh.srcIdx = srcIdx;
                sh.src = src;
                // `if (blacklist_[currentSource_]) return {}` -- indexed with the DOUBLED id (ocw_seperator.cpp:334, quirk A.7.1)
                if ((long long)src < G.n && (G.blStamp ? stampedBefore(G, src, srcIdx) : P.blacklist[src] != 0)) sh.skip = 1;
                if ((long long)(src >> 1) >= G.n) sh.skip = 1;
            }
        }
        __syncthreads();
        if (sh.wi >= total) break;
        if (sh.skip) continue;
        const uint32_t src = sh.src;
        const int32_t srcIdx = sh.srcIdx;
        const uint32_t target = src ^ 1u;
        // ---- distances_[source] = 0, hopDistances_[source] = 0, heap.push(source)   (:54-57) ----------------------------
        if (threadIdx.x == 0) {
            Entry e = emptyEntry();
            e.key = src >> 1;
            if (src & 1u) {
                e.d1 = 0.0;
                e.h1 = 0;
            } else {
                e.d0 = 0.0;
                e.h0 = 0;
            }
            storeEntry(&M.tab[src >> 1], e);
            M.touched[0] = src >> 1;
            X.qpos[src] = 0;
            X.qpos[src ^ 1u] = kNoQpos;
            HeapEnt k0;
            k0.hi = 0ull;
            k0.lo = (unsigned long long)src;
            M.heap[0] = k0;
            sh.best = P.bestInit;
            sh.bestHop = 0;
            sh.histN = 0;
            sh.poolN = 1;
            sh.touchedN = 1;
            sh.arcs = 0;
            sh.settled = 0;
            sh.overflow = 0;
            sh.done = 0;
        }
        __syncthreads();

        RCLK_DECL;
        while (!sh.done && !sh.overflow) {
            RCLK_RESET;
            // ---- the minimal key of the pool: 128-bit lexicographic minimum over the live keys -------------------------------
            const uint32_t poolN = sh.poolN;
            unsigned long long mhi = ~0ull, mlo = ~0ull;
            for (uint32_t i0 = threadIdx.x; i0 < poolN; i0 += 4 * kRunThreads) { // four keys per thread in flight: the scan is latency
                HeapEnt k[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t i = i0 + j * kRunThreads;
                    k[j].hi = kDeadKey;
                    k[j].lo = ~0ull;
                    if (i < poolN) k[j] = M.heap[i];
                }
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (k[j].hi != kDeadKey && keyLess(k[j].hi, k[j].lo, mhi, mlo)) {
                        mhi = k[j].hi;
                        mlo = k[j].lo;
                    }
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                const unsigned long long ohi = __shfl_xor_sync(0xFFFFFFFFu, mhi, off), olo = __shfl_xor_sync(0xFFFFFFFFu, mlo, off);
                if (keyLess(ohi, olo, mhi, mlo)) {
                    mhi = ohi;
                    mlo = olo;
                }
            }
            if ((threadIdx.x & 31) == 0) {
                redHi[threadIdx.x >> 5] = mhi;
                redLo[threadIdx.x >> 5] = mlo;
            }
            __syncthreads();
            if (threadIdx.x == 0) {
                for (int w = 1; w < kRunThreads / 32; ++w)
                    if (keyLess(redHi[w], redLo[w], mhi, mlo)) {
                        mhi = redHi[w];
                        mlo = redLo[w];
                    }
                sh.minHi = mhi;
                sh.minLo = mlo;
                sh.runLen = 0;
                // loop condition of ocw_seperator.cpp:88-90 for the first member of the run
                const double dmin = __longlong_as_double((long long)mhi);
                if (mhi == ~0ull || (uint32_t)mlo == target || 2 * dmin > sh.best) sh.done = 1;
                sh.runD = dmin;
                sh.runH = (uint32_t)(mlo >> 32);
            }
            __syncthreads();
            if (sh.done) break;
            RCLK(4);
            // ---- collect the run: every live key with the minimal (dist, hops); the keys leave the pool -------------------
            {
                const unsigned long long dhi = sh.minHi;
                const uint32_t hmin = sh.runH;
                for (uint32_t i0 = threadIdx.x; i0 < poolN; i0 += 4 * kRunThreads) {
                    HeapEnt k[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint32_t i = i0 + j * kRunThreads;
                        k[j].hi = kDeadKey;
                        k[j].lo = ~0ull;
                        if (i < poolN) k[j] = M.heap[i];
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint32_t i = i0 + j * kRunThreads;
                        if (i < poolN && k[j].hi == dhi && (uint32_t)(k[j].lo >> 32) == hmin) {
                            const uint32_t pos = atomicAdd(&sh.runLen, 1u);
                            if (pos < X.runCap) X.run[pos] = (uint32_t)k[j].lo;
                            M.heap[i].hi = kDeadKey;
                            X.qpos[(uint32_t)k[j].lo] = kNoQpos;
                        }
                    }
                }
            }
            __syncthreads();
            uint32_t runLen = sh.runLen;
            SMS_CHK(runLen >= 1 && runLen <= X.runCap);
            RCLK(5);
            // ---- in id order; the target ends the run in front of itself (:89) ------------------------------------------------
            const uint32_t nwords = (uint32_t)((2 * G.n + 31) >> 5);
            if (runLen > 1 && (runLen > 4 * kRunOfferCap || nwords <= 16 * runLen)) {
                // a long run: the members are distinct ids below 2n, so "sorting" is marking a bitmap and reading it in order
                for (uint32_t i = threadIdx.x; i < runLen; i += blockDim.x) {
                    const uint32_t id = X.run[i];
                    atomicOr(&X.bits[id >> 5], 1u << (id & 31));
                }
                __syncthreads();
                const uint32_t per = (nwords + blockDim.x - 1) / blockDim.x;
                const uint32_t w0 = min(threadIdx.x * per, nwords), w1 = min(w0 + per, nwords);
                uint32_t cntb = 0;
                for (uint32_t w = w0; w < w1; ++w) cntb += __popc(X.bits[w]);
                uint32_t incl = cntb;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) {
                    const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, off);
                    if ((threadIdx.x & 31) >= off) incl += o;
                }
                __shared__ uint32_t bSum[kRunThreads / 32];
                if ((threadIdx.x & 31) == 31) bSum[threadIdx.x >> 5] = incl;
                __syncthreads();
                uint32_t pos = incl - cntb;
                for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) pos += bSum[w];
                for (uint32_t w = w0; w < w1; ++w) {
                    uint32_t bits = X.bits[w];
                    if (!bits) continue;
                    X.bits[w] = 0u; // left clean for the next run
                    while (bits) {
                        const int b = __ffs(bits) - 1;
                        bits &= bits - 1;
                        X.run[pos++] = 32u * w + (uint32_t)b;
                    }
                }
                __syncthreads();
            } else if (runLen > 1) {
                const uint32_t np = nextPow2(runLen);
                if (np <= 4 * kRunOfferCap) {
                    uint32_t *k32 = reinterpret_cast<uint32_t *>(S.evD); // 2 cap doubles = 4 cap ids of scratch
                    for (uint32_t i = threadIdx.x; i < np; i += blockDim.x) k32[i] = i < runLen ? X.run[i] : 0xFFFFFFFFu;
                    // 32-bit bitonic network in shared memory
                    for (uint32_t size = 2; size <= np; size <<= 1)
                        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
                            __syncthreads();
                            for (uint32_t t = threadIdx.x; t < (np >> 1); t += blockDim.x) {
                                const uint32_t lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                                const bool up = (lo & size) == 0;
                                const uint32_t a = k32[lo], b = k32[hi];
                                if ((a > b) == up) {
                                    k32[lo] = b;
                                    k32[hi] = a;
                                }
                            }
                        }
                    __syncthreads();
                    for (uint32_t i = threadIdx.x; i < runLen; i += blockDim.x) X.run[i] = k32[i];
                } else {
                    for (uint32_t i = runLen + threadIdx.x; i < np; i += blockDim.x) X.run[i] = 0xFFFFFFFFu;
                    bitonicSortGlobal32(X.run, np);
                }
                __syncthreads();
            }
            RCLK(6);
            bool endsAtTarget = false;
            {
                // position of the target in the sorted run, if it is a member (binary search by every thread: uniform)
                uint32_t lo = 0, hi = runLen;
                while (lo < hi) {
                    const uint32_t mid = (lo + hi) >> 1;
                    if (X.run[mid] < target)
                        lo = mid + 1;
                    else
                        hi = mid;
                }
                if (lo < runLen && X.run[lo] == target) {
                    endsAtTarget = true;
                    runLen = lo; // members behind the target are never extracted; the search ends when the target is on top
                }
            }
            // ---- the members, chunk by chunk ----------------------------------------------------------------------------------
            uint32_t m0 = 0;
            while (m0 < runLen && !sh.done && !sh.overflow) {
                // settled bits + prefix of the degrees of the next members, as many as fit a chunk
                const uint32_t window = min(runLen - m0, kRunOfferCap);
                __syncthreads();
                for (uint32_t i = threadIdx.x; i < window; i += blockDim.x) { // ids, rows and degrees first (in parallel), then the prefix
                    const uint32_t id = X.run[m0 + i];
                    const long long rb = G.rowptr[id >> 1];
                    S.mCur[i] = id;
                    S.mRow[i] = (uint32_t)rb;
                    S.offs[i] = (uint32_t)(G.rowptr[(id >> 1) + 1] - rb);
                }
                __syncthreads();
                {
                    // exclusive prefix of the degrees over the window (block scan), and how many members fit a chunk
                    constexpr int IW = kRunOfferCap / kRunThreads;
                    __shared__ uint32_t wSum[kRunThreads / 32];
                    uint32_t dg[IW], tsum = 0;
                    const uint32_t deg0 = S.offs[0]; // (degree of the first member: needed when its row alone exceeds a chunk)
#pragma unroll
                    for (int i = 0; i < IW; ++i) {
                        const uint32_t q = threadIdx.x * IW + i;
                        dg[i] = q < window ? S.offs[q] : 0u;
                        tsum += dg[i];
                    }
                    uint32_t incl = tsum;
#pragma unroll
                    for (int off = 1; off < 32; off <<= 1) {
                        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, off);
                        if ((threadIdx.x & 31) >= off) incl += o;
                    }
                    if ((threadIdx.x & 31) == 31) wSum[threadIdx.x >> 5] = incl;
                    if (threadIdx.x == 0) {
                        sh.nseg = 0;
                        sh.nev = 0;
                    }
                    __syncthreads();
                    uint32_t run0 = incl - tsum;
                    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) run0 += wSum[w];
                    uint32_t fit = 0, last = 0;
#pragma unroll
                    for (int i = 0; i < IW; ++i) {
                        const uint32_t q = threadIdx.x * IW + i;
                        if (q < window) {
                            S.offs[q] = run0;
                            run0 += dg[i];
                            if (run0 <= kRunOfferCap) {
                                fit++;
                                last = run0;
                            }
                        }
                    }
                    if (fit) {
                        atomicAdd(&sh.nseg, fit);  // members whose rows fit completely (a prefix of the window: the sums are monotone)
                        atomicMax(&sh.nev, last);  // arcs of those members
                    }
                    __syncthreads();
                    if (threadIdx.x == 0) {
                        if (sh.nseg == 0) { // the first row alone is longer than a chunk: processed in pieces
                            sh.nseg = 1;
                            sh.nev = deg0;
                        }
                    }
                    __syncthreads();
                    if (threadIdx.x == 0) S.offs[sh.nseg] = sh.nev;
                    __syncthreads();
                    if (sh.nev <= kRunOfferCap) { // rank of the member of every arc of the chunk
                        const uint32_t nmem = sh.nseg;
                        for (uint32_t r = threadIdx.x; r < nmem; r += blockDim.x)
                            for (uint32_t j = S.offs[r]; j < S.offs[r + 1]; ++j) S.aRank[j] = (uint16_t)r;
                    }
                }
                __syncthreads();
                RCLK(7);
                const uint32_t cnt = sh.nseg, tot = sh.nev;
                const bool lastOfRun = (m0 + cnt >= runLen) && !endsAtTarget;
                __syncthreads();
                const uint32_t snapPool = sh.poolN, snapTouched = sh.touchedN;
                const long long snapArcs = sh.arcs;
                if (tot <= kRunOfferCap) {
                    if (tot > 0) runChunk(G, M, X, sh, S, cnt, 0, 0, tot, srcIdx, snapPool, lastOfRun);
                    __syncthreads();
                    int c = tot > 0 ? sh.cutoff : -1;
                    if (c >= 0) {
                        // the search ends behind member m0 + c: redo the chunk up to it (no cutoff possible this time)
                        __syncthreads();
                        if (threadIdx.x == 0) {
                            sh.touchedN = snapTouched;
                            sh.arcs = snapArcs;
                        }
                        __syncthreads();
                        const uint32_t cnt2 = (uint32_t)c + 1;
                        const uint32_t tot2 = S.offs[cnt2];
                        if (tot2 > 0) runChunk(G, M, X, sh, S, cnt2, 0, 0, tot2, srcIdx, sh.poolN, true);
                        __syncthreads();
                        SMS_CHK(sh.cutoff < 0);
                        if (threadIdx.x == 0) {
                            sh.settled += cnt2;
                            sh.done = 1;
                        }
                        for (uint32_t i = threadIdx.x; i < cnt2; i += blockDim.x) {
                            const uint32_t id = X.run[m0 + i];
                            atomicOr(&M.tab[id >> 1].flags, 1u << (id & 1u));
                        }
                        __syncthreads();
                        break;
                    }
                    if (threadIdx.x == 0) sh.settled += cnt;
                    for (uint32_t i = threadIdx.x; i < cnt; i += blockDim.x) {
                        const uint32_t id = X.run[m0 + i];
                        atomicOr(&M.tab[id >> 1].flags, 1u << (id & 1u));
                    }
                } else {
                    // one member with a long row: its arcs in pieces (offers of one member never meet at a vertex)
                    for (uint32_t a0 = 0; a0 < tot && !sh.overflow; a0 += kRunOfferCap) {
                        const uint32_t a1 = min(a0 + kRunOfferCap, tot);
                        __syncthreads();
                        runChunk(G, M, X, sh, S, 1, a0, a1, a1 - a0, srcIdx, sh.poolN, lastOfRun);
                        __syncthreads();
                    }
                    if (threadIdx.x == 0) {
                        sh.settled += 1;
                        const uint32_t id = X.run[m0];
                        M.tab[id >> 1].flags |= 1u << (id & 1u);
                    }
                }
                __syncthreads();
                RCLK_RESET;
                m0 += cnt;
                // the loop condition in front of the next member (:90); within a chunk the replay has tested it already
                if (threadIdx.x == 0 && 2 * sh.runD > sh.best && (m0 < runLen || endsAtTarget)) sh.done = 1;
                __syncthreads();
            }
            if (endsAtTarget) {
                __syncthreads();
                if (threadIdx.x == 0) sh.done = 1;
                __syncthreads();
            }
        }
        __syncthreads();

        RCLK_RESET;
        // ---- everything after the search: the code shared with the other engines, by the first warp --------------------------
        SrcState S;
        S.best = sh.best;
        S.bestHop = (double)sh.bestHop;
        S.histN = sh.histN;
        S.heapN = 0;
        S.touchedN = min(sh.touchedN, M.caps.touchedCap);
        S.arcs = sh.arcs;
        S.settled = sh.settled;
        S.steps = 0;
        S.loads = 0;
        S.overflow = sh.overflow;
        __syncthreads();
        if (threadIdx.x < 32) {
            if (!S.overflow && S.histN > 0) {
                const int nsel = selectIntermediates<32>(T, M, S, maxInter, sel);
                RCLK(10);
                if (!S.overflow) {
                    Emitter E;
                    E.R = &R;
                    E.srcIdx = srcIdx;
                    E.seq = 0;
                    E.tier = P.tier;
#pragma unroll
                    for (int r = 0; r < 16; ++r) sel[r] = T.shfl(sel[r], 0); // lane 0 holds the selection
                    walksOfSource(G, M, S, E, sel, nsel, src, srcIdx);
                    RCLK_RESET;
                }
                S.overflow = T.shfl(S.overflow, 0);
            }
            if (!S.overflow && G.blStampNext) verbatimSourceDone<32>(T, G, src, srcIdx, S.histN == 0);
            if (!S.overflow) {
                if (P.exportLabels) {
                    for (uint32_t i = T.lane; i < S.touchedN; i += 32) {
                        const Entry e = loadEntry(&M.tab[M.touched[i]]);
#pragma unroll
                        for (int sd = 0; sd < 2; ++sd) {
                            const double dd = sd ? e.d1 : e.d0;
                            if (dd >= kUnlabelled) continue;
                            const long long li = (long long)atomicAdd(&R.counters[5], 1ull);
                            if (li < R.labCap) {
                                R.labSrc[li] = srcIdx;
                                R.labNode[li] = (int32_t)(2 * e.key + sd);
                                R.labDist[li] = dd;
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
                    accArcs += S.arcs;
                    accSettled += S.settled;
                    accSources++;
                }
            } else if (T.lane == 0) {
                const unsigned long long k = atomicAdd(P.ovfCount, 1ull);
                P.ovfList[k] = srcIdx;
            }
        }
        __syncthreads();
        RCLK(8);
        // leave the table clean for the next source
        const uint32_t tn = min(sh.touchedN, M.caps.touchedCap);
        for (uint32_t i = threadIdx.x; i < tn; i += blockDim.x) storeEntry(&M.tab[M.touched[i]], emptyEntry());
        RCLK(9);
    }
    if (threadIdx.x == 0 && accSources) {
        atomicAdd(&R.counters[2], (unsigned long long)accArcs);
        atomicAdd(&R.counters[3], (unsigned long long)accSettled);
        atomicAdd(&R.counters[P.runSlot], (unsigned long long)accSources);
    }
}

} // namespace smsgpu
