// TEST INFRASTRUCTURE ONLY -- thin C wrapper around the UNMODIFIED reference separator.
//
// This translation unit is OUR code; everything it drives is compiled by oracle/Makefile straight from the
// sources where they lie under /root/reference (src/graph/ocw_seperator.cpp, src/graph/odd_closed_walk.cpp,
// src/graph/small_ccs.cpp + four NetworKit translation units).  Outputs go to oracle/_ref/ only
// (git-ignored, shipped to the GPU box as a prebuilt .so).  No reference source is copied into this repo.
//
// Same C API as ocw_oracle_capi.cpp with the prefix ref_, so tests can diff restatement vs reference and
// bench.py can time the reference itself (cpu_baseline.kind = "reference").
#include <algorithm>
#include <array>
#include <cstdint>
#include <optional>
#include <ranges>
#include <stdexcept>
#include <unordered_set>
#include <vector>

#include "networkit/graph/Graph.hpp"
#include "tlx/container/d_ary_addressable_int_heap.hpp"

#include "sms/graph/odd_closed_walk.hpp"
// The label export (distances of the last Dijkstra run) needs OcwSeparator's private arrays.  Access
// specifiers do not change layout; all standard / third-party headers are already included above.
#define private public
#include "sms/graph/ocw_seperator.hpp"
#undef private
#include "sms/conshdlr/holes.hpp"
#include "sms/conshdlr/triangles.hpp"
#include "sms/graph/small_ccs.hpp"

#include <chrono>
#include "sweep_common.hpp"

namespace {

NetworKit::Graph buildGraph(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev) {
    NetworKit::Graph g(n, /*weighted*/ true, /*directed*/ false);
    for (int64_t e = 0; e < m; ++e) g.addEdge(eu[e], ev[e], 1.0);
    g.indexEdges();
    return g;
}

struct RefEngine {
    NetworKit::Graph g;
    sms::OcwSeparator sep;
    RefEngine(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev) : g(buildGraph(n, m, eu, ev)), sep(g) {}
    void updateEdgeWeight(int64_t e, double w) { sep.updateEdgeWeight(e, w); }
    void resetBlacklist(const uint8_t *bl0) {
        const int64_t n = static_cast<int64_t>(g.numberOfNodes());
        for (int64_t u = 0; u < n; ++u) sep.updateBlacklist(u, bl0[u] != 0); // public API, as the handler does
    }
    std::vector<sms::OddClosedWalk> getViolatedOcws(int64_t s) { return sep.getViolatedOcws(s); }
    void appendLabels(sweep::Result &r) {
        for (size_t v = 0; v < sep.distances_.size(); ++v)
            if (sep.distances_[v] < sep.kInfDist_
                || sep.hopDistances_[v] != static_cast<int>(g.numberOfNodes())) {
                r.labelNode.push_back(static_cast<int32_t>(v));
                r.labelDist.push_back(sep.distances_[v]);
                r.labelHops.push_back(sep.hopDistances_[v]);
            }
    }
    void appendStats(sweep::Result &r) {
        for (int i = 0; i < 6; ++i) r.srcStats.push_back(0);
    }
};

sms::OddClosedWalk walkFromArrays(const int32_t *nodes, const uint8_t *stay, int64_t len) {
    sms::OddClosedWalk w(nodes[0]);
    for (int64_t i = 0; i < len; ++i) w.addEdge(nodes[(i + 1) % len], stay[i] != 0);
    return w;
}
sweep::Result *packWalks(const std::vector<sms::OddClosedWalk> &walks) {
    auto *r = new sweep::Result();
    for (auto &w : walks) {
        for (uint64_t i = 0; i < w.size(); ++i) {
            r->node.push_back(static_cast<int32_t>(w.getIthNode(i)));
            r->stay.push_back(w.ithEdgeIsStay(i) ? 1 : 0);
        }
        r->cutOff.push_back(static_cast<int64_t>(r->node.size()));
        r->cutSrc.push_back(0);
    }
    return r;
}

} // namespace

extern "C" {

// NetworKit's own adjacency order and edge ids, to pin csrFromEdgeList of the restatement and the product's CSR builder
void ref_csr(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, int64_t *rowptr, int32_t *col, int32_t *eid) {
    auto g = buildGraph(n, m, eu, ev);
    int64_t pos = 0;
    for (int64_t u = 0; u < n; ++u) {
        rowptr[u] = pos;
        for (unsigned i = 0; i < g.degree(u); ++i) {
            auto [v, id] = g.getIthNeighborWithId(u, i);
            col[pos] = static_cast<int32_t>(v);
            eid[pos] = static_cast<int32_t>(id);
            ++pos;
        }
    }
    rowptr[n] = pos;
}

void *ref_sweep(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x, const uint8_t *bl0,
                const int32_t *sources, int64_t nsrc, int mode, int wantLabels, int wantCuts) {
    RefEngine eng(n, m, eu, ev);
    return sweep::run(eng, n, x, m, bl0, sources, nsrc, mode, wantLabels, wantCuts);
}
void ref_sizes(void *h, int64_t *out) { sweep::sizes(static_cast<sweep::Result *>(h), out); }
double ref_seconds(void *h) { return static_cast<sweep::Result *>(h)->seconds; }
void ref_timing(void *h, double *out) { sweep::timing(static_cast<sweep::Result *>(h), out); }
void ref_copy_cuts(void *h, int64_t *cutOff, int32_t *cutSrc, int32_t *node, uint8_t *stay) {
    sweep::copyCuts(static_cast<sweep::Result *>(h), cutOff, cutSrc, node, stay);
}
void ref_copy_labels(void *h, int64_t *labelOff, int32_t *node, double *dist, int32_t *hops) {
    sweep::copyLabels(static_cast<sweep::Result *>(h), labelOff, node, dist, hops);
}
void ref_copy_stats(void *h, int64_t *stats) {
    auto *r = static_cast<sweep::Result *>(h);
    std::copy(r->srcStats.begin(), r->srcStats.end(), stats);
}
void ref_free(void *h) { delete static_cast<sweep::Result *>(h); }

void *ref_simple_odd_cycles(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x,
                            const int32_t *nodes, const uint8_t *stay, int64_t len) {
    RefEngine eng(n, m, eu, ev);
    for (int64_t e = 0; e < m; ++e) eng.updateEdgeWeight(e, x[e]);
    return packWalks(eng.sep.simpleOddCycles(walkFromArrays(nodes, stay, len)));
}
void *ref_extract_cycles(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x,
                         const int32_t *nodes, const uint8_t *stay, int64_t len) {
    RefEngine eng(n, m, eu, ev);
    for (int64_t e = 0; e < m; ++e) eng.updateEdgeWeight(e, x[e]);
    return packWalks(eng.sep.extractCycles(walkFromArrays(nodes, stay, len)));
}
void *ref_most_violated(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x, int64_t source) {
    RefEngine eng(n, m, eu, ev);
    for (int64_t e = 0; e < m; ++e) eng.updateEdgeWeight(e, x[e]);
    std::vector<sms::OddClosedWalk> out;
    if (auto w = eng.sep.getMostViolatedOcw(source)) out.push_back(*w);
    return packWalks(out);
}
int ref_odd_selection(const int32_t *nodes, const double *weights, int64_t k, uint8_t *stayOut) {
    std::vector<NetworKit::node> nv(nodes, nodes + k);
    std::vector<double> wv(weights, weights + k);
    auto w = sms::violatedOddSelectionClosedWalk(nv, wv);
    if (!w) return 0;
    for (int64_t i = 0; i < k; ++i) stayOut[i] = w->ithEdgeIsStay(i) ? 1 : 0;
    return 1;
}
void ref_small_ccs(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, int64_t *counts, int32_t *tri,
                   int32_t *holes) {
    auto g = buildGraph(n, m, eu, ev);
    sms::SmallChordlessCycles scc(g);
    scc.run();
    counts[0] = static_cast<int64_t>(scc.triangles.size());
    counts[1] = static_cast<int64_t>(scc.fourHoles.size());
    if (tri)
        for (size_t i = 0; i < scc.triangles.size(); ++i)
            for (int k = 0; k < 3; ++k) tri[3 * i + k] = static_cast<int32_t>(scc.triangles[i][k]);
    if (holes)
        for (size_t i = 0; i < scc.fourHoles.size(); ++i)
            for (int k = 0; k < 4; ++k) holes[4 * i + k] = static_cast<int32_t>(scc.fourHoles[i][k]);
}
void *ref_small_cycle_scan(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x, int K,
                           const int32_t *cycles, int64_t ncycles) {
    auto g = buildGraph(n, m, eu, ev);
    std::vector<sms::OddClosedWalk> out;
    // Both separators construct an EMPTY weighted graph and rely on setWeight() inserting the edge
    // (NetworKit Graph::setWeight adds a missing edge, Graph.cpp:856-877).
    // Timed like one round of the handler (conshdlr_triangles.cpp:86-93): the m updateWeights (set-up, once per round) and
    // getViolated... (the scan); registering the list (scip_initsol) is not part of a round.
    std::vector<double> xe(m);
    for (int64_t e = 0; e < m; ++e) xe[e] = x[g.edgeId(eu[e], ev[e])];
    double tSetup = 0, tScan = 0;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto secs = [](auto a, auto b) { return std::chrono::duration<double>(b - a).count(); };
    if (K == 3) {
        sms::TriangleSeparator ts(&g);
        for (int64_t e = 0; e < m; ++e) ts.updateWeights(eu[e], ev[e], xe[e]); // inserts the edges
        for (int64_t i = 0; i < ncycles; ++i) ts.addTriangle(cycles[3 * i], cycles[3 * i + 1], cycles[3 * i + 2]);
        auto t0 = now();
        for (int64_t e = 0; e < m; ++e) ts.updateWeights(eu[e], ev[e], xe[e]);
        auto t1 = now();
        out = ts.getViolatedTriangles();
        tSetup = secs(t0, t1);
        tScan = secs(t1, now());
    } else {
        sms::HoleSeparator hs(&g);
        for (int64_t e = 0; e < m; ++e) hs.updateWeights(eu[e], ev[e], xe[e]);
        for (int64_t i = 0; i < ncycles; ++i)
            hs.addFourHole(cycles[4 * i], cycles[4 * i + 1], cycles[4 * i + 2], cycles[4 * i + 3]);
        auto t0 = now();
        for (int64_t e = 0; e < m; ++e) hs.updateWeights(eu[e], ev[e], xe[e]);
        auto t1 = now();
        out = hs.getViolatedFourHoles();
        tSetup = secs(t0, t1);
        tScan = secs(t1, now());
    }
    auto *r = packWalks(out);
    r->secondsSetup = tSetup;
    r->secondsSources = tScan;
    r->seconds = tSetup + tScan;
    r->sourcesTimed = ncycles;
    return r;
}

} // extern "C"
