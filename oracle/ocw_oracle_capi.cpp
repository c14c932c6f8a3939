// TEST INFRASTRUCTURE ONLY -- C API (ctypes-friendly) over the CPU restatement in ocw_oracle.hpp.
// Built by oracle/Makefile into oracle/liborc.so.  Never used by the product path.
#include "ocw_oracle.hpp"
#include "sweep_common.hpp"

namespace {

struct OrcEngine {
    orc::Csr g;
    orc::OcwOracle sep;
    std::vector<int64_t> mutated;
    OrcEngine(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, bool canonical = false)
        : g(orc::csrFromEdgeList(n, m, eu, ev)), sep(g, canonical) {}
    void updateEdgeWeight(int64_t e, double w) { sep.updateEdgeWeight(e, w); }
    void resetBlacklist(const uint8_t *bl0) {
        for (int64_t u = 0; u < g.n; ++u) sep.updateBlacklist(u, bl0[u] != 0);
    }
    std::vector<orc::Walk> getViolatedOcws(int64_t s) { return sep.getViolatedOcws(s); }
    void appendLabels(sweep::Result &r) {
        for (auto &l : sep.lastLabels()) {
            r.labelNode.push_back(static_cast<int32_t>(l.node));
            r.labelDist.push_back(l.dist);
            r.labelHops.push_back(l.hops);
        }
    }
    void appendStats(sweep::Result &r) {
        auto &s = sep.lastStats();
        for (int64_t v : {s.arcsScanned, s.settled, s.labelled, s.heapMax, s.intermediates, s.maxWalkLen})
            r.srcStats.push_back(v);
    }
};

// cheaper mode-I reset for big graphs: only undo what the last source changed
struct OrcEngineFast : OrcEngine {
    using OrcEngine::OrcEngine;
    std::vector<uint8_t> base;
    bool primed = false;
    void resetBlacklist(const uint8_t *bl0) {
        if (!primed) {
            OrcEngine::resetBlacklist(bl0);
            base.assign(bl0, bl0 + g.n);
            primed = true;
            return;
        }
        const auto &cur = sep.blacklist();
        // stems and integral components are usually tiny compared with n; a word-wise diff is cheap
        for (int64_t u = 0; u < g.n; ++u)
            if (cur[u] != base[u]) sep.updateBlacklist(u, base[u] != 0);
    }
};

orc::Walk walkFromArrays(const int32_t *nodes, const uint8_t *stay, int64_t len) {
    orc::Walk w(nodes[0]);
    for (int64_t i = 0; i < len; ++i) w.addEdge(nodes[(i + 1) % len], stay[i] != 0);
    return w;
}
sweep::Result *packWalks(const std::vector<orc::Walk> &walks) {
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

void orc_csr(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, int64_t *rowptr, int32_t *col, int32_t *eid) {
    auto g = orc::csrFromEdgeList(n, m, eu, ev);
    std::copy(g.rowptr.begin(), g.rowptr.end(), rowptr);
    std::copy(g.col.begin(), g.col.end(), col);
    std::copy(g.eid.begin(), g.eid.end(), eid);
}

void *orc_sweep(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x, const uint8_t *bl0,
                const int32_t *sources, int64_t nsrc, int mode, int wantLabels, int wantCuts) {
    // mode bit 0: 0 = I (independent sources), 1 = V (verbatim sequential); bit 1: canonical tie-breaking
    OrcEngineFast eng(n, m, eu, ev, (mode & 2) != 0);
    return sweep::run(eng, n, x, m, bl0, sources, nsrc, mode & 1, wantLabels, wantCuts);
}
// The sequential sweep (mode V) computed as the fixed point of independent sweeps over a time-stamped blacklist
// (ocw_oracle.hpp "time-stamped blacklist"): the CPU statement of the device's SMSGPU_MODE_VERBATIM.  info[0] = sweeps
// run, info[1] = sources whose walk list changed after the first sweep (summed over sweeps), info[2] = sources run in
// the last sweep, info[3] = vertices blacklisted at the end.  The result must equal orc_sweep(mode V) -- tests check it.
void *orc_sweep_fixpoint(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x, const uint8_t *bl0,
                         const int32_t *sources, int64_t nsrc, int canonical, int maxSweeps, int64_t *info) {
    OrcEngine eng(n, m, eu, ev, canonical != 0);
    for (int64_t e = 0; e < m; ++e) eng.updateEdgeWeight(e, x[e]);
    eng.sep.stampBegin(bl0);
    sweep::Result *r = nullptr;
    std::vector<uint64_t> prevSig(nsrc, 0), sig(nsrc, 0);
    int64_t sweeps = 0, changed = 0;
    auto t0 = std::chrono::steady_clock::now();
    for (;;) {
        delete r;
        r = new sweep::Result();
        eng.sep.stampSweepBegin(bl0);
        int64_t ran = 0;
        for (int64_t k = 0; k < nsrc; ++k) {
            const int64_t s = sources[k];
            sig[k] = 0;
            const bool skip = (s < n) ? bl0[s] != 0 : false;
            if (!skip) {
                eng.sep.stampNow(k);
                auto walks = eng.getViolatedOcws(s);
                ran += eng.sep.lastStats().settled > 0;
                for (auto &w : walks) {
                    uint64_t h = 1469598103934665603ull;
                    for (uint64_t i = 0; i < w.size(); ++i) {
                        r->node.push_back(static_cast<int32_t>(w.getIthNode(i)));
                        r->stay.push_back(w.ithEdgeIsStay(i) ? 1 : 0);
                        h = (h ^ (uint64_t)(2 * w.getIthNode(i) + w.ithEdgeIsStay(i))) * 1099511628211ull;
                    }
                    sig[k] = sig[k] * 31 + h;
                    r->cutOff.push_back(static_cast<int64_t>(r->node.size()));
                    r->cutSrc.push_back(static_cast<int32_t>(k));
                    r->nWalks += 1;
                    r->totalLen += static_cast<int64_t>(w.size());
                }
                eng.appendStats(*r);
            } else {
                for (int i = 0; i < 6; ++i) r->srcStats.push_back(0);
            }
        }
        ++sweeps;
        if (sweeps > 1)
            for (int64_t k = 0; k < nsrc; ++k) changed += sig[k] != prevSig[k];
        prevSig = sig;
        info[2] = ran;
        if (eng.sep.stampSweepEnd() || sweeps >= maxSweeps) break;
    }
    r->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    info[0] = sweeps;
    info[1] = changed;
    int64_t nbl = 0;
    for (auto v : eng.sep.stamps()) nbl += v != orc::OcwOracle::kNever;
    info[3] = nbl;
    return r;
}
void orc_sizes(void *h, int64_t *out) { sweep::sizes(static_cast<sweep::Result *>(h), out); }
double orc_seconds(void *h) { return static_cast<sweep::Result *>(h)->seconds; }
void orc_timing(void *h, double *out) { sweep::timing(static_cast<sweep::Result *>(h), out); }
void orc_copy_cuts(void *h, int64_t *cutOff, int32_t *cutSrc, int32_t *node, uint8_t *stay) {
    sweep::copyCuts(static_cast<sweep::Result *>(h), cutOff, cutSrc, node, stay);
}
void orc_copy_labels(void *h, int64_t *labelOff, int32_t *node, double *dist, int32_t *hops) {
    sweep::copyLabels(static_cast<sweep::Result *>(h), labelOff, node, dist, hops);
}
void orc_copy_stats(void *h, int64_t *stats) {
    auto *r = static_cast<sweep::Result *>(h);
    std::copy(r->srcStats.begin(), r->srcStats.end(), stats);
}
void orc_free(void *h) { delete static_cast<sweep::Result *>(h); }

// --- single functions, for the reference's known-answer tests --------------------------------------------
// walk given as len nodes (closed implicitly) + len stay flags; x by edge id
void *orc_simple_odd_cycles(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x,
                            const int32_t *nodes, const uint8_t *stay, int64_t len) {
    OrcEngine eng(n, m, eu, ev);
    for (int64_t e = 0; e < m; ++e) eng.updateEdgeWeight(e, x[e]);
    return packWalks(eng.sep.simpleOddCycles(walkFromArrays(nodes, stay, len)));
}
void *orc_extract_cycles(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x,
                         const int32_t *nodes, const uint8_t *stay, int64_t len) {
    OrcEngine eng(n, m, eu, ev);
    for (int64_t e = 0; e < m; ++e) eng.updateEdgeWeight(e, x[e]);
    return packWalks(eng.sep.extractCycles(walkFromArrays(nodes, stay, len)));
}
void *orc_most_violated(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x, int64_t source) {
    OrcEngine eng(n, m, eu, ev);
    for (int64_t e = 0; e < m; ++e) eng.updateEdgeWeight(e, x[e]);
    std::vector<orc::Walk> out;
    if (auto w = eng.sep.getMostViolatedOcw(source)) out.push_back(*w);
    return packWalks(out);
}
// violatedOddSelectionClosedWalk on an explicit cycle; returns 1 and fills stayOut[k] if violated
int orc_odd_selection(const int32_t *nodes, const double *weights, int64_t k, uint8_t *stayOut) {
    std::vector<int64_t> nv(nodes, nodes + k);
    std::vector<double> wv(weights, weights + k);
    auto w = orc::violatedOddSelectionClosedWalk(nv, wv);
    if (!w) return 0;
    for (int64_t i = 0; i < k; ++i) stayOut[i] = w->ithEdgeIsStay(i) ? 1 : 0;
    return 1;
}
// SmallChordlessCycles: counts first (tri==nullptr), then fill
void orc_small_ccs(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, int64_t *counts, int32_t *tri,
                   int32_t *holes) {
    auto g = orc::csrFromEdgeList(n, m, eu, ev);
    auto r = orc::computeSmallCCs(g);
    counts[0] = static_cast<int64_t>(r.triangles.size());
    counts[1] = static_cast<int64_t>(r.fourHoles.size());
    if (tri)
        for (size_t i = 0; i < r.triangles.size(); ++i)
            for (int k = 0; k < 3; ++k) tri[3 * i + k] = static_cast<int32_t>(r.triangles[i][k]);
    if (holes)
        for (size_t i = 0; i < r.fourHoles.size(); ++i)
            for (int k = 0; k < 4; ++k) holes[4 * i + k] = static_cast<int32_t>(r.fourHoles[i][k]);
}
// TriangleSeparator / HoleSeparator scan over an explicit list (K = 3 or 4); x by edge id, clamped inside
void *orc_small_cycle_scan(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev, const double *x, int K,
                           const int32_t *cycles, int64_t ncycles) {
    auto g = orc::csrFromEdgeList(n, m, eu, ev);
    std::vector<double> xc(x, x + m);
    for (auto &v : xc) v = std::min(std::max(0., v), 1.);
    std::vector<orc::Walk> out;
    auto t0 = std::chrono::steady_clock::now();
    if (K == 3) {
        std::vector<std::array<int64_t, 3>> c(ncycles);
        for (int64_t i = 0; i < ncycles; ++i)
            for (int k = 0; k < 3; ++k) c[i][k] = cycles[3 * i + k];
        t0 = std::chrono::steady_clock::now();
        out = orc::violatedSmallCycles<3>(g, c, xc);
    } else {
        std::vector<std::array<int64_t, 4>> c(ncycles);
        for (int64_t i = 0; i < ncycles; ++i)
            for (int k = 0; k < 4; ++k) c[i][k] = cycles[4 * i + k];
        t0 = std::chrono::steady_clock::now();
        out = orc::violatedSmallCycles<4>(g, c, xc);
    }
    auto *r = packWalks(out);
    r->secondsSources = r->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    r->sourcesTimed = ncycles;
    return r;
}

} // extern "C"
