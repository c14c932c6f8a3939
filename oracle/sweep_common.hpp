// TEST INFRASTRUCTURE ONLY -- shared sweep driver + C API plumbing for the two CPU checkers:
//   * oracle/ocw_oracle_capi.cpp : the restatement (ocw_oracle.hpp), symbols prefixed orc_
//   * oracle/ref_driver.cpp      : the UNMODIFIED reference compiled from /root/reference, prefixed ref_
// Both expose the same C functions so tests can diff them.  Never used by the product path.
//
// The sweep reproduces what ConshdlrOscw::sepaOddCycles does without SCIP
// (/root/reference/src/conshdlr/conshdlr_oscw.cpp:174-195): weights by edge id, blacklist, then the
// sources in order, skipping sources the handler would skip (root / locally fixed == initially blacklisted).
//   mode 0 ("I"): the blacklist is re-initialised before every source (sources independent)
//   mode 1 ("V"): verbatim sequential sweep, blacklist side effects carry over
#pragma once
#include <chrono>
#include <cstdint>
#include <cstring>
#include <vector>

namespace sweep {

struct Result {
    // cuts (closed walks): nodes[cut_off[k] .. cut_off[k+1]) are the walk's nodes WITHOUT the repeated end node,
    // stay[i] is the type of the edge leaving nodes[i]
    std::vector<int64_t> cutOff{0};
    std::vector<int32_t> cutSrc; // index into the sources array
    std::vector<int32_t> node;
    std::vector<uint8_t> stay;
    // labels per source (only if requested)
    std::vector<int64_t> labelOff{0};
    std::vector<int32_t> labelNode;
    std::vector<double> labelDist;
    std::vector<int32_t> labelHops;
    // per-source stats: arcs, settled, labelled, heapMax, intermediates, maxWalkLen (zeros for the reference engine)
    std::vector<int64_t> srcStats;
    double seconds = 0;
    // the same time split up (bench.py's CPU baseline): once-per-round set-up (m x updateEdgeWeight + blacklist, the
    // part of conshdlr_oscw.cpp:174-185 that does not scale with the number of sources) and the sources themselves
    double secondsSetup = 0, secondsSources = 0, sumSqSource = 0, maxSource = 0;
    int64_t sourcesTimed = 0;
    int64_t nWalks = 0, totalLen = 0; // counted even when cuts are not stored (timing runs)
};

template <class Engine>
Result *run(Engine &eng, int64_t n, const double *x, int64_t m, const uint8_t *bl0, const int32_t *sources,
            int64_t nsrc, int mode, int wantLabels, int wantCuts) {
    auto *r = new Result();
    auto t0 = std::chrono::steady_clock::now();
    for (int64_t e = 0; e < m; ++e) eng.updateEdgeWeight(e, x[e]);
    eng.resetBlacklist(bl0);
    r->secondsSetup = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (int64_t k = 0; k < nsrc; ++k) {
        const int64_t s = sources[k];
        bool skip = (s < n) ? bl0[s] != 0 : false; // handler-level skip (conshdlr_oscw.cpp:191-193)
        if (!skip) {
            if (mode == 0) eng.resetBlacklist(bl0);
            const auto ts = std::chrono::steady_clock::now();
            auto walks = eng.getViolatedOcws(s);
            const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - ts).count();
            r->secondsSources += dt;
            r->sumSqSource += dt * dt;
            r->maxSource = dt > r->maxSource ? dt : r->maxSource;
            r->sourcesTimed += 1;
            if (wantCuts) {
                for (auto &w : walks) {
                    for (uint64_t i = 0; i < w.size(); ++i) {
                        r->node.push_back(static_cast<int32_t>(w.getIthNode(i)));
                        r->stay.push_back(w.ithEdgeIsStay(i) ? 1 : 0);
                    }
                    r->cutOff.push_back(static_cast<int64_t>(r->node.size()));
                    r->cutSrc.push_back(static_cast<int32_t>(k));
                }
            }
            for (auto &w : walks) {
                r->nWalks += 1;
                r->totalLen += static_cast<int64_t>(w.size());
            }
            if (wantLabels) eng.appendLabels(*r);
            eng.appendStats(*r);
        } else {
            for (int i = 0; i < 6; ++i) r->srcStats.push_back(0);
        }
        if (wantLabels) r->labelOff.push_back(static_cast<int64_t>(r->labelNode.size()));
    }
    r->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return r;
}

inline void sizes(const Result *r, int64_t *out) {
    out[0] = static_cast<int64_t>(r->cutSrc.size());
    out[1] = static_cast<int64_t>(r->node.size());
    out[2] = static_cast<int64_t>(r->labelNode.size());
    out[3] = static_cast<int64_t>(r->srcStats.size() / 6);
    out[4] = r->nWalks;
    out[5] = r->totalLen;
}
inline void timing(const Result *r, double *out) {
    out[0] = r->secondsSetup;
    out[1] = r->secondsSources;
    out[2] = r->sumSqSource;
    out[3] = r->maxSource;
    out[4] = static_cast<double>(r->sourcesTimed);
}
inline void copyCuts(const Result *r, int64_t *cutOff, int32_t *cutSrc, int32_t *node, uint8_t *stay) {
    std::memcpy(cutOff, r->cutOff.data(), r->cutOff.size() * sizeof(int64_t));
    if (!r->cutSrc.empty()) std::memcpy(cutSrc, r->cutSrc.data(), r->cutSrc.size() * sizeof(int32_t));
    if (!r->node.empty()) {
        std::memcpy(node, r->node.data(), r->node.size() * sizeof(int32_t));
        std::memcpy(stay, r->stay.data(), r->stay.size());
    }
}
inline void copyLabels(const Result *r, int64_t *labelOff, int32_t *node, double *dist, int32_t *hops) {
    std::memcpy(labelOff, r->labelOff.data(), r->labelOff.size() * sizeof(int64_t));
    if (!r->labelNode.empty()) {
        std::memcpy(node, r->labelNode.data(), r->labelNode.size() * sizeof(int32_t));
        std::memcpy(dist, r->labelDist.data(), r->labelDist.size() * sizeof(double));
        std::memcpy(hops, r->labelHops.data(), r->labelHops.size() * sizeof(int32_t));
    }
}

} // namespace sweep
