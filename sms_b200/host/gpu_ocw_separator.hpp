// sms::GpuOcwSeparator -- host-side C++ mirror of sms::OcwSeparator over the C ABI of libsmsgpu.so.
//
// Reference interface mirrored (CharJon/SMS, paths relative to its root):
//   include/sms/graph/ocw_seperator.hpp:14-41   class OcwSeparator: ctor(const NetworKit::Graph&), updateWeights,
//                                               updateEdgeWeight, updateBlacklist, getViolatedOcws
//   include/sms/graph/odd_closed_walk.hpp:21-123 class OddClosedWalk: size, getIthNode, ithEdgeIsStay, numCrosses,
//                                               isValid, isSimple
// Same names, argument meaning and error behaviour (std::runtime_error where the reference throws:
// src/graph/ocw_seperator.cpp:16-19,26-27).  Added: separateRound(), the batch call that replaces the per-vertex
// loop of ConshdlrOscw::sepaOddCycles (src/conshdlr/conshdlr_oscw.cpp:188-195).
//
// Header-only; depends on include/smsgpu.h and the C++17 standard library.  The constructor template accepts any
// graph type with NetworKit::Graph's read interface (upperNodeIdBound, numberOfEdges, degree,
// getIthNeighborWithId, hasEdgeIds, isDirected), so the same header compiles against the real NetworKit inside
// SMS and against the small stand-in used by this repository's headless tests (host/shim/networkit).
// There is no CPU path: every query ends in libsmsgpu.so and throws if no CUDA device is present.
#ifndef SMS_GPU_OCW_SEPARATOR_HPP
#define SMS_GPU_OCW_SEPARATOR_HPP

#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "smsgpu.h"

namespace sms {

// Result type with the read interface of sms::OddClosedWalk (odd_closed_walk.hpp:21-123).  nodes_ holds size()+1
// entries, the last one repeating the first (as OddClosedWalk::addEdge leaves it); edgeIds_ is the extra the GPU
// path returns so that the handler can use ProbDataRootedMc::edgeIdToVar (mc_rooted.hpp:71) instead of the
// adjacency scan behind edgeToVar(u, v).
class GpuOddClosedWalk {
public:
    GpuOddClosedWalk() = default;
    GpuOddClosedWalk(const int32_t *nodes, const int32_t *eids, const uint8_t *stay, int64_t len, double violation, int32_t src = 0)
        : nodes_(nodes, nodes + len), edgeIds_(eids, eids + len), isStay_(stay, stay + len), violation_(violation), src_(src) {
        nodes_.push_back(nodes_.front());
        numCrosses_ = static_cast<unsigned>(std::count(isStay_.begin(), isStay_.end(), uint8_t(0)));
    }
    uint64_t size() const { return isStay_.size(); }                          // odd_closed_walk.hpp:41
    uint64_t getIthNode(uint64_t i) const { return static_cast<uint64_t>(nodes_[i]); } // :43-46, valid for i <= size()
    bool ithEdgeIsStay(uint64_t i) const { return isStay_[i] != 0; }          // :48-51
    bool ithEdgeIsCross(uint64_t i) const { return isStay_[i] == 0; }
    unsigned numCrosses() const { return numCrosses_; }                       // :60
    uint64_t ithEdgeId(uint64_t i) const { return static_cast<uint64_t>(edgeIds_[i]); }
    double violation() const { return violation_; }                           // 1 - length under the round's x*
    int32_t sourceIndex() const { return src_; } // index into the round's source list (dedup: the first finder)
    bool isValid() const { return nodes_.size() >= 3 && nodes_.front() == nodes_.back() && (numCrosses_ % 2 == 1); } // :67-69
    bool isSimple() const {                                                   // :86-96
        std::vector<int32_t> s(nodes_.begin(), nodes_.end() - 1);
        std::sort(s.begin(), s.end());
        return std::adjacent_find(s.begin(), s.end()) == s.end();
    }

private:
    std::vector<int32_t> nodes_, edgeIds_;
    std::vector<uint8_t> isStay_;
    unsigned numCrosses_ = 0;
    double violation_ = 0.;
    int32_t src_ = 0;
};

class GpuOcwSeparator {
public:
    // OcwSeparator::OcwSeparator (src/graph/ocw_seperator.cpp:9-23): rejects graphs that are directed, not compact
    // or without edge ids with std::runtime_error; uploads the CSR once (adjacency in the graph's own order).
    template <class GraphT>
    explicit GpuOcwSeparator(const GraphT &g, int device = 0) {
        if (g.isDirected()) throw std::runtime_error("Graph is directed");
        if (!g.hasEdgeIds()) throw std::runtime_error("Graph has no edge ids.");
        if (g.upperNodeIdBound() != g.numberOfNodes()) throw std::runtime_error("Graph is not compact.");
        const int64_t n = static_cast<int64_t>(g.upperNodeIdBound());
        std::vector<int64_t> rowptr(n + 1, 0);
        for (int64_t u = 0; u < n; ++u) rowptr[u + 1] = rowptr[u] + static_cast<int64_t>(g.degree(u));
        std::vector<int32_t> col(rowptr[n]), eid(rowptr[n]);
        for (int64_t u = 0; u < n; ++u)
            for (uint64_t i = 0; i < g.degree(u); ++i) {
                auto [v, id] = g.getIthNeighborWithId(u, i); // insertion order, Graph.hpp:1666-1671
                col[rowptr[u] + i] = static_cast<int32_t>(v);
                eid[rowptr[u] + i] = static_cast<int32_t>(id);
            }
        init(n, rowptr[n] / 2, rowptr.data(), col.data(), eid.data(), device);
    }
    // same, from CSR arrays in the numbering contract of smsgpu.h
    GpuOcwSeparator(int64_t n, int64_t m, const int64_t *rowptr, const int32_t *col, const int32_t *eid, int device = 0) {
        init(n, m, rowptr, col, eid, device);
    }
    ~GpuOcwSeparator() { smsgpu_destroy(ctx_); }
    GpuOcwSeparator(const GpuOcwSeparator &) = delete;
    GpuOcwSeparator &operator=(const GpuOcwSeparator &) = delete;

    // ocw_seperator.cpp:25-31 -- edgeId(u, v) is the first adjacency entry of u holding v (Graph.cpp:119-127)
    void updateWeights(uint64_t u, uint64_t v, double w) {
        for (int64_t a = rowptr_[u]; a < rowptr_[u + 1]; ++a)
            if (static_cast<uint64_t>(col_[a]) == v) return updateEdgeWeight(static_cast<uint64_t>(eid_[a]), w);
        throw std::runtime_error("Edge does not exist");
    }
    // ocw_seperator.cpp:33-35 (the clamp to [0, 1] itself runs on the device, kernel K1)
    void updateEdgeWeight(uint64_t idx, double ew) {
        weights_.at(idx) = ew;
        dirty_ = true;
    }
    // ocw_seperator.hpp:25-28
    void updateBlacklist(uint64_t u, bool blacklisted) {
        blacklist_.at(u) = blacklisted ? 1 : 0;
        dirty_ = true;
    }
    // whole-round input in one call: x* by edge id (m values) and the blacklist (n bytes)
    void setRoundInput(const double *xByEid, const uint8_t *blacklist) {
        std::copy(xByEid, xByEid + m_, weights_.begin());
        std::copy(blacklist, blacklist + n_, blacklist_.begin());
        dirty_ = true;
    }

    // ocw_seperator.cpp:330-361 for ONE source (a double-cover id, quirk A.7.1), raw list in emission order
    std::vector<GpuOddClosedWalk> getViolatedOcws(uint64_t source) {
        const int32_t s = static_cast<int32_t>(source);
        return separateRound(&s, 1, /*dedup=*/false);
    }
    // The loop of conshdlr_oscw.cpp:188-195 in one call.  sources == nullptr: doubled ids 0..n-1 as the handler
    // issues them; sources whose id is blacklisted are skipped on the device (ocw_seperator.cpp:334).
    // mode: SMSGPU_MODE_INDEPENDENT / SMSGPU_MODE_VERBATIM; minEfficacy: device pre-filter (0 = keep everything)
    std::vector<GpuOddClosedWalk> separateRound(const int32_t *sources = nullptr, int64_t nsrc = 0, bool dedup = true,
                                                int mode = SMSGPU_MODE_INDEPENDENT, double minEfficacy = 0.) {
        flush();
        check(smsgpu_set_sources(ctx_, sources, nsrc, 0, 1));
        smsgpu_params p;
        smsgpu_default_params(&p);
        p.dedup = dedup ? 1 : 0;
        p.mode = mode;
        p.min_efficacy = minEfficacy;
        check(smsgpu_separate(ctx_, &p));
        smsgpu_cuts c;
        check(smsgpu_get_cuts(ctx_, &c));
        std::vector<GpuOddClosedWalk> out;
        out.reserve(static_cast<size_t>(c.ncuts));
        for (int64_t k = 0; k < c.ncuts; ++k)
            out.emplace_back(c.node + c.off[k], c.eid + c.off[k], c.is_stay + c.off[k], c.off[k + 1] - c.off[k], c.violation[k], c.src[k]);
        return out;
    }
    smsgpu_stats stats() const {
        smsgpu_stats s;
        smsgpu_get_stats(ctx_, &s);
        return s;
    }
    // ---- small cycles (K6) on this context: used by GpuSmallCycleSeparator below -------------------------------------
    // SmallChordlessCycles::computeSmallCCs (src/graph/small_ccs.cpp:18-43) on the device; both lists get registered
    std::pair<int64_t, int64_t> enumerateSmallCycles(bool wantHoles) {
        int64_t nt = 0, nh = 0;
        check(smsgpu_small_cycles_enumerate(ctx_, wantHoles ? 1 : 0, &nt, &nh));
        return {nt, nh};
    }
    void setSmallCycles(int k, const std::vector<int32_t> &flat) { check(smsgpu_small_cycles_set(ctx_, k, static_cast<int64_t>(flat.size()) / k, flat.data())); }
    std::vector<GpuOddClosedWalk> scanSmallCycles(int k) {
        flush();
        check(smsgpu_small_cycles_scan(ctx_, k));
        smsgpu_cuts c;
        check(smsgpu_get_small_cycle_cuts(ctx_, k, &c));
        std::vector<GpuOddClosedWalk> out;
        out.reserve(static_cast<size_t>(c.ncuts));
        for (int64_t i = 0; i < c.ncuts; ++i)
            out.emplace_back(c.node + c.off[i], c.eid + c.off[i], c.is_stay + c.off[i], c.off[i + 1] - c.off[i], c.violation[i], c.src[i]);
        return out;
    }
    smsgpu_ctx *context() const { return ctx_; }
    int64_t numberOfNodes() const { return n_; }
    int64_t numberOfEdges() const { return m_; }

private:
    void init(int64_t n, int64_t m, const int64_t *rowptr, const int32_t *col, const int32_t *eid, int device) {
        n_ = n;
        m_ = m;
        rowptr_.assign(rowptr, rowptr + n + 1);
        col_.assign(col, col + 2 * m);
        eid_.assign(eid, eid + 2 * m);
        weights_.assign(static_cast<size_t>(m), 0.);
        blacklist_.assign(static_cast<size_t>(n), 0);
        if (smsgpu_create(&ctx_, n, m, rowptr, col, eid, device) != SMSGPU_OK)
            throw std::runtime_error(std::string("smsgpu_create: ") + smsgpu_last_error(nullptr));
    }
    void flush() {
        if (!dirty_) return;
        check(smsgpu_set_round_input(ctx_, weights_.data(), blacklist_.data()));
        dirty_ = false;
    }
    void check(int rc) const {
        if (rc != SMSGPU_OK) throw std::runtime_error(std::string("libsmsgpu: ") + smsgpu_last_error(ctx_));
    }

    smsgpu_ctx *ctx_ = nullptr;
    int64_t n_ = 0, m_ = 0;
    std::vector<int64_t> rowptr_;
    std::vector<int32_t> col_, eid_;
    std::vector<double> weights_;
    std::vector<uint8_t> blacklist_;
    bool dirty_ = true;
};

// sms::TriangleSeparator (include/sms/conshdlr/triangles.hpp:12-54) for K = 3 and sms::HoleSeparator
// (include/sms/conshdlr/holes.hpp) for K = 4 over the C ABI: same constructor argument, updateWeights / addTriangle /
// addFourHole / getViolatedTriangles / getViolatedFourHoles; the walks come back in list order like the reference's.
// Added: enumerate(), the device version of the SmallChordlessCycles run the handlers do in scip_initsol.
template <int K>
class GpuSmallCycleSeparator {
    static_assert(K == 3 || K == 4, "triangles or four-holes");

public:
    template <class GraphT>
    explicit GpuSmallCycleSeparator(GraphT const *const originalGraph, int device = 0) : gpu_(*originalGraph, device) {}
    void updateWeights(uint64_t u, uint64_t v, double w) { gpu_.updateWeights(u, v, w); } // clamped on the device like triangles.hpp:27
    void updateEdgeWeight(uint64_t edgeId, double w) { gpu_.updateEdgeWeight(edgeId, w); }
    void addTriangle(uint64_t u, uint64_t v, uint64_t w) {
        static_assert(K == 3, "addTriangle is the triangle separator's");
        for (uint64_t a : {u, v, w}) list_.push_back(static_cast<int32_t>(a));
        registered_ = false;
    }
    void addFourHole(uint64_t a, uint64_t b, uint64_t c, uint64_t d) {
        static_assert(K == 4, "addFourHole is the hole separator's");
        for (uint64_t v : {a, b, c, d}) list_.push_back(static_cast<int32_t>(v));
        registered_ = false;
    }
    // scip_initsol: the whole list from the device (replaces the SmallChordlessCycles loop + addTriangle calls)
    int64_t enumerate() {
        const auto cnt = gpu_.enumerateSmallCycles(K == 4);
        list_.clear();
        registered_ = true; // the enumeration registered it
        return K == 3 ? cnt.first : cnt.second;
    }
    std::vector<GpuOddClosedWalk> getViolated() {
        if (!registered_) {
            gpu_.setSmallCycles(K, list_);
            registered_ = true;
        }
        return gpu_.scanSmallCycles(K);
    }
    std::vector<GpuOddClosedWalk> getViolatedTriangles() { return getViolated(); }
    std::vector<GpuOddClosedWalk> getViolatedFourHoles() { return getViolated(); }
    GpuOcwSeparator &context() { return gpu_; }

private:
    GpuOcwSeparator gpu_;
    std::vector<int32_t> list_;
    bool registered_ = false;
};
using GpuTriangleSeparator = GpuSmallCycleSeparator<3>;
using GpuHoleSeparator = GpuSmallCycleSeparator<4>;

} // namespace sms

#endif // SMS_GPU_OCW_SEPARATOR_HPP
