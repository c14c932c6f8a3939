// TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the reference odd-cycle separation path.
//
// This file is the parity oracle for the CUDA path in sms_b200/csrc.  It must never be linked,
// imported or executed by the product path (libsmsgpu.so / sms_b200 package); only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
//
// It restates, function by function, the algorithm of the reference (paths relative to /root/reference):
//   src/graph/ocw_seperator.cpp, include/sms/graph/ocw_seperator.hpp      (sms::OcwSeparator)
//   src/graph/odd_closed_walk.cpp, include/sms/graph/odd_closed_walk.hpp  (sms::OddClosedWalk, violatedOddSelectionClosedWalk)
//   include/sms/conshdlr/triangles.hpp, include/sms/conshdlr/holes.hpp    (TriangleSeparator, HoleSeparator)
//   src/graph/small_ccs.cpp                                               (SmallChordlessCycles)
//   extern/networkit/networkit/cpp/graph/Graph.cpp:131-170,472-520        (adjacency order + edge ids)
//   extern/networkit/extlibs/tlx/tlx/container/d_ary_addressable_int_heap.hpp (binary addressable heap, tie rules)
//
// Parity status: PINNED.  tests/test_oracle_vs_reference.py checks this restatement against
// (a) every known-answer test the reference holds for this path (ocw_separator_gtest.cpp,
// odd_closed_walk_gtest.cpp, triangles_gtest.cpp, holes_gtest.cpp, small_ccs_gtest.cpp) and
// (b) the unmodified reference compiled into oracle/_ref/libsmsref.so, walk-for-walk and
// label-for-label (bitwise), including tie-heavy LP-like x*.
//
// The only intended deviation from the reference: per-source O(n) resets are replaced by
// touched-list resets (same values, same semantics, lets the oracle run on 1M-node graphs).
//
// Two tie modes (constructor flag `canonical`):
//   canonical=false : ties resolved exactly as the reference does (tlx binary heap sift rules, std::sort).
//                     Bitwise equal to oracle/_ref on every input, ties included.
//   canonical=true  : the CANONICAL tie-breaking the CUDA path implements (BASELINE.json north_star:
//                     "identical after canonical tie-breaking"): the priority queue pops by the strict
//                     total order (dist, hops, doubled id); the intermediate list and the chord list are
//                     sorted stably.  Identical to the reference whenever no two queue entries share
//                     (dist, hops) -- e.g. iid U(0,1) x* -- which tests/test_oracle_vs_reference.py checks.
#pragma once
#include <algorithm>
#include <array>
#include <cstdint>
#include <optional>
#include <set>
#include <tuple>
#include <stdexcept>
#include <vector>

namespace orc {

// ---------------------------------------------------------------------------------------------
// Graph in NetworKit numbering.
// ---------------------------------------------------------------------------------------------
struct Csr {
    int64_t n = 0, m = 0;
    std::vector<int64_t> rowptr; // n+1
    std::vector<int32_t> col;    // 2m (self loops not supported), adjacency in addEdge insertion order
    std::vector<int32_t> eid;    // 2m, NetworKit edge id of every adjacency entry

    int64_t degree(int64_t u) const { return rowptr[u + 1] - rowptr[u]; }
    // Graph::indexInOutEdgeArray (Graph.cpp:119-127): first adjacency position holding v.
    int64_t arcOf(int64_t u, int64_t v) const {
        for (int64_t a = rowptr[u]; a < rowptr[u + 1]; ++a)
            if (col[a] == v) return a;
        return -1;
    }
    bool hasEdge(int64_t u, int64_t v) const { return arcOf(u, v) >= 0; }
    int64_t edgeId(int64_t u, int64_t v) const { // Graph.cpp:187-199
        auto a = arcOf(u, v);
        if (a < 0) throw std::runtime_error("Edge does not exist");
        return eid[a];
    }
};

// Graph::addEdge (Graph.cpp:472-520): undirected, append v to adj[u] and u to adj[v], in call order.
// Graph::indexEdges (Graph.cpp:131-170): ids by u ascending, adjacency position ascending, where u >= v;
// the opposite direction copies the id.
inline Csr csrFromEdgeList(int64_t n, int64_t m, const int32_t *eu, const int32_t *ev) {
    Csr g;
    g.n = n;
    g.m = m;
    g.rowptr.assign(n + 1, 0);
    for (int64_t e = 0; e < m; ++e) {
        if (eu[e] == ev[e]) throw std::runtime_error("self loops are not supported");
        if (eu[e] < 0 || ev[e] < 0 || eu[e] >= n || ev[e] >= n) throw std::runtime_error("vertex out of range");
        g.rowptr[eu[e] + 1]++;
        g.rowptr[ev[e] + 1]++;
    }
    for (int64_t u = 0; u < n; ++u) g.rowptr[u + 1] += g.rowptr[u];
    g.col.assign(2 * m, -1);
    g.eid.assign(2 * m, -1);
    std::vector<int64_t> fill(g.rowptr.begin(), g.rowptr.end() - 1);
    std::vector<int64_t> pairArc(2 * m, -1); // position of the opposite entry
    for (int64_t e = 0; e < m; ++e) {
        int64_t a = fill[eu[e]]++, b = fill[ev[e]]++;
        g.col[a] = ev[e];
        g.col[b] = eu[e];
        pairArc[a] = b;
        pairArc[b] = a;
    }
    int64_t omega = 0;
    for (int64_t u = 0; u < n; ++u)
        for (int64_t a = g.rowptr[u]; a < g.rowptr[u + 1]; ++a)
            if (u >= g.col[a]) g.eid[a] = static_cast<int32_t>(omega++);
    // the reference copies ids via indexInOutEdgeArray(v,u) = FIRST matching entry; identical to the
    // paired entry on simple graphs (multi-edges are rejected by the product and not generated in tests).
    for (int64_t a = 0; a < 2 * m; ++a)
        if (g.eid[a] < 0) g.eid[a] = g.eid[pairArc[a]];
    return g;
}

// ---------------------------------------------------------------------------------------------
// OddClosedWalk (odd_closed_walk.hpp:21-123)
// ---------------------------------------------------------------------------------------------
struct Walk {
    std::vector<int64_t> nodes; // nodes_[0..size()], nodes_[size()] == nodes_[0] for a closed walk
    std::vector<uint8_t> stay;  // isStay_[0..size())
    uint64_t numCrosses = 0;

    explicit Walk(int64_t start) { nodes.push_back(start); }
    void addEdge(int64_t next, bool isStay) { // :37-41
        nodes.push_back(next);
        stay.push_back(isStay ? 1 : 0);
        numCrosses += 1 - (isStay ? 1 : 0);
    }
    uint64_t size() const { return nodes.size() - 1; }
    int64_t getIthNode(uint64_t i) const { return nodes[i]; }
    bool ithEdgeIsStay(uint64_t i) const { return stay[i] != 0; }
    bool isValid() const { // :73-81
        uint64_t c = 0;
        for (auto s : stay) c += 1 - s;
        return nodes[0] == nodes[size()] && stay.size() == size() && c == numCrosses && (numCrosses % 2 == 1);
    }
    bool isValid(const Csr &g) const { // :83-90
        bool ok = true;
        for (uint64_t i = 0; i < size(); ++i) ok &= g.hasEdge(nodes[i], nodes[i + 1]);
        return ok && isValid();
    }
    bool isSimple() const { // :58-71
        std::vector<int64_t> s(nodes.begin(), nodes.end() - 1);
        std::sort(s.begin(), s.end());
        return std::adjacent_find(s.begin(), s.end()) == s.end();
    }
    Walk extract(uint64_t start, uint64_t end) const { // :98-106
        Walk w(nodes[start]);
        for (uint64_t i = start + 1; i <= end; ++i) w.addEdge(nodes[i], stay[i - 1]);
        return w;
    }
    Walk splitOnChord(uint64_t start, uint64_t end, bool inner, bool cross) const { // odd_closed_walk.cpp:67-90
        if (inner) {
            Walk w(nodes[start]);
            for (uint64_t i = start + 1; i <= end; ++i) w.addEdge(nodes[i], stay[i - 1]);
            w.addEdge(nodes[start], !cross);
            return w;
        }
        Walk w(nodes[0]);
        for (uint64_t i = 1; i <= start; ++i) w.addEdge(nodes[i], stay[i - 1]);
        w.addEdge(nodes[end], !cross);
        for (uint64_t i = end + 1; i < nodes.size(); ++i) w.addEdge(nodes[i], stay[i - 1]);
        return w;
    }
};

// violatedOddSelectionClosedWalk (odd_closed_walk.cpp:8-65): closed-form most violated odd subset on a fixed cycle.
inline std::optional<Walk> violatedOddSelectionClosedWalk(const std::vector<int64_t> &nodes,
                                                          const std::vector<double> &weights) {
    const uint64_t kUnassigned = UINT64_MAX;
    double curLength = 0;
    std::vector<bool> crossEdges(nodes.size(), false);
    double lightestCross = 1;
    uint64_t lightestCrossIndex = kUnassigned;
    double heaviestStay = 0;
    uint64_t heaviestStayIndex = kUnassigned;
    int numCrossSelected = 0;
    for (uint64_t i = 0; i < nodes.size(); i++) {
        if (weights[i] > 0.5) {
            curLength += 1 - weights[i];
            crossEdges[i] = true;
            numCrossSelected += 1;
            if (weights[i] < lightestCross) {
                lightestCross = weights[i];
                lightestCrossIndex = i;
            }
        } else {
            curLength += weights[i];
            crossEdges[i] = false;
            if (weights[i] > heaviestStay) {
                heaviestStay = weights[i];
                heaviestStayIndex = i;
            }
        }
        if (curLength >= 1) return {};
    }
    if (numCrossSelected % 2 == 0) {
        if ((lightestCrossIndex != kUnassigned) && (curLength + 2 * lightestCross - 1 < 1)) {
            crossEdges[lightestCrossIndex] = false;
        } else if ((heaviestStayIndex != kUnassigned) && (curLength - 2 * heaviestStay + 1 < 1)) {
            crossEdges[heaviestStayIndex] = true;
        } else {
            return {};
        }
    }
    Walk w(nodes[0]);
    for (unsigned i = 1; i < nodes.size(); ++i) w.addEdge(nodes[i], !crossEdges[i - 1]);
    w.addEdge(nodes[0], !crossEdges.back());
    return w;
}

// ---------------------------------------------------------------------------------------------
// Binary addressable heap with the sift rules of tlx::DAryAddressableIntHeap<.,2,.>
// (d_ary_addressable_int_heap.hpp:121-134 push, :153-169 remove, :185-190 extract_top, :206-215 update,
//  :305-317 sift_up, :321-350 sift_down).  Ties: sift_up moves past parents that are NOT strictly
// smaller; sift_down moves only below strictly smaller children; left child wins ties.
// ---------------------------------------------------------------------------------------------
template <class Less>
class TlxBinaryHeap {
public:
    explicit TlxBinaryHeap(Less less, std::vector<int64_t> &handles) : less_(less), handles_(handles) {}
    bool empty() const { return heap_.empty(); }
    size_t size() const { return heap_.size(); }
    int64_t top() const { return heap_[0]; }
    bool contains(int64_t key) const { return handles_[key] >= 0; }
    void push(int64_t key) {
        handles_[key] = static_cast<int64_t>(heap_.size());
        heap_.push_back(key);
        siftUp(heap_.size() - 1);
    }
    void update(int64_t key) {
        if (handles_[key] < 0)
            push(key);
        else if (handles_[key] && less_(heap_[handles_[key]], heap_[parent(handles_[key])]))
            siftUp(handles_[key]);
        else
            siftDown(handles_[key]);
    }
    int64_t extractTop() {
        int64_t topItem = heap_[0];
        remove(topItem);
        return topItem;
    }
    void remove(int64_t key) {
        size_t h = handles_[key];
        std::swap(heap_[h], heap_.back());
        handles_[heap_[h]] = h;
        handles_[heap_.back()] = -1;
        heap_.pop_back();
        if (h < heap_.size()) {
            if (h && less_(heap_[h], heap_[parent(h)]))
                siftUp(h);
            else
                siftDown(h);
        }
    }
    // leaves handles_ clean for the next run
    void clear() {
        for (auto k : heap_) handles_[k] = -1;
        heap_.clear();
    }
    size_t maxSize = 0;

private:
    static size_t parent(size_t k) { return (k - 1) / 2; }
    void siftUp(size_t k) {
        maxSize = std::max(maxSize, heap_.size());
        int64_t value = heap_[k];
        size_t p = parent(k);
        while (k > 0 && !less_(heap_[p], value)) {
            heap_[k] = heap_[p];
            handles_[heap_[k]] = k;
            k = p, p = parent(k);
        }
        handles_[value] = k;
        heap_[k] = value;
    }
    void siftDown(size_t k) {
        int64_t value = heap_[k];
        while (true) {
            size_t l = 2 * k + 1;
            if (l >= heap_.size()) break;
            size_t c = l;
            size_t right = std::min(heap_.size(), c + 2);
            while (++l < right)
                if (less_(heap_[l], heap_[c])) c = l;
            if (!less_(heap_[c], value)) break;
            heap_[k] = heap_[c];
            handles_[heap_[k]] = k;
            k = c;
        }
        handles_[value] = k;
        heap_[k] = value;
    }
    Less less_;
    std::vector<int64_t> heap_;
    std::vector<int64_t> &handles_;
};

// ---------------------------------------------------------------------------------------------
// OcwSeparator (ocw_seperator.hpp:14-85, ocw_seperator.cpp)
// ---------------------------------------------------------------------------------------------
struct SourceStats {
    int64_t arcsScanned = 0;  // numberOfEdgesSeen (ocw_seperator.cpp:86,100)
    int64_t settled = 0;      // heap extractions
    int64_t labelled = 0;     // distinct double-cover nodes that ever received a label
    int64_t heapMax = 0;      // max heap size
    int64_t intermediates = 0;
    int64_t maxWalkLen = 0;
    double finalBestDist = 0;
};

class OcwOracle {
public:
    explicit OcwOracle(const Csr &g, bool canonical = false)
        : n_(g.n), g_(g), canonical_(canonical), weights_(g.m, 0.), currentSource_(-1), blacklist_(g.n, 0) { // :9-23
        distances_.assign(2 * n_ + 1, kInfDist_);
        hopDistances_.assign(2 * n_ + 1, static_cast<int>(2 * n_ + 2));
        previous_.assign(2 * n_ + 1, -1);
        handles_.assign(2 * n_ + 1, -1);
        // first runDijsktra resets hop distances to n_ (quirk A.7.5); emulate the reset values lazily
        std::fill(hopDistances_.begin(), hopDistances_.end(), static_cast<int>(n_));
    }

    void updateWeights(int64_t u, int64_t v, double w) { // :25-31
        weights_[g_.edgeId(u, v)] = std::min(std::max(0., w), 1.);
    }
    void updateEdgeWeight(int64_t idx, double ew) { weights_[idx] = std::min(std::max(0., ew), 1.); } // :33-35
    void updateBlacklist(int64_t u, bool b) { blacklist_[u] = b; }                                      // hpp:25-28
    bool blacklisted(int64_t u) const { return blacklist_[u]; }
    const std::vector<uint8_t> &blacklist() const { return blacklist_; }

    std::vector<Walk> getViolatedOcws(int64_t source) { // :330-361
        currentSource_ = source;
        last_ = SourceStats{};
        std::vector<Walk> violated;
        if (isBl(currentSource_)) return violated; // indexed with the DOUBLED id (quirk A.7.1)
        runDijkstra();
        if (intermediateVertices_.empty()) return violated;
        std::vector<Walk> startOcws;
        for (int i = 0; i < static_cast<int>(intermediateVertices_.size()) && i < 10; ++i)
            startOcws.emplace_back(intermediateVertexToOcw(intermediateVertices_[i]));
        for (const auto &start : startOcws) {
            last_.maxWalkLen = std::max<int64_t>(last_.maxWalkLen, start.size());
            for (const auto &simple : simpleOddCycles(start))
                for (auto &c : extractCycles(simple)) violated.push_back(c);
        }
        return violated;
    }

    std::optional<Walk> getMostViolatedOcw(int64_t source) { // :190-200
        currentSource_ = source;
        last_ = SourceStats{};
        if (blacklist_[currentSource_]) return {};
        runDijkstra();
        if (intermediateVertices_.empty()) return {};
        return intermediateVertexToOcw(intermediateVertices_.back());
    }

    std::vector<Walk> simpleOddCycles(const Walk &ocw) const { // :202-227
        std::vector<Walk> out;
        std::vector<std::pair<int64_t, int64_t>> prevPos; // sparse stand-in for vector<int64>(n_, -1)
        auto get = [&](int64_t v) -> int64_t {
            for (auto &p : prevPos)
                if (p.first == v) return p.second;
            return -1;
        };
        auto set = [&](int64_t v, int64_t pos) {
            for (auto &p : prevPos)
                if (p.first == v) {
                    p.second = pos;
                    return;
                }
            prevPos.emplace_back(v, pos);
        };
        int64_t lastLeftPos = -1;
        for (uint64_t i = 0; i < ocw.size() + 1; i++) {
            const auto v = ocw.getIthNode(i);
            if (lastLeftPos < get(v)) {
                lastLeftPos = get(v);
                if (i - lastLeftPos > 2) out.push_back(ocw.extract(lastLeftPos, i));
            }
            set(v, static_cast<int64_t>(i));
        }
        if (out.empty()) out.push_back(ocw);
        return out;
    }

    std::vector<Walk> extractCycles(const Walk &ocw) const { // :229-328
        std::vector<Walk> result;
        const uint64_t cycleLength = ocw.size();
        std::vector<int> numCrossingUpTo(cycleLength + 1, 0);
        std::vector<double> pathLengthUpTo(cycleLength + 1, 0.);
        int numCrossings = 0;
        double pathLength = 0.;
        for (unsigned i = 1; i <= cycleLength; i++) {
            auto cur = ocw.getIthNode(i), prev = ocw.getIthNode(i - 1);
            numCrossings += (ocw.ithEdgeIsStay(i - 1) ? 0 : 1);
            auto e = g_.edgeId(cur, prev);
            pathLength += ocw.ithEdgeIsStay(i - 1) ? weights_[e] : 1 - weights_[e];
            numCrossingUpTo[i] = numCrossings;
            pathLengthUpTo[i] = pathLength;
        }
        struct Chord {
            uint64_t start, end, length;
            bool crossing, inner;
        };
        std::vector<Chord> goodChords;
        for (uint64_t i = 2; i < cycleLength; i++) {
            auto cur = ocw.getIthNode(i);
            for (int64_t a = g_.rowptr[cur]; a < g_.rowptr[cur + 1]; ++a) {
                int64_t nb = g_.col[a];
                if (nb == ocw.getIthNode(i - 1) || nb == ocw.getIthNode(i + 1)) continue;
                uint64_t j = UINT64_MAX;
                for (uint64_t k = 0; k < i; k++)
                    if (ocw.getIthNode(k) == nb) {
                        j = k;
                        break;
                    }
                if (j == UINT64_MAX) continue;
                uint64_t innerCrossingDif = numCrossingUpTo[i] - numCrossingUpTo[j];
                double innerLen = pathLengthUpTo[i] - pathLengthUpTo[j];
                double outerLen = pathLengthUpTo[cycleLength] - innerLen;
                double x = weights_[g_.edgeId(ocw.getIthNode(i), ocw.getIthNode(j))];
                if (innerCrossingDif % 2 == 0) {
                    if (innerLen - x < 0) goodChords.push_back({j, i, i - j + 1, true, true});
                    if (outerLen + x < 1) goodChords.push_back({j, i, cycleLength - i + j + 1, false, false});
                } else {
                    if (innerLen + x < 1) goodChords.push_back({j, i, i - j + 1, false, true});
                    if (outerLen - x < 0) goodChords.push_back({j, i, cycleLength - i + j + 1, true, false});
                }
            }
        }
        if (canonical_)
            std::stable_sort(goodChords.begin(), goodChords.end(), [](auto &a, auto &b) { return a.length < b.length; });
        else
            std::sort(goodChords.begin(), goodChords.end(), [](auto &a, auto &b) { return a.length < b.length; });
        // coveredNodes: vector<bool>(n_) indexed by POSITION when read and by NODE ID when written (quirk A.7.3)
        std::vector<int64_t> covered;
        auto isCovered = [&](int64_t k) { return std::find(covered.begin(), covered.end(), k) != covered.end(); };
        for (auto c : goodChords) {
            if (!isCovered(c.start) && !isCovered(c.end)) {
                auto w = ocw.splitOnChord(c.start, c.end, c.inner, c.crossing);
                result.push_back(w);
                for (uint64_t i = 1; i < w.size() - 1; i++) covered.push_back(w.getIthNode(i));
            }
        }
        if (result.empty()) return {ocw};
        return result;
    }

    // labels of the last run (for the 1e-9 distance parity check): all doubled nodes with dist < kInf
    struct Label {
        int64_t node;
        double dist;
        int hops;
    };
    std::vector<Label> lastLabels() const {
        std::vector<Label> out;
        for (auto v : touched_) out.push_back({v, distances_[v], hopDistances_[v]});
        std::sort(out.begin(), out.end(), [](auto &a, auto &b) { return a.node < b.node; });
        return out;
    }
    const SourceStats &lastStats() const { return last_; }
    double weight(int64_t e) const { return weights_[e]; }

    static constexpr double kInfDist_ = 1.0 + 10e-4; // hpp:45

private:
    static int64_t twinId(int64_t u) { return u ^ 1; }
    static int64_t toSide(int64_t u, bool side) { return 2 * u + side; }
    static int64_t originalId(int64_t u) { return u >> 1; }
    static bool side(int64_t u) { return u & 1; }
    static bool sameSide(int64_t u, int64_t v) { return !(side(u) ^ side(v)); }

    void runDijkstra() { // :37-118
        const int64_t target = twinId(currentSource_);
        // reset (touched-list stand-in for the two std::fill calls at :41-42)
        for (auto v : touched_) {
            distances_[v] = kInfDist_;
            hopDistances_[v] = static_cast<int>(n_);
        }
        touched_.clear();
        intermediateVertices_.clear();

        auto distSmaller = [this](int64_t u, int64_t v) {
            return distances_[u] == distances_[v] ? hopDistances_[u] < hopDistances_[v]
                                                  : distances_[u] < distances_[v];
        };
        TlxBinaryHeap<decltype(distSmaller)> heap(distSmaller, handles_);
        // canonical queue: strict total order (dist, hops, doubled id); handles_[v] >= 0 marks membership
        using CKey = std::tuple<double, int, int64_t>;
        std::set<CKey> cq;
        size_t cqMax = 0;
        auto qEmpty = [&] { return canonical_ ? cq.empty() : heap.empty(); };
        auto qTop = [&]() -> int64_t { return canonical_ ? std::get<2>(*cq.begin()) : heap.top(); };
        auto qPop = [&]() -> int64_t {
            if (!canonical_) return heap.extractTop();
            int64_t v = std::get<2>(*cq.begin());
            cq.erase(cq.begin());
            handles_[v] = -1;
            return v;
        };
        auto qPush = [&](int64_t v) {
            if (!canonical_) return heap.push(v);
            cq.emplace(distances_[v], hopDistances_[v], v);
            handles_[v] = 0;
            cqMax = std::max(cqMax, cq.size());
        };

        double bestDist = 1.0 - 1e-6;
        double bestHopDist = 0;

        auto touch = [this](int64_t v) {
            if (distances_[v] == kInfDist_ && hopDistances_[v] == static_cast<int>(n_)) touched_.push_back(v);
        };
        touch(currentSource_);
        distances_[currentSource_] = 0.;
        hopDistances_[currentSource_] = 0;
        qPush(currentSource_);

        auto updatePaths = [&](int64_t t, double newDist, double newHopDist, int64_t pred) { // :59-84
            if (distances_[t] == newDist ? newHopDist < hopDistances_[t] : newDist < distances_[t]) {
                touch(t);
                if (canonical_ && handles_[t] >= 0) cq.erase(CKey(distances_[t], hopDistances_[t], t));
                distances_[t] = newDist;
                hopDistances_[t] = static_cast<int>(newHopDist);
                previous_[t] = pred;
                if (canonical_)
                    qPush(t);
                else if (heap.contains(t))
                    heap.update(t);
                else
                    heap.push(t);
                auto tw = twinId(t);
                auto D = distances_[t] + distances_[tw];
                auto H = hopDistances_[t] + hopDistances_[tw];
                if (D < bestDist || ((D == bestDist) && (H < bestHopDist))) {
                    bestDist = D;
                    bestHopDist = H;
                    intermediateVertices_.push_back(t);
                } else if ((D == bestDist) && (H < 2 * bestHopDist)) {
                    intermediateVertices_.push_back(t);
                }
            }
        };

        while (!qEmpty() && qTop() != target && (2 * distances_[qTop()] <= bestDist)) { // :88-90
            int64_t current = qPop();
            int64_t cur = originalId(current);
            last_.settled++;
            for (int64_t a = g_.rowptr[cur]; a < g_.rowptr[cur + 1]; ++a) { // :96-111
                int64_t nb = g_.col[a];
                if (isBl(nb)) continue;
                last_.arcsScanned++;
                const double newHopDist = hopDistances_[current] + 1;
                const double stayW = weights_[g_.eid[a]];
                updatePaths(toSide(nb, side(current)), distances_[current] + stayW, newHopDist, current);
                const double crossW = 1.0 - stayW;
                updatePaths(toSide(nb, !side(current)), distances_[current] + crossW, newHopDist, current);
            }
        }
        for (auto &k : cq) handles_[std::get<2>(k)] = -1;
        last_.finalBestDist = bestDist;
        last_.heapMax = static_cast<int64_t>(canonical_ ? cqMax : heap.maxSize);
        heap.clear();
        last_.labelled = static_cast<int64_t>(touched_.size());
        last_.intermediates = static_cast<int64_t>(intermediateVertices_.size());
        if (intermediateVertices_.empty()) blacklistIntReachable(originalId(currentSource_)); // :113-114
        auto byFinalLabel = [this](auto u, auto v) { // :115-117
            return distances_[u] == distances_[v] ? hopDistances_[u] < hopDistances_[v]
                                                  : distances_[u] < distances_[v];
        };
        if (canonical_)
            std::stable_sort(intermediateVertices_.begin(), intermediateVertices_.end(), byFinalLabel);
        else
            std::sort(intermediateVertices_.begin(), intermediateVertices_.end(), byFinalLabel);
    }

    struct ParityEdge {
        int64_t u, v;
        bool isStay;
    };
    std::vector<ParityEdge> getShortestPath(int64_t v) { // :120-136
        std::vector<ParityEdge> path;
        auto cur = v;
        while (cur != currentSource_) {
            auto prev = previous_[cur];
            path.push_back({originalId(cur), originalId(prev), sameSide(cur, prev)});
            cur = prev;
        }
        return path;
    }

    Walk intermediateVertexToOcw(int64_t inter) { // :138-188
        auto path = getShortestPath(inter);
        auto twinPath = getShortestPath(twinId(inter));
        if (twinPath.empty()) {
            Walk w(originalId(inter));
            for (auto &e : path) w.addEdge(e.v, e.isStay);
            return w;
        }
        size_t endPath = path.size() - 1, endTwin = twinPath.size() - 1;
        while ((path[endPath].v == twinPath[endTwin].v) && (endPath > 0) && (endTwin > 0)) { // :164-169
            endPath--;
            endTwin--;
            setBl(path[endPath].v); // stem blacklisting side effect
        }
        if (path[endPath].v != twinPath[endTwin].v) {
            endPath++;
            endTwin++;
        }
        Walk w(originalId(inter));
        for (size_t i = 0; i <= endPath; i++) w.addEdge(path[i].v, path[i].isStay);
        for (int64_t i = static_cast<int64_t>(endTwin); i >= 0; i--) w.addEdge(twinPath[i].u, twinPath[i].isStay);
        return w;
    }

    void blacklistIntReachable(int64_t source) { // :363-382
        setBl(source);
        std::vector<int64_t> stack{source};
        while (!stack.empty()) {
            auto next = stack.back();
            stack.pop_back();
            for (int64_t a = g_.rowptr[next]; a < g_.rowptr[next + 1]; ++a) {
                auto w = weights_[g_.eid[a]];
                // stamp mode: a vertex stamped <= now in THIS sweep is blacklisted as well (by an earlier source, or by
                // this very flood: the reference's DFS reads its own writes)
                if ((w == 0.0 || w == 1.0) && !isBl(g_.col[a]) && !(stampMode_ && stampNext_[g_.col[a]] <= now_)) {
                    setBl(g_.col[a]);
                    stack.push_back(g_.col[a]);
                }
            }
        }
    }

    // ---- time-stamped blacklist (the device's SMSGPU_MODE_VERBATIM, restated on the CPU) -------------------------------
    // The sequential sweep's blacklist is a function of time: vertex v is blacklisted for the source at position t of
    // the sweep iff some source at a position < t blacklisted it (or it was blacklisted initially).  With
    // stamp[v] = position of the FIRST source that blacklists v (-1: initially, kNever: never) the sweep becomes a
    // fixed-point problem: every source t reads "stamp[v] < t" and writes min(stamp[v], t).  Sources are then independent
    // GIVEN the stamps, and iterating whole sweeps (each reading the stamps of the previous sweep) converges to the
    // sequential result: after k sweeps the first k sources are exact, in practice after a handful.
    bool isBl(int64_t v) const { return stampMode_ ? stampCur_[v] < now_ : blacklist_[v] != 0; }
    void setBl(int64_t v) {
        if (stampMode_) {
            if (now_ < stampNext_[v]) stampNext_[v] = now_;
        } else {
            blacklist_[v] = true;
        }
    }

public:
    static constexpr int64_t kNever = INT64_MAX;
    void stampBegin(const uint8_t *bl0) { // before the first sweep
        stampMode_ = true;
        stampCur_.assign(n_, kNever);
        for (int64_t v = 0; v < n_; ++v)
            if (bl0[v]) stampCur_[v] = -1;
        stampNext_ = stampCur_;
    }
    void stampSweepBegin(const uint8_t *bl0) { // every sweep writes a fresh stamp array
        for (int64_t v = 0; v < n_; ++v) stampNext_[v] = bl0[v] ? -1 : kNever;
    }
    bool stampSweepEnd() { // true: the sweep reproduced the stamps it read -> fixed point
        const bool same = stampNext_ == stampCur_;
        stampCur_ = stampNext_;
        return same;
    }
    void stampNow(int64_t t) { now_ = t; }
    const std::vector<int64_t> &stamps() const { return stampCur_; }

private:
    bool stampMode_ = false;
    int64_t now_ = 0;
    std::vector<int64_t> stampCur_, stampNext_;

    const int64_t n_;
    const Csr &g_;
    const bool canonical_;
    std::vector<double> weights_;
    std::vector<double> distances_;
    std::vector<int> hopDistances_;
    std::vector<int64_t> previous_;
    int64_t currentSource_;
    std::vector<uint8_t> blacklist_;
    std::vector<int64_t> intermediateVertices_;
    std::vector<int64_t> handles_;
    std::vector<int64_t> touched_;
    SourceStats last_;
};

// ---------------------------------------------------------------------------------------------
// SmallChordlessCycles::computeSmallCCs (small_ccs.cpp:18-43)
// ---------------------------------------------------------------------------------------------
struct SmallCcs {
    std::vector<std::array<int64_t, 3>> triangles;
    std::vector<std::array<int64_t, 4>> fourHoles;
};
inline SmallCcs computeSmallCCs(const Csr &g) {
    SmallCcs r;
    for (int64_t u = 0; u < g.n; ++u)
        for (int64_t a = g.rowptr[u]; a < g.rowptr[u + 1]; ++a) {
            int64_t x = g.col[a];
            if (x <= u) continue;
            for (int64_t b = g.rowptr[u]; b < g.rowptr[u + 1]; ++b) {
                int64_t y = g.col[b];
                if (y <= x) continue;
                if (g.hasEdge(x, y)) {
                    r.triangles.push_back({x, u, y});
                } else {
                    for (int64_t c = g.rowptr[x]; c < g.rowptr[x + 1]; ++c) {
                        int64_t z = g.col[c];
                        if (z <= u || g.hasEdge(u, z)) continue;
                        if (g.hasEdge(z, y)) r.fourHoles.push_back({x, u, y, z});
                    }
                }
            }
        }
    return r;
}

// TriangleSeparator::getViolatedTriangles (triangles.hpp:34-53) / HoleSeparator::getViolatedFourHoles (holes.hpp:38-54).
// x is indexed by edge id and must already be clamped to [0,1] (updateWeights clamps, triangles.hpp:22-30).
template <size_t K>
inline std::vector<Walk> violatedSmallCycles(const Csr &g, const std::vector<std::array<int64_t, K>> &cycles,
                                             const std::vector<double> &x) {
    std::vector<Walk> out;
    for (auto &c : cycles) {
        std::vector<int64_t> nodes(c.begin(), c.end());
        std::vector<double> vals(K);
        for (size_t i = 0; i + 1 < K; ++i) vals[i] = x[g.edgeId(c[i], c[i + 1])];
        vals[K - 1] = x[g.edgeId(c[K - 1], c[0])]; // weight(w,u) for triangles, weight(a,d) for holes (same edge)
        auto w = violatedOddSelectionClosedWalk(nodes, vals);
        if (w) out.push_back(*w);
    }
    return out;
}

} // namespace orc
