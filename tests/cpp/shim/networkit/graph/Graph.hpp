// HEADLESS-TEST STAND-IN for <networkit/graph/Graph.hpp>: the handful of NetworKit::Graph members the
// cycleparity handler and GpuOcwSeparator read.  Inside SMS the real NetworKit header is used instead (this
// directory is simply not put on the include path).  Semantics follow NetworKit's documented behaviour:
//   * addEdge appends v to u's adjacency and u to v's (insertion order; Graph.cpp:472-520)
//   * indexEdges numbers the edges by ascending u, ascending adjacency position, taken where u >= v (Graph.cpp:131-170)
//   * edgeId(u, v) is the id stored at the first adjacency position of u holding v (Graph.cpp:119-127, 187-199)
#pragma once
#include <cstdint>
#include <stdexcept>
#include <utility>
#include <vector>

namespace NetworKit {

using node = uint64_t;
using index = uint64_t;
using edgeid = uint64_t;
using edgeweight = double;
using count = uint64_t;

struct Edge {
    node u, v;
};

class Graph {
public:
    explicit Graph(count n = 0, bool weighted = false, bool directed = false) : adj_(n), ids_(n), weighted_(weighted), directed_(directed) {}
    count numberOfNodes() const { return adj_.size(); }
    count upperNodeIdBound() const { return adj_.size(); }
    count numberOfEdges() const { return m_; }
    count upperEdgeIdBound() const { return indexed_ ? m_ : 0; }
    bool isDirected() const { return directed_; }
    bool isWeighted() const { return weighted_; }
    bool hasEdgeIds() const { return indexed_; }
    bool hasNode(node u) const { return u < adj_.size(); }
    count degree(node u) const { return adj_[u].size(); }
    void addEdge(node u, node v, edgeweight = 1.0) {
        adj_[u].push_back(v);
        ids_[u].push_back(0);
        adj_[v].push_back(u);
        ids_[v].push_back(0);
        ++m_;
        indexed_ = false;
    }
    void indexEdges(bool = false) {
        edgeid next = 0;
        for (node u = 0; u < adj_.size(); ++u)
            for (index i = 0; i < adj_[u].size(); ++i) {
                const node v = adj_[u][i];
                if (v <= u) {
                    ids_[u][i] = next;
                    for (index j = 0; j < adj_[v].size(); ++j)
                        if (adj_[v][j] == u) {
                            ids_[v][j] = next;
                            break;
                        }
                    ++next;
                }
            }
        indexed_ = true;
    }
    bool hasEdge(node u, node v) const {
        for (node w : adj_[u])
            if (w == v) return true;
        return false;
    }
    edgeid edgeId(node u, node v) const {
        for (index i = 0; i < adj_[u].size(); ++i)
            if (adj_[u][i] == v) return ids_[u][i];
        throw std::runtime_error("Edge does not exist");
    }
    node getIthNeighbor(node u, index i) const { return adj_[u][i]; }
    std::pair<node, edgeid> getIthNeighborWithId(node u, index i) const { return {adj_[u][i], ids_[u][i]}; }
    const std::vector<node> &neighborRange(node u) const { return adj_[u]; }
    std::vector<node> nodeRange() const {
        std::vector<node> r(adj_.size());
        for (node u = 0; u < r.size(); ++u) r[u] = u;
        return r;
    }
    // edges in NetworKit's iteration order: u ascending, adjacency position ascending, reported where u >= v
    std::vector<Edge> edgeRange() const {
        std::vector<Edge> r;
        for (node u = 0; u < adj_.size(); ++u)
            for (node v : adj_[u])
                if (v <= u) r.push_back({u, v});
        return r;
    }

private:
    std::vector<std::vector<node>> adj_;
    std::vector<std::vector<edgeid>> ids_;
    count m_ = 0;
    bool weighted_, directed_, indexed_ = false;
};

} // namespace NetworKit
