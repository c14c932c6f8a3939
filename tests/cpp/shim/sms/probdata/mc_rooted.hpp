// HEADLESS-TEST STAND-IN for sms/probdata/mc_rooted.hpp (+ mc_edges.hpp): the members of ProbDataEdgesMc /
// ProbDataRootedMc the handlers call -- edgeToVar, edgeIdToVar, nodeToVar, getRoot, getMcGraph
// (include/sms/probdata/mc_rooted.hpp:61-99, mc_edges.hpp).  Inside SMS the real classes are used, unchanged.
#pragma once
#include <memory>
#include <vector>

#include "networkit/graph/Graph.hpp"
#include "objscip/objscip.h"

class ProbDataEdgesMc : public scip::ObjProbData {
public:
    virtual SCIP_VAR *edgeToVar(NetworKit::node u, NetworKit::node v) const = 0;
    virtual SCIP_VAR *edgeIdToVar(NetworKit::index edgeId) const = 0;
    virtual const NetworKit::Graph *getMcGraph() const = 0;
};

class ProbDataRootedMc : public ProbDataEdgesMc {
public:
    // one binary variable per edge (index = edge id) and per non-root node (index = m + node)
    ProbDataRootedMc(SCIP *, const NetworKit::Graph *g, NetworKit::node root) : mcGraph_(g), root_(root) {
        const auto m = g->upperEdgeIdBound(), n = g->upperNodeIdBound();
        for (NetworKit::index e = 0; e < m; ++e) {
            vars_.push_back(std::make_unique<SCIP_Var>());
            vars_.back()->index = static_cast<int>(e);
            edgeIdToVar_.push_back(vars_.back().get());
        }
        for (NetworKit::node u = 0; u < n; ++u) {
            vars_.push_back(std::make_unique<SCIP_Var>());
            vars_.back()->index = static_cast<int>(m + u);
            nodeToVar_.push_back(vars_.back().get());
        }
    }
    SCIP_VAR *edgeToVar(NetworKit::node u, NetworKit::node v) const override { return edgeIdToVar(mcGraph_->edgeId(u, v)); }
    SCIP_VAR *edgeIdToVar(NetworKit::index edgeId) const override { return edgeIdToVar_[edgeId]; }
    SCIP_VAR *nodeToVar(NetworKit::node u) const { return nodeToVar_[u]; }
    NetworKit::node getRoot() const { return root_; }
    const NetworKit::Graph *getMcGraph() const override { return mcGraph_; }

private:
    const NetworKit::Graph *mcGraph_;
    NetworKit::node root_;
    std::vector<std::unique_ptr<SCIP_Var>> vars_;
    std::vector<SCIP_VAR *> edgeIdToVar_, nodeToVar_;
};
