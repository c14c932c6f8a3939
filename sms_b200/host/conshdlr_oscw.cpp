// Drop-in replacement for src/conshdlr/conshdlr_oscw.cpp of CharJon/SMS (see sms/conshdlr/conshdlr_oscw.hpp here).
// Same handler properties (reference conshdlr_oscw.cpp:12-30), same callbacks and result-code protocol, same rows
// ("osc", lhs -inf, rhs numCrosses-1, coefficient -1 stay / +1 cross, flags local FALSE / modifiable FALSE /
// removable TRUE).  What changes is sepaOddCycles: instead of n calls of OcwSeparator::getViolatedOcws it ships
// x* and the blacklist to the device once and gets the whole round's cycles back from libsmsgpu.so.
#include "sms/conshdlr/conshdlr_oscw.hpp"

#include <algorithm>
#include <stdexcept>

#include "sms/auxiliary/scip.hpp"
#include "sms/probdata/mc_rooted.hpp"

#define CONSHDLR_NAME "cycleparity"
#define CONSHDLR_DESC "cycle parity constraints"
#define CONSHDLR_SEPAPRIORITY (+600000)
#define CONSHDLR_ENFOPRIORITY (-600000)
#define CONSHDLR_CHECKPRIORITY (-600000)
#define CONSHDLR_SEPAFREQ 5
#define CONSHDLR_PROPFREQ 1
#define CONSHDLR_EAGERFREQ 100
#define CONSHDLR_MAXPREROUNDS (-1)
#define CONSHDLR_DELAYSEPA TRUE
#define CONSHDLR_DELAYPROP FALSE
#define CONSHDLR_NEEDSCONS TRUE
#define CONSHDLR_PROP_TIMING SCIP_PROPTIMING_BEFORELP
#define CONSHDLR_PRESOLTIMING SCIP_PRESOLTIMING_ALWAYS
#define CONSHDLR_SEPA_CUTPOOL TRUE

namespace sms {

ConshdlrOscw::ConshdlrOscw(Scip *scip, const NetworKit::Graph *sepGraph)
    : scip::ObjConshdlr(scip, CONSHDLR_NAME, CONSHDLR_DESC, CONSHDLR_SEPAPRIORITY, CONSHDLR_ENFOPRIORITY, CONSHDLR_CHECKPRIORITY,
                        CONSHDLR_SEPAFREQ, CONSHDLR_PROPFREQ, CONSHDLR_EAGERFREQ, CONSHDLR_MAXPREROUNDS, CONSHDLR_DELAYSEPA,
                        CONSHDLR_DELAYPROP, CONSHDLR_NEEDSCONS, CONSHDLR_PROP_TIMING, CONSHDLR_PRESOLTIMING),
      sepGraph_(sepGraph),
      ocwSeparator_(*sepGraph), // throws std::runtime_error like OcwSeparator's ctor, and when no CUDA device exists
      sepaCutpool_(CONSHDLR_SEPA_CUTPOOL),
      xByEid_(sepGraph->numberOfEdges(), 0.),
      blacklist_(sepGraph->upperNodeIdBound(), 0) {
    // the reference wraps this in SCIP_CALL_EXC_NO_THROW (print and exit); a failing registration is equally fatal here
    if (SCIPaddBoolParam(scip, "constraints/" CONSHDLR_NAME "/sepacutpool", "separate cuts into cutpool", &sepaCutpool_, FALSE,
                         CONSHDLR_SEPA_CUTPOOL, nullptr, nullptr) != SCIP_OKAY)
        throw std::runtime_error("SCIPaddBoolParam(constraints/cycleparity/sepacutpool) failed");
}

SCIP_RETCODE SCIPcreateConsOddCyclesFast(SCIP *scip, SCIP_CONS **cons, const char *name, const NetworKit::Graph * /*graph*/,
                                         SCIP_Bool initial, SCIP_Bool separate, SCIP_Bool enforce, SCIP_Bool check, SCIP_Bool propagate,
                                         SCIP_Bool local, SCIP_Bool modifiable, SCIP_Bool dynamic, SCIP_Bool removable) {
    SCIP_CONSHDLR *conshdlr = SCIPfindConshdlr(scip, CONSHDLR_NAME);
    if (conshdlr == nullptr) {
        SCIPerrorMessage("constraint handler not found\n");
        return SCIP_PLUGINNOTFOUND;
    }
    SCIP_CALL(SCIPcreateCons(scip, cons, name, conshdlr, nullptr, initial, separate, enforce, check, propagate, local, modifiable,
                             dynamic, removable, FALSE));
    return SCIP_OKAY;
}

SCIP_RETCODE SCIPcreateConsBasicOddCyclesFast(SCIP *scip, SCIP_CONS **cons, const char *name, const NetworKit::Graph *graph) {
    return SCIPcreateConsOddCyclesFast(scip, cons, name, graph, FALSE, TRUE, TRUE, TRUE, TRUE, FALSE, FALSE, FALSE, TRUE);
}

// CONSCHECK: an integral edge vector is a cut iff the labelling side[v] = side[u] xor x_uv is consistent
// (reference conshdlr_oscw.cpp:96-144).  Host code, not hot.
SCIP_RETCODE ConshdlrOscw::scip_check(SCIP *scip, SCIP_CONSHDLR *, SCIP_CONS **, int nconss, SCIP_SOL *sol, SCIP_Bool, SCIP_Bool,
                                      SCIP_Bool, SCIP_Bool, SCIP_RESULT *result) {
    auto *probData = dynamic_cast<ProbDataEdgesMc *>(SCIPgetObjProbData(scip));
    if (probData == nullptr) return SCIP_ERROR;
    const auto n = sepGraph_->upperNodeIdBound();
    for (int c = 0; c < nconss; ++c) {
        std::vector<int8_t> side(n, -1);
        std::vector<NetworKit::node> stack;
        for (NetworKit::node r = 0; r < n; ++r) {
            if (side[r] >= 0) continue;
            side[r] = 0;
            stack.push_back(r);
            while (!stack.empty()) {
                const auto u = stack.back();
                stack.pop_back();
                for (auto v : sepGraph_->neighborRange(u)) {
                    const int bit = SCIPgetSolVal(scip, sol, probData->edgeToVar(u, v)) < 0.5 ? 0 : 1;
                    if (side[v] < 0) {
                        side[v] = static_cast<int8_t>(side[u] ^ bit);
                        stack.push_back(v);
                    } else if (side[v] != (side[u] ^ bit)) {
                        *result = SCIP_INFEASIBLE;
                        return SCIP_OKAY;
                    }
                }
            }
        }
    }
    *result = SCIP_FEASIBLE;
    return SCIP_OKAY;
}

SCIP_RETCODE ConshdlrOscw::scip_sepasol(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss,
                                        SCIP_SOL *sol, SCIP_RESULT *result) {
    SCIP_CALL(sepaOddCycles(scip, conshdlr, conss, nconss, nusefulconss, sol, result));
    return SCIP_OKAY;
}

SCIP_RETCODE ConshdlrOscw::scip_sepalp(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss,
                                       SCIP_RESULT *result) {
    SCIP_CALL(sepaOddCyclesLp(scip, conshdlr, conss, nconss, nusefulconss, result));
    return SCIP_OKAY;
}

// One separation round (reference conshdlr_oscw.cpp:165-247).
SCIP_RETCODE ConshdlrOscw::sepaOddCycles(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **, int, int, SCIP_SOL *sol,
                                         SCIP_RESULT *result) {
    *result = SCIP_DIDNOTFIND;
    auto *probDataMc = dynamic_cast<ProbDataRootedMc *>(SCIPgetObjProbData(scip));
    if (probDataMc == nullptr) return SCIP_ERROR;

    // round input: x* by edge id (edgeIdToVar, mc_rooted.hpp:71 -- no adjacency scans) and the blacklist
    // {root} + {locally fixed node variables} (reference :175-185)
    const auto m = sepGraph_->numberOfEdges();
    for (NetworKit::index e = 0; e < m; ++e)
        xByEid_[e] = std::min(std::max(0., SCIPgetSolVal(scip, sol, probDataMc->edgeIdToVar(e))), 1.);
    const auto root = probDataMc->getRoot();
    for (auto u : sepGraph_->nodeRange())
        blacklist_[u] = (u == root || scipVarIsFixedLocal(scip, probDataMc->nodeToVar(u))) ? 1 : 0;

    std::vector<GpuOddClosedWalk> walks;
    try {
        ocwSeparator_.setRoundInput(xByEid_.data(), blacklist_.data());
        // sources 0..n-1 as the reference issues them (:188-195); blacklisted ones are skipped on the device
        walks = ocwSeparator_.separateRound(nullptr, 0, /*dedup=*/true, sweepMode_, deviceMinEfficacy_);
    } catch (const std::exception &e) {
        SCIPerrorMessage("cycleparity: GPU separation failed: %s\n", e.what());
        return SCIP_ERROR;
    }
    roundsRun_++;
    cutsFromDevice_ += walks.size();

    // rows (reference :188-243).  The list is ordered by (first-finding source, emission index).  The reference's protocol,
    // kept as it is: the loop over the sources stops in front of the next source once the result is CUTOFF (:188), the walks of
    // the current source are still processed, and a later efficacious row of that source that is feasible overwrites CUTOFF
    // with SEPARATED (:231-237); efficacy is always asked for the LP solution (sol = nullptr, :221/232) even in CONSSEPASOL.
    int32_t curSrc = -1;
    for (const auto &ocw : walks) {
        if (ocw.sourceIndex() != curSrc) {
            if (*result == SCIP_CUTOFF) break;
            curSrc = ocw.sourceIndex();
        }
        const SCIP_Real lhs = -SCIPinfinity(scip);
        const unsigned int rhs = ocw.numCrosses() - 1;
        SCIP_ROW *row;
        SCIP_CALL(SCIPcreateEmptyRowConshdlr(scip, &row, conshdlr, "osc", lhs, rhs, FALSE, FALSE, TRUE));
        SCIP_CALL(SCIPcacheRowExtensions(scip, row));
        for (uint64_t i = 0; i < ocw.size(); ++i) {
            SCIP_CALL(SCIPaddVarToRow(scip, row, probDataMc->edgeIdToVar(ocw.ithEdgeId(i)), ocw.ithEdgeIsStay(i) ? -1. : 1.));
            totalConstructedCycleLength_ += 1;
        }
        SCIP_CALL(SCIPflushRowExtensions(scip, row));
        if (sepaCutpool_) {
            if (SCIPisCutEfficacious(scip, nullptr, row)) {
                SCIP_CALL(SCIPaddPoolCut(scip, row));
                *result = SCIP_SEPARATED;
            }
        } else if (SCIPisCutEfficacious(scip, nullptr, row)) {
            SCIP_Bool infeasible;
            SCIP_CALL(SCIPaddRow(scip, row, TRUE, &infeasible));
            *result = infeasible ? SCIP_CUTOFF : SCIP_SEPARATED;
        }
        SCIP_CALL(SCIPreleaseRow(scip, &row));
    }
    return SCIP_OKAY;
}

SCIP_RETCODE ConshdlrOscw::sepaOddCyclesLp(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss,
                                           SCIP_RESULT *result) {
    return sepaOddCycles(scip, conshdlr, conss, nconss, nusefulconss, nullptr, result);
}

// CONSENFOLP (reference :261-272): the same round; "nothing found" means the LP solution is feasible for the handler
SCIP_RETCODE ConshdlrOscw::scip_enfolp(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss,
                                       SCIP_Bool, SCIP_RESULT *result) {
    *result = SCIP_INFEASIBLE;
    SCIP_CALL(sepaOddCyclesLp(scip, conshdlr, conss, nconss, nusefulconss, result));
    if (*result == SCIP_DIDNOTFIND) *result = SCIP_FEASIBLE;
    return SCIP_OKAY;
}

// CONSENFOPS (reference :275-283): not implemented there either
SCIP_RETCODE ConshdlrOscw::scip_enfops(SCIP *, SCIP_CONSHDLR *, SCIP_CONS **, int, int, SCIP_Bool, SCIP_Bool, SCIP_RESULT *result) {
    *result = SCIP_FEASIBLE;
    throw std::runtime_error("Not implemented: scip_enfops");
}

// CONSLOCK (reference :289-303): every edge variable can violate the constraint in both directions
SCIP_RETCODE ConshdlrOscw::scip_lock(SCIP *scip, SCIP_CONSHDLR *, SCIP_CONS *, SCIP_LOCKTYPE, int nlockspos, int nlocksneg) {
    auto *probDataMc = dynamic_cast<ProbDataEdgesMc *>(SCIPgetObjProbData(scip));
    if (probDataMc == nullptr) return SCIP_ERROR;
    for (auto e : probDataMc->getMcGraph()->edgeRange())
        SCIP_CALL(SCIPaddVarLocksType(scip, probDataMc->edgeToVar(e.u, e.v), SCIP_LOCKTYPE_MODEL, nlockspos + nlocksneg,
                                      nlockspos + nlocksneg));
    return SCIP_OKAY;
}

// CONSTRANS (reference :306-323): the constraint carries no data
SCIP_RETCODE ConshdlrOscw::scip_trans(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS *sourcecons, SCIP_CONS **targetcons) {
    SCIP_CALL(SCIPcreateCons(scip, targetcons, SCIPconsGetName(sourcecons), conshdlr, nullptr, SCIPconsIsInitial(sourcecons),
                             SCIPconsIsSeparated(sourcecons), SCIPconsIsEnforced(sourcecons), SCIPconsIsChecked(sourcecons),
                             SCIPconsIsPropagated(sourcecons), SCIPconsIsLocal(sourcecons), SCIPconsIsModifiable(sourcecons),
                             SCIPconsIsDynamic(sourcecons), SCIPconsIsRemovable(sourcecons), SCIPconsIsStickingAtNode(sourcecons)));
    return SCIP_OKAY;
}

// clone (reference :326-330): a sub-SCIP gets its own handler and therefore its own GPU context on the same graph
scip::ObjProbCloneable *ConshdlrOscw::clone(SCIP *scip, SCIP_Bool *valid) const {
    *valid = TRUE;
    return new ConshdlrOscw(scip, sepGraph_);
}

} // namespace sms
