// Drop-in replacement for src/conshdlr/conshdlr_triangles.cpp of CharJon/SMS (see sms/conshdlr/conshdlr_triangles.hpp
// here).  Same handler properties (reference conshdlr_triangles.cpp:17-35), same callbacks and result-code protocol,
// same rows ("oc", lhs -inf, rhs numCrosses-1, -1 stay / +1 cross, flags local FALSE / modifiable FALSE / removable
// TRUE, SCIPaddRow with forcecut = false).  What changes: scip_initsol asks the device for the triangle list
// (smsgpu_small_cycles_enumerate instead of SmallChordlessCycles + one addTriangle per triangle, reference :193-203), and
// sepaTrianglesLp ships x* by edge id once and gets the violated triangles back from the K6 scan instead of three
// adjacency scans and two vector allocations per triangle (triangles.hpp:34-53).
#include "sms/conshdlr/conshdlr_triangles.hpp"

#include <algorithm>
#include <stdexcept>

#include "sms/probdata/mc_rooted.hpp"

#define CONSHDLR_NAME "triangleparity"
#define CONSHDLR_DESC "chordless three cycle parity constraint"
#define CONSHDLR_SEPAPRIORITY (+800000)
#define CONSHDLR_ENFOPRIORITY (-600000)
#define CONSHDLR_CHECKPRIORITY (-600000)
#define CONSHDLR_SEPAFREQ (1)
#define CONSHDLR_PROPFREQ (1)
#define CONSHDLR_EAGERFREQ (100)
#define CONSHDLR_MAXPREROUNDS (-1)
#define CONSHDLR_DELAYSEPA FALSE
#define CONSHDLR_DELAYPROP FALSE
#define CONSHDLR_NEEDSCONS FALSE
#define CONSHDLR_PROP_TIMING SCIP_PROPTIMING_BEFORELP
#define CONSHDLR_PRESOLTIMING SCIP_PRESOLTIMING_ALWAYS
#define CONSHDLR_SEPA_CUTPOOL FALSE

namespace sms {

ConshdlrTriangles::ConshdlrTriangles(Scip *scip, const NetworKit::Graph *sepGraph)
    : scip::ObjConshdlr(scip, CONSHDLR_NAME, CONSHDLR_DESC, CONSHDLR_SEPAPRIORITY, CONSHDLR_ENFOPRIORITY, CONSHDLR_CHECKPRIORITY,
                        CONSHDLR_SEPAFREQ, CONSHDLR_PROPFREQ, CONSHDLR_EAGERFREQ, CONSHDLR_MAXPREROUNDS, CONSHDLR_DELAYSEPA,
                        CONSHDLR_DELAYPROP, CONSHDLR_NEEDSCONS, CONSHDLR_PROP_TIMING, CONSHDLR_PRESOLTIMING),
      sepGraph_(sepGraph),
      triangleSeparator_(sepGraph), // throws std::runtime_error when the graph is unusable or no CUDA device exists
      sepaCutpool_(CONSHDLR_SEPA_CUTPOOL) {
    if (SCIPaddBoolParam(scip, "conshdlr/" CONSHDLR_NAME "/sepa_cutpool", "separate cuts into cutpool", &sepaCutpool_, FALSE,
                         CONSHDLR_SEPA_CUTPOOL, nullptr, nullptr) != SCIP_OKAY)
        throw std::runtime_error("SCIPaddBoolParam(conshdlr/triangleparity/sepa_cutpool) failed");
}

// reference :53-60: the handler never rejects a solution
SCIP_RETCODE ConshdlrTriangles::scip_check(SCIP *, SCIP_CONSHDLR *, SCIP_CONS **, int, SCIP_SOL *, SCIP_Bool, SCIP_Bool, SCIP_Bool, SCIP_Bool,
                                           SCIP_RESULT *result) {
    *result = SCIP_FEASIBLE;
    return SCIP_OKAY;
}

// reference :62-66
SCIP_RETCODE ConshdlrTriangles::scip_sepasol(SCIP *, SCIP_CONSHDLR *, SCIP_CONS **, int, int, SCIP_SOL *, SCIP_RESULT *) {
    throw std::runtime_error("Not implemented: scip_sepasol.");
}

SCIP_RETCODE ConshdlrTriangles::scip_sepalp(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss,
                                            SCIP_RESULT *result) {
    SCIP_CALL(sepaTrianglesLp(scip, conshdlr, conss, nconss, nusefulconss, result));
    return SCIP_OKAY;
}

// One round (reference :73-132)
SCIP_RETCODE ConshdlrTriangles::sepaTrianglesLp(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **, int, int, SCIP_RESULT *result) {
    *result = SCIP_DIDNOTFIND;
    auto *probData = dynamic_cast<ProbDataEdgesMc *>(SCIPgetObjProbData(scip));
    if (probData == nullptr) return SCIP_ERROR;
    // x* by edge id, one pass (the reference: edgeToVar(u, v) + setWeight(u, v), adjacency scans each, :86-91)
    const auto m = sepGraph_->numberOfEdges();
    std::vector<GpuOddClosedWalk> violated;
    try {
        for (NetworKit::index e = 0; e < m; ++e)
            triangleSeparator_.updateEdgeWeight(e, std::min(std::max(0., SCIPgetSolVal(scip, nullptr, probData->edgeIdToVar(e))), 1.));
        violated = triangleSeparator_.getViolatedTriangles();
    } catch (const std::exception &e) {
        SCIPerrorMessage("triangleparity: GPU scan failed: %s\n", e.what());
        return SCIP_ERROR;
    }
    for (const auto &ocw : violated) {
        if (*result == SCIP_CUTOFF) break;
        const SCIP_Real lhs = -SCIPinfinity(scip);
        const unsigned int rhs = ocw.numCrosses() - 1;
        SCIP_ROW *row;
        SCIP_CALL(SCIPcreateEmptyRowConshdlr(scip, &row, conshdlr, "oc", lhs, rhs, FALSE, FALSE, TRUE));
        SCIP_CALL(SCIPcacheRowExtensions(scip, row));
        for (uint64_t i = 0; i < ocw.size(); ++i)
            SCIP_CALL(SCIPaddVarToRow(scip, row, probData->edgeIdToVar(ocw.ithEdgeId(i)), ocw.ithEdgeIsStay(i) ? -1. : 1.));
        SCIP_CALL(SCIPflushRowExtensions(scip, row));
        if (SCIPisCutEfficacious(scip, nullptr, row)) {
            *result = SCIP_SEPARATED;
            if (sepaCutpool_) {
                SCIP_CALL(SCIPaddPoolCut(scip, row));
            } else {
                SCIP_Bool infeasible;
                SCIP_CALL(SCIPaddRow(scip, row, FALSE, &infeasible));
                if (infeasible) *result = SCIP_CUTOFF;
            }
        }
        SCIP_CALL(SCIPreleaseRow(scip, &row));
    }
    return SCIP_OKAY;
}

// CONSENFOLP (reference :135-153)
SCIP_RETCODE ConshdlrTriangles::scip_enfolp(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss, SCIP_Bool,
                                            SCIP_RESULT *result) {
    *result = SCIP_INFEASIBLE;
    SCIP_RESULT resultSepa;
    SCIP_CALL(sepaTrianglesLp(scip, conshdlr, conss, nconss, nusefulconss, &resultSepa));
    *result = (resultSepa == SCIP_SEPARATED || resultSepa == SCIP_NEWROUND) ? SCIP_SEPARATED : SCIP_FEASIBLE;
    return SCIP_OKAY;
}

// CONSENFOPS (reference :156-163): not implemented there either
SCIP_RETCODE ConshdlrTriangles::scip_enfops(SCIP *, SCIP_CONSHDLR *, SCIP_CONS **, int, int, SCIP_Bool, SCIP_Bool, SCIP_RESULT *result) {
    *result = SCIP_FEASIBLE;
    throw std::runtime_error("Not implemented: scip_enfops");
}

// CONSLOCK (reference :165-169): no locks
SCIP_RETCODE ConshdlrTriangles::scip_lock(SCIP *, SCIP_CONSHDLR *, SCIP_CONS *, SCIP_LOCKTYPE, int, int) { return SCIP_OKAY; }

// CONSTRANS (reference :171-185)
SCIP_RETCODE ConshdlrTriangles::scip_trans(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS *sourcecons, SCIP_CONS **targetcons) {
    SCIP_CALL(SCIPcreateCons(scip, targetcons, SCIPconsGetName(sourcecons), conshdlr, nullptr, SCIPconsIsInitial(sourcecons),
                             SCIPconsIsSeparated(sourcecons), SCIPconsIsEnforced(sourcecons), SCIPconsIsChecked(sourcecons),
                             SCIPconsIsPropagated(sourcecons), SCIPconsIsLocal(sourcecons), SCIPconsIsModifiable(sourcecons),
                             SCIPconsIsDynamic(sourcecons), SCIPconsIsRemovable(sourcecons), SCIPconsIsStickingAtNode(sourcecons)));
    return SCIP_OKAY;
}

scip::ObjProbCloneable *ConshdlrTriangles::clone(SCIP *scip, SCIP_Bool *valid) const {
    *valid = TRUE;
    return new ConshdlrTriangles(scip, sepGraph_);
}

// CONSINITSOL (reference :193-203): the triangle list, from the device
SCIP_RETCODE ConshdlrTriangles::scip_initsol(SCIP *, SCIP_CONSHDLR *, SCIP_CONS **, int) {
    try {
        numTriangles_ = triangleSeparator_.enumerate();
    } catch (const std::exception &e) {
        SCIPerrorMessage("triangleparity: triangle enumeration failed: %s\n", e.what());
        return SCIP_ERROR;
    }
    return SCIP_OKAY;
}

} // namespace sms
