// Drop-in replacement for include/sms/conshdlr/conshdlr_oscw.hpp of CharJon/SMS: the SCIP constraint handler
// "cycleparity" with the same class name, constructor, handler properties, SCIP parameter and callbacks
// (reference: include/sms/conshdlr/conshdlr_oscw.hpp:21-258, src/conshdlr/conshdlr_oscw.cpp), whose separation
// engine is sms::GpuOcwSeparator (libsmsgpu.so) instead of sms::OcwSeparator.  Nothing else in SMS changes:
// BasicOcwSolver still does `new ConshdlrOscw(scip_, &graph_)` (src/solver/basic_ocw.cpp:155-161).
#ifndef SMS_CONSHDLR_OC_FAST_HPP
#define SMS_CONSHDLR_OC_FAST_HPP

#include <cstdint>

#include "networkit/graph/Graph.hpp"
#include "objscip/objscip.h"
#include "scip/cons_linear.h"

#include "gpu_ocw_separator.hpp"

namespace sms {

class ConshdlrOscw : public scip::ObjConshdlr {
private:
    NetworKit::Graph const *const sepGraph_;
    sms::GpuOcwSeparator ocwSeparator_; // owns the GPU context: CSR uploaded once, here
    SCIP_Bool sepaCutpool_;
    std::vector<double> xByEid_;        // round input staging: x* by edge id, blacklist by vertex
    std::vector<uint8_t> blacklist_;

    // Statistics (the reference keeps totalConstructedCycleLength_, conshdlr_oscw.hpp:30)
    uint64_t totalConstructedCycleLength_ = 0;
    uint64_t roundsRun_ = 0, cutsFromDevice_ = 0;
    int sweepMode_ = 0;
    double deviceMinEfficacy_ = 0.;

public:
    explicit ConshdlrOscw(Scip *scip, const NetworKit::Graph *sepGraph);
    ~ConshdlrOscw() override = default;

    SCIP_DECL_CONSCHECK(scip_check) override;
    SCIP_DECL_CONSENFOLP(scip_enfolp) override;
    SCIP_DECL_CONSENFOPS(scip_enfops) override;
    SCIP_DECL_CONSLOCK(scip_lock) override;
    SCIP_DECL_CONSTRANS(scip_trans) override;
    SCIP_DECL_CONSSEPASOL(scip_sepasol) override;
    SCIP_DECL_CONSSEPALP(scip_sepalp) override;
    scip::ObjProbCloneable *SCIP_DECL_CONSHDLRCLONE(clone) override;
    SCIP_Bool iscloneable() const override { return TRUE; }

    // SMSGPU_MODE_INDEPENDENT (default) or SMSGPU_MODE_VERBATIM: which sweep contract the device runs (include/smsgpu.h)
    void setSweepMode(int mode) { sweepMode_ = mode; }
    // efficacy pre-filter on the device (smsgpu_params::min_efficacy); BasicOcwSolver would pass the value it sets for
    // separating/minefficacyroot (src/solver/basic_ocw.cpp:244).  0 = off; SCIPisCutEfficacious still decides every survivor.
    void setDeviceMinEfficacy(double e) { deviceMinEfficacy_ = e; }
    uint64_t totalConstructedCycleLength() const { return totalConstructedCycleLength_; }
    uint64_t roundsRun() const { return roundsRun_; }
    uint64_t cutsFromDevice() const { return cutsFromDevice_; }

private:
    SCIP_RETCODE sepaOddCycles(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss, SCIP_SOL *sol,
                               SCIP_RESULT *result);
    SCIP_RETCODE sepaOddCyclesLp(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss,
                                 SCIP_RESULT *result);
};

// same two factory functions as the reference (conshdlr_oscw.cpp:48-91)
SCIP_RETCODE SCIPcreateConsOddCyclesFast(SCIP *scip, SCIP_CONS **cons, const char *name, const NetworKit::Graph *graph, SCIP_Bool initial,
                                         SCIP_Bool separate, SCIP_Bool enforce, SCIP_Bool check, SCIP_Bool propagate, SCIP_Bool local,
                                         SCIP_Bool modifiable, SCIP_Bool dynamic, SCIP_Bool removable);
SCIP_RETCODE SCIPcreateConsBasicOddCyclesFast(SCIP *scip, SCIP_CONS **cons, const char *name, const NetworKit::Graph *graph);

} // namespace sms

#endif // SMS_CONSHDLR_OC_FAST_HPP
