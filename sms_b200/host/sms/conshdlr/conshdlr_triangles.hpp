// Drop-in replacement for include/sms/conshdlr/conshdlr_triangles.hpp of CharJon/SMS: the SCIP constraint handler
// "triangleparity" with the same class name, constructor, handler properties, SCIP parameter and callbacks
// (reference: include/sms/conshdlr/conshdlr_triangles.hpp, src/conshdlr/conshdlr_triangles.cpp), whose separator is
// sms::GpuTriangleSeparator (libsmsgpu.so, kernel K6) instead of sms::TriangleSeparator, and whose triangle list comes
// from the device enumeration instead of SmallChordlessCycles.  BasicOcwSolver still does
// `new ConshdlrTriangles(scip_, &graph_)` (src/solver/basic_ocw.cpp).
#ifndef SMS_CONSHDLRTRIANGLES_GPU_HPP
#define SMS_CONSHDLRTRIANGLES_GPU_HPP

#include <cstdint>

#include "networkit/graph/Graph.hpp"
#include "objscip/objscip.h"

#include "gpu_ocw_separator.hpp"

namespace sms {

class ConshdlrTriangles : public scip::ObjConshdlr {
private:
    NetworKit::Graph const *const sepGraph_;
    sms::GpuTriangleSeparator triangleSeparator_; // owns a GPU context (graph replica uploaded once, here)
    SCIP_Bool sepaCutpool_;
    int64_t numTriangles_ = 0;

public:
    explicit ConshdlrTriangles(Scip *scip, const NetworKit::Graph *sepGraph);
    ~ConshdlrTriangles() override = default;

    SCIP_DECL_CONSCHECK(scip_check) override;
    SCIP_DECL_CONSENFOLP(scip_enfolp) override;
    SCIP_DECL_CONSENFOPS(scip_enfops) override;
    SCIP_DECL_CONSLOCK(scip_lock) override;
    SCIP_DECL_CONSTRANS(scip_trans) override;
    SCIP_DECL_CONSSEPASOL(scip_sepasol) override;
    SCIP_DECL_CONSSEPALP(scip_sepalp) override;
    SCIP_DECL_CONSINITSOL(scip_initsol) override;
    scip::ObjProbCloneable *SCIP_DECL_CONSHDLRCLONE(clone) override;
    SCIP_Bool iscloneable() const override { return TRUE; }

    int64_t numTriangles() const { return numTriangles_; }

private:
    SCIP_RETCODE sepaTrianglesLp(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss, SCIP_RESULT *result);
};

} // namespace sms

#endif // SMS_CONSHDLRTRIANGLES_GPU_HPP
