/* HEADLESS-TEST STAND-IN for <objscip/objscip.h>: the C++ wrapper classes the handlers derive from, with the
 * constructor signature and virtual callbacks of SCIP's objconshdlr.h / objprobdata.h / objcloneable.h.
 * SCIPincludeObjConshdlr registers the object with the fake SCIP so that SCIPfindConshdlr works. */
#pragma once
#include "scip/scip.h"

namespace scip {

struct ObjProbCloneable {
    virtual ~ObjProbCloneable() = default;
    virtual ObjProbCloneable *clone(SCIP *, SCIP_Bool *valid) const {
        if (valid) *valid = FALSE;
        return nullptr;
    }
    virtual SCIP_Bool iscloneable() const { return FALSE; }
};

class ObjProbData {
public:
    virtual ~ObjProbData() = default;
};

class ObjConshdlr : public ObjProbCloneable {
public:
    SCIP *scip_;
    std::string scip_name_, scip_desc_;
    const int scip_sepapriority_, scip_enfopriority_, scip_checkpriority_, scip_sepafreq_, scip_propfreq_, scip_eagerfreq_,
        scip_maxprerounds_;
    const SCIP_Bool scip_delaysepa_, scip_delayprop_, scip_needscons_;
    const SCIP_PROPTIMING scip_proptiming_;
    const SCIP_PRESOLTIMING scip_presoltiming_;

    ObjConshdlr(SCIP *scip, const char *name, const char *desc, int sepapriority, int enfopriority, int checkpriority, int sepafreq,
                int propfreq, int eagerfreq, int maxprerounds, SCIP_Bool delaysepa, SCIP_Bool delayprop, SCIP_Bool needscons,
                SCIP_PROPTIMING proptiming, SCIP_PRESOLTIMING presoltiming)
        : scip_(scip), scip_name_(name), scip_desc_(desc), scip_sepapriority_(sepapriority), scip_enfopriority_(enfopriority),
          scip_checkpriority_(checkpriority), scip_sepafreq_(sepafreq), scip_propfreq_(propfreq), scip_eagerfreq_(eagerfreq),
          scip_maxprerounds_(maxprerounds), scip_delaysepa_(delaysepa), scip_delayprop_(delayprop), scip_needscons_(needscons),
          scip_proptiming_(proptiming), scip_presoltiming_(presoltiming) {}

    virtual SCIP_DECL_CONSTRANS(scip_trans) = 0;
    virtual SCIP_DECL_CONSSEPALP(scip_sepalp) {
        (void)scip, (void)conshdlr, (void)conss, (void)nconss, (void)nusefulconss;
        *result = SCIP_DIDNOTRUN;
        return SCIP_OKAY;
    }
    virtual SCIP_DECL_CONSSEPASOL(scip_sepasol) {
        (void)scip, (void)conshdlr, (void)conss, (void)nconss, (void)nusefulconss, (void)sol;
        *result = SCIP_DIDNOTRUN;
        return SCIP_OKAY;
    }
    virtual SCIP_DECL_CONSINITSOL(scip_initsol) {
        (void)scip, (void)conshdlr, (void)conss, (void)nconss;
        return SCIP_OKAY;
    }
    virtual SCIP_DECL_CONSENFOLP(scip_enfolp) = 0;
    virtual SCIP_DECL_CONSENFOPS(scip_enfops) = 0;
    virtual SCIP_DECL_CONSCHECK(scip_check) = 0;
    virtual SCIP_DECL_CONSLOCK(scip_lock) = 0;
};

} // namespace scip

inline SCIP_RETCODE SCIPincludeObjConshdlr(SCIP *scip, scip::ObjConshdlr *obj, SCIP_Bool /*deleteobject*/) {
    auto *h = new SCIP_Conshdlr{obj->scip_name_, obj};
    scip->conshdlrs.push_back(h);
    return SCIP_OKAY;
}
