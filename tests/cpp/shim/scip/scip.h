/* HEADLESS-TEST STAND-IN for <scip/scip.h>: SCIP is not installed in this image (nor on the GPU box), so the
 * cycleparity / triangleparity handlers of this repository are compiled and exercised against this recording
 * fake.  It declares exactly the SCIP symbols the two reference handlers use (list taken
 * from src/conshdlr/conshdlr_oscw.cpp and conshdlr_triangles.cpp), with SCIP's own signatures, and implements
 * them as simple in-memory bookkeeping: rows that would go to the cut pool / the LP are appended to vectors the
 * test reads back.  Inside SMS the real SCIP headers are used instead (this directory is not on the include path).
 * Efficacy follows SCIP's definition for an LP solution: (activity - rhs) / ||a||_2, efficacious iff
 * > separating/minefficacyroot (1e-6 in src/solver/basic_ocw.cpp:244). */
#pragma once
#include <cmath>
#include <cstdio>
#include <map>
#include <string>
#include <utility>
#include <vector>

typedef double SCIP_Real;
typedef unsigned int SCIP_Bool;
#ifndef TRUE
#define TRUE 1u
#define FALSE 0u
#endif

enum SCIP_Retcode { SCIP_OKAY = 1, SCIP_ERROR = 0, SCIP_PLUGINNOTFOUND = -11 };
typedef enum SCIP_Retcode SCIP_RETCODE;
enum SCIP_Result {
    SCIP_DIDNOTRUN = 1, SCIP_DELAYED = 2, SCIP_DIDNOTFIND = 3, SCIP_FEASIBLE = 4, SCIP_INFEASIBLE = 5, SCIP_UNBOUNDED = 6,
    SCIP_CUTOFF = 7, SCIP_SEPARATED = 8, SCIP_NEWROUND = 9
};
typedef enum SCIP_Result SCIP_RESULT;
enum SCIP_LockType { SCIP_LOCKTYPE_MODEL = 0, SCIP_LOCKTYPE_CONFLICT = 1 };
typedef enum SCIP_LockType SCIP_LOCKTYPE;
typedef unsigned int SCIP_PROPTIMING;
typedef unsigned int SCIP_PRESOLTIMING;
#define SCIP_PROPTIMING_BEFORELP 0x001u
#define SCIP_PRESOLTIMING_ALWAYS 0x01Cu

struct SCIP_Var {
    int index = 0;
    SCIP_Real lbLocal = 0., ubLocal = 1.;
    std::string name;
    int nlocksdown = 0, nlocksup = 0;
};
typedef struct SCIP_Var SCIP_VAR;
struct SCIP_Sol {
    std::vector<SCIP_Real> val; /* by variable index */
};
typedef struct SCIP_Sol SCIP_SOL;
struct SCIP_Conshdlr {
    std::string name;
    void *obj = nullptr; /* scip::ObjConshdlr* */
};
typedef struct SCIP_Conshdlr SCIP_CONSHDLR;
struct SCIP_ConsData;
typedef struct SCIP_ConsData SCIP_CONSDATA;
struct SCIP_Cons {
    std::string name;
    SCIP_CONSHDLR *hdlr = nullptr;
    SCIP_Bool initial, separate, enforce, check, propagate, local, modifiable, dynamic, removable, stickingatnode;
};
typedef struct SCIP_Cons SCIP_CONS;
struct SCIP_Row {
    std::string name;
    SCIP_CONSHDLR *origin = nullptr;
    SCIP_Real lhs = 0., rhs = 0.;
    SCIP_Bool local = FALSE, modifiable = FALSE, removable = FALSE;
    std::vector<std::pair<SCIP_VAR *, SCIP_Real>> terms;
    int nuses = 1;
};
typedef struct SCIP_Row SCIP_ROW;

namespace scip {
class ObjProbData;
}

struct Scip {
    std::vector<SCIP_Real> lpsol;               /* current LP solution by variable index (sol == nullptr) */
    scip::ObjProbData *probdata = nullptr;
    std::vector<SCIP_CONSHDLR *> conshdlrs;
    std::map<std::string, SCIP_Bool *> boolParams;
    SCIP_Real minEfficacy = 1e-6;               /* separating/minefficacyroot, basic_ocw.cpp:244 */
    std::vector<SCIP_ROW> poolCuts, lpRows;     /* what SCIPaddPoolCut / SCIPaddRow received, in call order */
    long nRowsCreated = 0, nRowsReleased = 0, nAddVar = 0;
    bool nextAddRowInfeasible = false;          /* test hook for the CUTOFF branch */
    long infeasibleAtAddRow = -1, nAddRow = 0;  /* test hook: the k-th SCIPaddRow call (0-based) reports infeasible */
};
typedef struct Scip SCIP;

#define SCIPerrorMessage(...) std::fprintf(stderr, __VA_ARGS__)
#define SCIP_CALL(x)                                       \
    do {                                                   \
        SCIP_RETCODE scip_call_rc_ = (x);                  \
        if (scip_call_rc_ != SCIP_OKAY) return scip_call_rc_; \
    } while (0)

inline SCIP_Real SCIPinfinity(SCIP *) { return 1e20; }
inline SCIP_RETCODE SCIPaddBoolParam(SCIP *scip, const char *name, const char *, SCIP_Bool *valueptr, SCIP_Bool, SCIP_Bool defaultvalue,
                                     void *, void *) {
    if (valueptr) *valueptr = defaultvalue;
    scip->boolParams[name] = valueptr;
    return SCIP_OKAY;
}
inline SCIP_CONSHDLR *SCIPfindConshdlr(SCIP *scip, const char *name) {
    for (auto *h : scip->conshdlrs)
        if (h->name == name) return h;
    return nullptr;
}
inline SCIP_RETCODE SCIPcreateCons(SCIP *, SCIP_CONS **cons, const char *name, SCIP_CONSHDLR *conshdlr, SCIP_CONSDATA *, SCIP_Bool initial,
                                   SCIP_Bool separate, SCIP_Bool enforce, SCIP_Bool check, SCIP_Bool propagate, SCIP_Bool local,
                                   SCIP_Bool modifiable, SCIP_Bool dynamic, SCIP_Bool removable, SCIP_Bool stickingatnode) {
    *cons = new SCIP_Cons{name, conshdlr, initial, separate, enforce, check, propagate, local, modifiable, dynamic, removable, stickingatnode};
    return SCIP_OKAY;
}
inline SCIP_RETCODE SCIPreleaseCons(SCIP *, SCIP_CONS **cons) {
    delete *cons;
    *cons = nullptr;
    return SCIP_OKAY;
}
inline const char *SCIPconsGetName(SCIP_CONS *c) { return c->name.c_str(); }
inline SCIP_Bool SCIPconsIsInitial(SCIP_CONS *c) { return c->initial; }
inline SCIP_Bool SCIPconsIsSeparated(SCIP_CONS *c) { return c->separate; }
inline SCIP_Bool SCIPconsIsEnforced(SCIP_CONS *c) { return c->enforce; }
inline SCIP_Bool SCIPconsIsChecked(SCIP_CONS *c) { return c->check; }
inline SCIP_Bool SCIPconsIsPropagated(SCIP_CONS *c) { return c->propagate; }
inline SCIP_Bool SCIPconsIsLocal(SCIP_CONS *c) { return c->local; }
inline SCIP_Bool SCIPconsIsModifiable(SCIP_CONS *c) { return c->modifiable; }
inline SCIP_Bool SCIPconsIsDynamic(SCIP_CONS *c) { return c->dynamic; }
inline SCIP_Bool SCIPconsIsRemovable(SCIP_CONS *c) { return c->removable; }
inline SCIP_Bool SCIPconsIsStickingAtNode(SCIP_CONS *c) { return c->stickingatnode; }

inline scip::ObjProbData *SCIPgetObjProbData(SCIP *scip) { return scip->probdata; }
inline SCIP_Real SCIPgetSolVal(SCIP *scip, SCIP_SOL *sol, SCIP_VAR *var) { return sol ? sol->val[var->index] : scip->lpsol[var->index]; }
inline SCIP_Real SCIPvarGetLbLocal(SCIP_VAR *v) { return v->lbLocal; }
inline SCIP_Real SCIPvarGetUbLocal(SCIP_VAR *v) { return v->ubLocal; }
inline int SCIPvarGetIndex(SCIP_VAR *v) { return v->index; }
inline SCIP_Bool SCIPisEQ(SCIP *, SCIP_Real a, SCIP_Real b) { return std::fabs(a - b) <= 1e-9 ? TRUE : FALSE; } /* SCIP epsilon */
inline SCIP_RETCODE SCIPaddVarLocksType(SCIP *, SCIP_VAR *var, SCIP_LOCKTYPE, int nlocksdown, int nlocksup) {
    var->nlocksdown += nlocksdown;
    var->nlocksup += nlocksup;
    return SCIP_OKAY;
}

inline SCIP_RETCODE SCIPcreateEmptyRowConshdlr(SCIP *scip, SCIP_ROW **row, SCIP_CONSHDLR *conshdlr, const char *name, SCIP_Real lhs,
                                               SCIP_Real rhs, SCIP_Bool local, SCIP_Bool modifiable, SCIP_Bool removable) {
    *row = new SCIP_Row{name, conshdlr, lhs, rhs, local, modifiable, removable, {}, 1};
    scip->nRowsCreated++;
    return SCIP_OKAY;
}
inline SCIP_RETCODE SCIPcacheRowExtensions(SCIP *, SCIP_ROW *) { return SCIP_OKAY; }
inline SCIP_RETCODE SCIPflushRowExtensions(SCIP *, SCIP_ROW *) { return SCIP_OKAY; }
inline SCIP_RETCODE SCIPaddVarToRow(SCIP *scip, SCIP_ROW *row, SCIP_VAR *var, SCIP_Real val) {
    row->terms.emplace_back(var, val);
    scip->nAddVar++;
    return SCIP_OKAY;
}
inline SCIP_Real SCIPgetCutEfficacy(SCIP *scip, SCIP_SOL *sol, SCIP_ROW *row) {
    SCIP_Real act = 0., nrm = 0.;
    for (auto &t : row->terms) {
        act += t.second * SCIPgetSolVal(scip, sol, t.first);
        nrm += t.second * t.second;
    }
    return nrm > 0. ? (act - row->rhs) / std::sqrt(nrm) : 0.;
}
inline SCIP_Bool SCIPisCutEfficacious(SCIP *scip, SCIP_SOL *sol, SCIP_ROW *row) {
    return SCIPgetCutEfficacy(scip, sol, row) > scip->minEfficacy ? TRUE : FALSE;
}
inline SCIP_RETCODE SCIPaddPoolCut(SCIP *scip, SCIP_ROW *row) {
    scip->poolCuts.push_back(*row);
    return SCIP_OKAY;
}
inline SCIP_RETCODE SCIPaddRow(SCIP *scip, SCIP_ROW *row, SCIP_Bool, SCIP_Bool *infeasible) {
    scip->lpRows.push_back(*row);
    *infeasible = (scip->nextAddRowInfeasible || scip->nAddRow == scip->infeasibleAtAddRow) ? TRUE : FALSE;
    scip->nAddRow++;
    return SCIP_OKAY;
}
inline SCIP_RETCODE SCIPreleaseRow(SCIP *scip, SCIP_ROW **row) {
    delete *row;
    *row = nullptr;
    scip->nRowsReleased++;
    return SCIP_OKAY;
}

/* callback signatures, as in scip/type_cons.h */
#define SCIP_DECL_CONSCHECK(x)                                                                                              \
    SCIP_RETCODE x(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, SCIP_SOL *sol, SCIP_Bool checkintegrality, \
                   SCIP_Bool checklprows, SCIP_Bool printreason, SCIP_Bool completely, SCIP_RESULT *result)
#define SCIP_DECL_CONSENFOLP(x) \
    SCIP_RETCODE x(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss, SCIP_Bool solinfeasible, SCIP_RESULT *result)
#define SCIP_DECL_CONSENFOPS(x)                                                                                              \
    SCIP_RETCODE x(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss, SCIP_Bool solinfeasible, \
                   SCIP_Bool objinfeasible, SCIP_RESULT *result)
#define SCIP_DECL_CONSLOCK(x) \
    SCIP_RETCODE x(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS *cons, SCIP_LOCKTYPE locktype, int nlockspos, int nlocksneg)
#define SCIP_DECL_CONSTRANS(x) SCIP_RETCODE x(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS *sourcecons, SCIP_CONS **targetcons)
#define SCIP_DECL_CONSSEPALP(x) \
    SCIP_RETCODE x(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss, SCIP_RESULT *result)
#define SCIP_DECL_CONSSEPASOL(x) \
    SCIP_RETCODE x(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss, int nusefulconss, SCIP_SOL *sol, SCIP_RESULT *result)
#define SCIP_DECL_CONSINITSOL(x) SCIP_RETCODE x(SCIP *scip, SCIP_CONSHDLR *conshdlr, SCIP_CONS **conss, int nconss)
#define SCIP_DECL_CONSHDLRCLONE(x) x(SCIP *scip, SCIP_Bool *valid) const
