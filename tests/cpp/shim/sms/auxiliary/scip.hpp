// HEADLESS-TEST STAND-IN for sms/auxiliary/scip.hpp: the one helper the handlers call.
// scipVarIsFixedLocal restates src/auxiliary/scip.cpp:70-72 of the reference.
#pragma once
#include "scip/scip.h"
inline bool scipVarIsFixedLocal(SCIP *scip, SCIP_VAR *var) { return SCIPisEQ(scip, SCIPvarGetLbLocal(var), SCIPvarGetUbLocal(var)); }
