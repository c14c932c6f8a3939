/* HEADLESS-TEST STAND-IN: forwards to the recording fake in scip/scip.h */
#pragma once
#include "scip/scip.h"
