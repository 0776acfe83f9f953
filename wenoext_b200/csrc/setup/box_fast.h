// Fast setup for structured hex boxes: translation-invariant cell classes are
// computed once with the generic algorithm (weno_setup.cpp) on a small
// representative mesh and instantiated for every cell of the class. Produces the
// same HostPlan shape as buildPlanGeneric.
#pragma once
#include "weno_setup.h"

namespace wenoext {
HostPlan buildPlanBoxFast(const BoxSpec& spec, const int32_t nProc[3], int32_t myProc,
                          const SetupParams& prm);
}
