#include "box_fast.h"

namespace wenoext {
HostPlan buildPlanBoxFast(const BoxSpec& spec, const int32_t nProc[3], int32_t myProc,
                          const SetupParams& prm) {
    // TODO(class cache): generic path for now
    PolyMesh gm = makeBoxMesh(spec);
    const int nProcs = nProc[0] * nProc[1] * nProc[2];
    std::vector<int32_t> cellProc;
    if (nProcs > 1) cellProc = decomposeBox(spec, nProc);
    return buildPlanGeneric(gm, cellProc, nProcs, myProc, prm);
}
}  // namespace wenoext
