// extern "C" wrappers of the host setup stand-in (see include/wenoext_b200.h).
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/wenoext_b200.h"
#include "plan_internal.h"
#include "setup/box_fast.h"
#include "setup/weno_setup.h"

using namespace wenoext;

struct wenoext_setup {
    HostPlan hp;
    std::vector<wenoext_patch> patches;
    std::vector<int64_t> sendHaloStart;
    std::vector<int32_t> sendHaloCellID;
};

static BoxSpec toBoxSpec(const wenoext_box_spec* b) {
    BoxSpec s;
    for (int i = 0; i < 3; ++i) {
        s.n[i] = b->n[i];
        s.lo[i] = b->lo[i];
        s.hi[i] = b->hi[i];
    }
    for (int i = 0; i < 6; ++i) s.sideType[i] = b->sideType[i];
    s.tetSplit = b->tetSplit;
    return s;
}

extern "C" int wenoext_setup_box(const wenoext_box_spec* box, const wenoext_setup_params* prm,
                                 int32_t myProc, wenoext_setup** out) {
    if (!box || !prm || !out) {
        setError("setup_box: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    *out = nullptr;
    try {
        BoxSpec spec = toBoxSpec(box);
        int32_t np[3] = {std::max(1, box->nProc[0]), std::max(1, box->nProc[1]),
                         std::max(1, box->nProc[2])};
        const int nProcs = np[0] * np[1] * np[2];
        if (myProc < 0 || myProc >= nProcs) throw std::runtime_error("myProc out of range");
        SetupParams sp;
        sp.polOrder = prm->polOrder;
        sp.extendRatio = prm->extendRatio > 0 ? prm->extendRatio : 2.5;
        sp.bestConditioned = prm->bestConditioned != 0;
        sp.maxCondition = prm->maxCondition > 0 ? prm->maxCondition : 1e-5;
        sp.checkCondition = prm->checkCondition != 0;
        sp.nThreads = prm->nThreads;
        auto s = std::make_unique<wenoext_setup>();
        if (prm->useClassCache && !spec.tetSplit) {
            s->hp = buildPlanBoxFast(spec, np, myProc, sp);
        } else {
            PolyMesh gm = makeBoxMesh(spec);
            std::vector<int32_t> cellProc;
            if (nProcs > 1) cellProc = decomposeBox(spec, np);
            s->hp = buildPlanGeneric(gm, cellProc, nProcs, myProc, sp);
        }
        *out = s.release();
    } catch (const std::exception& e) {
        setError(std::string("setup_box: ") + e.what());
        return WENOEXT_ERR_SETUP;
    }
    return 0;
}

extern "C" int wenoext_setup_destroy(wenoext_setup* s) {
    delete s;
    return 0;
}

extern "C" int wenoext_setup_desc(wenoext_setup* s, wenoext_plan_desc* d) {
    if (!s || !d) {
        setError("setup_desc: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    HostPlan& h = s->hp;
    memset(d, 0, sizeof(*d));
    d->polOrder = h.polOrder;
    d->nDvt = h.nDvt;
    for (int i = 0; i < 3; ++i) d->dim[i] = h.dim[i];
    d->nCells = h.nCells;
    d->nFaces = h.nFaces;
    d->nInternalFaces = h.nInternalFaces;
    d->owner = h.owner.data();
    d->neighbour = h.neighbour.data();
    d->cellFaceStart = h.cellFaceStart.data();
    d->cellFaces = h.cellFaces.data();
    s->patches.clear();
    for (const Patch& p : h.patches) s->patches.push_back({p.type, p.start, p.size, p.nbr});
    d->nPatches = (int32_t)s->patches.size();
    d->patches = s->patches.data();
    d->nStencils = (int64_t)h.stencilMatrix.size();
    d->cellStencilStart = h.cellStencilStart.data();
    d->stencilEntryStart = h.stencilEntryStart.data();
    d->stencilsID = h.stencilsID.data();
    d->cellToProcMap = h.cellToProcMap.empty() ? nullptr : h.cellToProcMap.data();
    d->stencilMatrix = h.stencilMatrix.data();
    d->nUniqueMatrices = (int32_t)h.matrixCols.size();
    d->matrixStart = h.matrixStart.data();
    d->matrixCols = h.matrixCols.data();
    d->matrixData = h.matrixData.data();
    d->B = h.B.data();
    d->intBasTrans = h.intBasTrans.data();
    d->refFacAr = h.refFacAr.data();
    d->nProcs = h.nProcs;
    d->myProc = h.myProc;
    d->sendProcList = h.sendProcList.data();
    d->receiveProcList = h.receiveProcList.data();
    d->receiveHaloSize = h.receiveHaloSize.data();
    s->sendHaloStart.assign(h.nProcs + 1, 0);
    s->sendHaloCellID.clear();
    for (int p = 0; p < h.nProcs; ++p) {
        const auto& l = h.sendHaloCellIDList[p];
        s->sendHaloCellID.insert(s->sendHaloCellID.end(), l.begin(), l.end());
        s->sendHaloStart[p + 1] = (int64_t)s->sendHaloCellID.size();
    }
    d->sendHaloStart = s->sendHaloStart.data();
    d->sendHaloCellID = s->sendHaloCellID.data();
    d->patchNbrCell = h.patchNbrCell.empty() ? nullptr : h.patchNbrCell.data();
    return 0;
}

extern "C" int wenoext_setup_halo_request(wenoext_setup* s, int32_t proc, const int32_t** ids,
                                          int64_t* n) {
    if (!s || !ids || !n || proc < 0 || proc >= s->hp.nProcs) {
        setError("setup_halo_request: bad argument");
        return WENOEXT_ERR_INVALID;
    }
    *ids = s->hp.requestHaloCellID[proc].data();
    *n = (int64_t)s->hp.requestHaloCellID[proc].size();
    return 0;
}

extern "C" int wenoext_setup_set_send_halo(wenoext_setup* s, int32_t proc, const int32_t* ids,
                                           int64_t n) {
    if (!s || proc < 0 || proc >= s->hp.nProcs || (n > 0 && !ids)) {
        setError("setup_set_send_halo: bad argument");
        return WENOEXT_ERR_INVALID;
    }
    for (int64_t i = 0; i < n; ++i)
        if (ids[i] < 0 || ids[i] >= s->hp.nCells) {
            setError("setup_set_send_halo: cell id out of range");
            return WENOEXT_ERR_INVALID;
        }
    setSendHalo(s->hp, proc, ids, n);
    return 0;
}

extern "C" int wenoext_setup_geometry(wenoext_setup* s, const double** cellCentres,
                                      const double** faceAreaVec) {
    if (!s) return WENOEXT_ERR_INVALID;
    if (cellCentres) *cellCentres = s->hp.cellCentres.data();
    if (faceAreaVec) *faceAreaVec = s->hp.faceAreaVec.data();
    return 0;
}
