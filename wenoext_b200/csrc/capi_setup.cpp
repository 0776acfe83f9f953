// extern "C" wrappers of the host-side descriptor producers (include/wenoext_setup.h):
// libwenoext_setup.so, plain C++ without CUDA. Kept apart from the product library so that the
// CPU-only users of a descriptor (the reference build, the oracle) never map the GPU library.
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include <dlfcn.h>

#include "../../include/wenoext_b200.h"
#include "../../include/wenoext_setup.h"
#include "setup/pinv_device.h"
#include "setup/box_fast.h"
#include "setup/weno_setup.h"
#include "setup/wenobase_cache.h"

using namespace wenoext;

namespace {
thread_local std::string g_setupError;
void setError(const std::string& s) { g_setupError = s; }
}  // namespace

extern "C" const char* wenoext_setup_last_error(void) { return g_setupError.c_str(); }

// The optional device batch of the pseudo-inverses (wenoext_setup_params::gpuPinv) lives in the product
// library; it is looked up at run time in libwenoext_b200.so next to this library, so that nothing here
// links against CUDA.
static void* productLibrary() {
    static void* handle = nullptr;
    if (!handle) {
        Dl_info info{};
        std::string dir = ".";
        if (dladdr((const void*)&wenoext_setup_last_error, &info) && info.dli_fname) {
            const std::string self(info.dli_fname);
            const size_t slash = self.rfind('/');
            if (slash != std::string::npos) dir = self.substr(0, slash);
        }
        handle = dlopen((dir + "/libwenoext_b200.so").c_str(), RTLD_NOW | RTLD_LOCAL);
        if (!handle) handle = dlopen("libwenoext_b200.so", RTLD_NOW | RTLD_LOCAL);
    }
    return handle;
}

int wenoext::pinvBatchedDevice(int device, int32_t nMat, int32_t n, const int32_t* rows, const int64_t* start,
                               const double* A, double minCondition, double* Ainv, int32_t* nZeros,
                               double* condition, int32_t* converged, std::string& err) {
    typedef int (*Fn)(int32_t, int32_t, const int32_t*, const int64_t*, const double*, double, double*, int32_t*,
                      double*, int32_t*, int);
    typedef const char* (*ErrFn)(void);
    void* handle = productLibrary();
    Fn fn = handle ? (Fn)dlsym(handle, "wenoext_pinv_batched") : nullptr;
    if (!fn) {
        err = "gpuPinv: libwenoext_b200.so (wenoext_pinv_batched) not found next to libwenoext_setup.so";
        return WENOEXT_ERR_SETUP;
    }
    const int rc = fn(nMat, n, rows, start, A, minCondition, Ainv, nZeros, condition, converged, device);
    if (rc) {
        ErrFn ef = (ErrFn)dlsym(handle, "wenoext_last_error");
        err = ef ? ef() : "wenoext_pinv_batched failed";
    }
    return rc;
}

// wenoext_ls_pinv_batched of the product library (csrc/ls_rows.cu), found the same way
int wenoext::lsPinvBatchedDevice(int device, const LsBatchHost& h, double minCondition, double* Ainv,
                                 int32_t* nZeros, double* condition, int32_t* converged, std::string& err) {
    typedef int (*Fn)(const wenoext_ls_batch*, double, double*, int32_t*, double*, int32_t*, double*, double*, int);
    typedef const char* (*ErrFn)(void);
    void* handle = productLibrary();
    Fn fn = handle ? (Fn)dlsym(handle, "wenoext_ls_pinv_batched") : nullptr;
    if (!fn) {
        err = "gpuPinv: libwenoext_b200.so (wenoext_ls_pinv_batched) not found next to libwenoext_setup.so";
        return WENOEXT_ERR_SETUP;
    }
    wenoext_ls_batch b{};
    b.polOrder = h.polOrder;
    b.nDvt = h.nDvt;
    for (int k = 0; k < 3; ++k) b.dim[k] = h.dim[k];
    b.nPoints = (int32_t)(h.points.size() / 3);
    b.nSrcCells = (int32_t)h.cellVolumes.size();
    b.points = h.points.data();
    b.triStart = h.triStart.data();
    b.triPoints = h.triPoints.data();
    b.cellCentres = h.cellCentres.data();
    b.cellVolumes = h.cellVolumes.data();
    b.nFrames = (int32_t)h.frameCell.size();
    b.JInv = h.JInv.data();
    b.refPoint = h.refPoint.data();
    b.refDet = h.refDet.data();
    b.frameCell = h.frameCell.data();
    b.volI = h.volI.data();
    b.nPairs = (int64_t)h.pairCell.size();
    b.pairFrame = h.pairFrame.data();
    b.pairCell = h.pairCell.data();
    b.nMat = (int32_t)h.rows.size();
    b.rows = h.rows.data();
    b.start = h.start.data();
    b.rowPair = h.rowPair.data();
    const int rc = fn(&b, minCondition, Ainv, nZeros, condition, converged, nullptr, nullptr, device);
    if (rc) {
        ErrFn ef = (ErrFn)dlsym(handle, "wenoext_last_error");
        err = ef ? ef() : "wenoext_ls_pinv_batched failed";
    }
    return rc;
}

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
        sp.gpuPinvDevice = prm->gpuPinv ? prm->device : -1;
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

// ---- constant/WENOBase<order>/ cache ----------------------

extern "C" int wenoext_cache_write(wenoext_setup* s, const char* dir, int labelBytes, int simdDoubles) {
    if (!s || !dir) {
        setError("cache_write: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    try {
        CacheFormat fmt;
        fmt.labelBytes = labelBytes ? labelBytes : 4;
        fmt.simdDoubles = simdDoubles ? simdDoubles : 8;
        writeWenoBaseCache(s->hp, dir, fmt);
    } catch (const std::exception& e) {
        setError(std::string("cache_write: ") + e.what());
        return WENOEXT_ERR_SETUP;
    }
    return 0;
}

extern "C" int wenoext_cache_read(const char* dir, const wenoext_mesh_desc* m, int32_t myProc,
                                  int labelBytes, wenoext_setup** out) {
    if (!dir || !m || !out) {
        setError("cache_read: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    *out = nullptr;
    try {
        auto s = std::make_unique<wenoext_setup>();
        HostPlan& h = s->hp;
        CacheFormat fmt;
        fmt.labelBytes = labelBytes;
        readWenoBaseCache(dir, myProc, h, fmt);
        if (m->nCells != h.nCells)
            throw std::runtime_error("the cache holds " + std::to_string(h.nCells) + " cells, the mesh " +
                                     std::to_string(m->nCells));
        if (m->nFaces != h.nFaces)
            throw std::runtime_error("the cache holds " + std::to_string(h.nFaces) + " faces, the mesh " +
                                     std::to_string(m->nFaces));
        if (m->nInternalFaces < 0 || m->nInternalFaces > m->nFaces || !m->owner ||
            (m->nInternalFaces > 0 && !m->neighbour))
            throw std::runtime_error("bad mesh addressing");
        if (myProc < 0 || myProc >= std::max(1, h.nProcs)) throw std::runtime_error("myProc out of range");
        h.nInternalFaces = m->nInternalFaces;
        h.owner.assign(m->owner, m->owner + m->nFaces);
        h.neighbour.assign(m->neighbour, m->neighbour + m->nInternalFaces);
        for (int f = 0; f < m->nFaces; ++f)
            if (h.owner[f] < 0 || h.owner[f] >= h.nCells) throw std::runtime_error("owner out of range");
        for (int f = 0; f < m->nInternalFaces; ++f)
            if (h.neighbour[f] < 0 || h.neighbour[f] >= h.nCells) throw std::runtime_error("neighbour out of range");
        if (m->cellFaceStart && m->cellFaces) {
            h.cellFaceStart.assign(m->cellFaceStart, m->cellFaceStart + h.nCells + 1);
            h.cellFaces.assign(m->cellFaces, m->cellFaces + h.cellFaceStart[h.nCells]);
        } else {
            // primitiveMesh::calcCells (OpenFOAM): owned faces first, then neighboured ones
            h.cellFaceStart.assign((size_t)h.nCells + 1, 0);
            for (int f = 0; f < m->nFaces; ++f) ++h.cellFaceStart[h.owner[f] + 1];
            for (int f = 0; f < m->nInternalFaces; ++f) ++h.cellFaceStart[h.neighbour[f] + 1];
            for (int c = 0; c < h.nCells; ++c) h.cellFaceStart[c + 1] += h.cellFaceStart[c];
            h.cellFaces.assign((size_t)h.cellFaceStart[h.nCells], 0);
            std::vector<int32_t> fill(h.cellFaceStart.begin(), h.cellFaceStart.end() - 1);
            for (int f = 0; f < m->nFaces; ++f) h.cellFaces[fill[h.owner[f]]++] = f;
            for (int f = 0; f < m->nInternalFaces; ++f) h.cellFaces[fill[h.neighbour[f]]++] = f;
        }
        h.patches.clear();
        for (int i = 0; i < m->nPatches; ++i) {
            Patch p;
            p.type = m->patches[i].type;
            p.start = m->patches[i].start;
            p.size = m->patches[i].size;
            p.nbr = m->patches[i].nbr;
            if (p.size < 0 || p.start < m->nInternalFaces || p.start + p.size > m->nFaces)
                throw std::runtime_error("patch " + std::to_string(i) + " outside the boundary faces");
            h.patches.push_back(p);
        }
        if (m->patchNbrCell) h.patchNbrCell.assign(m->patchNbrCell, m->patchNbrCell + (m->nFaces - m->nInternalFaces));
        h.requestHaloCellID.assign((size_t)std::max(1, h.nProcs), {});
        if (h.nProcs <= 1) {  // a serial cache has one-entry processor lists
            h.nProcs = 1;
            h.sendProcList.assign(1, -1);
            h.receiveProcList.assign(1, -1);
            h.receiveHaloSize.assign(1, 0);
            h.sendHaloCellIDList.assign(1, {});
        }
        *out = s.release();
    } catch (const std::exception& e) {
        setError(std::string("cache_read: ") + e.what());
        return WENOEXT_ERR_SETUP;
    }
    return 0;
}

extern "C" int wenoext_setup_mesh(wenoext_setup* s, wenoext_mesh_desc* m) {
    if (!s || !m) {
        setError("setup_mesh: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    HostPlan& h = s->hp;
    memset(m, 0, sizeof(*m));
    m->nCells = h.nCells;
    m->nFaces = h.nFaces;
    m->nInternalFaces = h.nInternalFaces;
    m->owner = h.owner.data();
    m->neighbour = h.neighbour.data();
    m->cellFaceStart = h.cellFaceStart.data();
    m->cellFaces = h.cellFaces.data();
    s->patches.clear();
    for (const Patch& p : h.patches) s->patches.push_back({p.type, p.start, p.size, p.nbr});
    m->nPatches = (int32_t)s->patches.size();
    m->patches = s->patches.data();
    m->patchNbrCell = h.patchNbrCell.empty() ? nullptr : h.patchNbrCell.data();
    return 0;
}

extern "C" int wenoext_case_read(const char* caseDir, int32_t polOrder, int32_t myProc, int labelBytes,
                                 wenoext_setup** out) {
    if (!caseDir || !out) {
        setError("case_read: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    *out = nullptr;
    try {
        auto s = std::make_unique<wenoext_setup>();
        HostPlan& h = s->hp;
        CacheFormat fmt;
        fmt.labelBytes = labelBytes;
        const std::string dir(caseDir);
        readWenoBaseCache(dir + "/constant/WENOBase" + std::to_string(polOrder), myProc, h, fmt);
        if (h.polOrder != polOrder)
            throw std::runtime_error("the cache was written for order " + std::to_string(h.polOrder));
        const int32_t nCellsCache = h.nCells, nFacesCache = h.nFaces;
        readPolyMeshAddressing(dir + "/constant/polyMesh", h);
        if (h.nCells != nCellsCache || h.nFaces != nFacesCache)
            throw std::runtime_error("cache (" + std::to_string(nCellsCache) + " cells, " + std::to_string(nFacesCache) +
                                     " faces) and polyMesh (" + std::to_string(h.nCells) + ", " +
                                     std::to_string(h.nFaces) + ") do not belong together");
        if (myProc < 0 || myProc >= std::max(1, h.nProcs)) throw std::runtime_error("myProc out of range");
        h.requestHaloCellID.assign((size_t)std::max(1, h.nProcs), {});
        if (h.nProcs <= 1) {
            h.nProcs = 1;
            h.sendProcList.assign(1, -1);
            h.receiveProcList.assign(1, -1);
            h.receiveHaloSize.assign(1, 0);
            h.sendHaloCellIDList.assign(1, {});
        }
        *out = s.release();
    } catch (const std::exception& e) {
        setError(std::string("case_read: ") + e.what());
        return WENOEXT_ERR_SETUP;
    }
    return 0;
}
