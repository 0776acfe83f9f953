// Host stand-in for the reference's once-only WENOBase preprocessing
// (libWENOEXT/WENOBase/WENOBase.C:674-938). OpenFOAM is not available here, so
// the meshes come from polymesh.h; every step below cites what it restates.
// The result (HostPlan) keeps the shapes of WENOBase's accessors
// (WENOBase.H:319-397) flattened to CSR, and is the input contract of the
// per-step CUDA path (include/wenoext_b200.h: wenoext_plan_desc).
#pragma once
#include <cstdint>
#include <vector>

#include <map>

#include "geometry_weno.h"
#include "polymesh.h"

namespace wenoext {

// WENOBase::Cell (WENOBase.H:62-67)
enum : int32_t { CELL_LOCAL = -1, CELL_DELETED = -4, CELL_EMPTY = -10 };

struct SetupParams {
    int32_t polOrder = 2;
    // WENODict keys and defaults (WENOBase.C:728-735)
    double extendRatio = 2.5;
    bool bestConditioned = false;
    double maxCondition = 1e-5;
    bool checkCondition = true;
    int32_t nThreads = 0;  // 0: hardware concurrency
    // >= 0: the pseudo-inverses of all stencils are formed in one batch on this CUDA device
    // (pinv_device.h) instead of per stencil on the host; not with bestConditioned, whose
    // search over stencil sizes stays on the host
    int32_t gpuPinvDevice = -1;
};

struct HostPlan {
    int32_t polOrder = 0, nDvt = 0;
    int32_t dim[3] = {0, 0, 0};  // dimList()[cell], identical for all cells
    int32_t nCells = 0, nFaces = 0, nInternalFaces = 0;
    // mesh addressing of the rank-local mesh
    std::vector<int32_t> owner, neighbour, cellFaceStart, cellFaces;
    std::vector<Patch> patches;
    // stencilsID()[cell][stencil][j] / cellToProcMap()[cell][stencil][j], CSR
    std::vector<int64_t> cellStencilStart;   // [nCells+1] -> stencil index
    std::vector<int64_t> stencilEntryStart;  // [nStencils+1] -> entry index
    std::vector<int32_t> stencilsID;
    std::vector<int32_t> cellToProcMap;
    // LSmatrix()[cell][stencil]: index into the de-duplicated pool (matrixDB), -1 if deleted
    std::vector<int32_t> stencilMatrix;  // [nStencils]
    std::vector<int64_t> matrixStart;    // [nUnique+1] offsets into matrixData
    std::vector<int32_t> matrixCols;     // [nUnique]; rows = nDvt, row-major, unpadded
    std::vector<double> matrixData;
    std::vector<double> B;            // [nCells][nDvt][nDvt]
    std::vector<double> intBasTrans;  // [nFaces][2][(polOrder+1)^3], List3D index
    std::vector<double> refFacAr;     // [nFaces]
    // halo lists, indexed by processor id, -1 = no traffic (WENOBase.H:122-138)
    int32_t nProcs = 1, myProc = 0;
    std::vector<int32_t> sendProcList, receiveProcList, receiveHaloSize;
    std::vector<std::vector<int32_t>> sendHaloCellIDList;
    // what this rank asks of each other rank (their processor-local cell ids, in
    // halo-id order); the owner's sendHaloCellIDList is filled from it
    // (distributeStencils, WENOBase.C:334-407)
    std::vector<std::vector<int32_t>> requestHaloCellID;
    // for each boundary face of a coupled patch: flat value index of the cell across
    // (cyclic: local id; processor: nCells + haloOffset(proc) + haloId), else -1
    std::vector<int32_t> patchNbrCell;
    // geometry kept for tests / synthetic fields (not read by the hot path)
    std::vector<double> cellCentres;  // [nCells][3]
    std::vector<double> faceAreaVec;  // [nFaces][3]  Sf
    // statistics like matrixDB::info (matrixDB.C:232-273)
    int64_t nMatricesTotal = 0;
};

// Pieces of the generic path shared with the structured-box fast path (box_fast.cpp)
namespace detail {
// Result of steps 1-4 of WENOBase::WENOBase for one cell, in GLOBAL cell ids.
struct CellSetup {
    std::vector<int32_t> central;                // sorted central list before the cut
    std::vector<std::vector<int32_t>> globalID;  // [stencil][j]; {CELL_DELETED} if deleted
    std::vector<std::vector<double>> Ainv;       // [stencil] nDvt x (size-1), row-major
    std::vector<uint8_t> pending;                // [stencil] 1: matrix and pseudo-inverse still to come (finishPending)
    bool empty = false;                          // Cell::empty
    List3D volI;
    Mat3 JInv;
    Vec3 refPoint;
    double refDet = 0;
};
CellSetup computeCell(const PolyMesh& gm, int g, int nDvt, const int dim[3], const SetupParams& prm);
int computeNDvt(const PolyMesh& m, int polOrder, int dim[3]);
// Deferred LS matrices and pseudo-inverses (SetupParams::gpuPinvDevice): one device batch over all pending
// stencils of `cells`, then the per-stencil decisions of calcMatrix / LSMatrixCheck.
void finishPending(std::vector<CellSetup>& cells, const PolyMesh& gm, const std::vector<int32_t>& cellGlobal,
                   int nDvt, const int dim[3], const SetupParams& prm);
// matrixDB stand-in (libWENOEXT/WENOBase/matrixDB.C:75-199): tolerance-based
// de-duplication, same key, same search window, same comparison.
struct MatrixPool {
    std::multimap<int32_t, int32_t> db;
    HostPlan* plan = nullptr;
    int nDvt = 0;
    int32_t insertNew(int32_t key, const std::vector<double>& A);
    int32_t similar(const std::vector<double>& A);
};
}  // namespace detail

// Generic path: any PolyMesh, any decomposition. `cellProc` may be empty (serial).
HostPlan buildPlanGeneric(const PolyMesh& globalMesh, const std::vector<int32_t>& cellProc,
                          int32_t nProcs, int32_t myProc, const SetupParams& prm);

// After every rank has built its plan: hand rank `from`'s request list to the
// owner `plan` (what Pstream does in distributeStencils).
void setSendHalo(HostPlan& plan, int32_t from, const int32_t* ids, int64_t n);

}  // namespace wenoext
