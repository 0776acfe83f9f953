// OpenFOAM-side flattening of the WENOBase singleton into a wenoext_plan_desc.
// Compiles against OpenFOAM (v5-8 / v1912-v2012) and the WENOEXT headers: add to libWENOEXT/ and link
// with -lwenoext_b200. In this repository it is compiled and run against the OpenFOAM
// stand-in oracle/of_stub (`make -C shim check`, tests/test_shim.py).
#include <map>
#include <vector>

#include "WENOBase.H"
#include "wenoB200Registry.H"
#include "wenoext_b200.h"
#include "wenoext_setup.h"
#include "processorFvPatch.H"
#include "cyclicFvPatch.H"
#include "emptyFvPatch.H"

namespace Foam { namespace wenoB200 {

struct Flat   // owns the temporaries the descriptor points to
{
    std::vector<int32_t> owner, neighbour, cellFaceStart, cellFaces;
    std::vector<wenoext_patch> patches;
    std::vector<int64_t> cellStencilStart, stencilEntryStart, matrixStart, sendHaloStart;
    std::vector<int32_t> stencilsID, cellToProcMap, stencilMatrix, matrixCols;
    std::vector<int32_t> sendProc, recvProc, recvSize, sendHaloCellID;
    std::vector<double>  matrixData, B, intBasTrans, refFacAr;
    wenoext_plan_desc d{};
};

wenoext_plan* planFromBase(const fvMesh& mesh, const WENOBase& wb, const label polOrder, const int device)
{
    Flat f;
    const label nCells = mesh.nCells(), nFaces = mesh.nFaces(), nInt = mesh.nInternalFaces();
    const label K = wb.degreesOfFreedom();                       // WENOBase.H:319-397 accessors
    // fvMesh::owner() is the internal-face slice; the ABI wants the owner of every face
    f.owner.assign(mesh.faceOwner().begin(), mesh.faceOwner().end());
    f.neighbour.assign(mesh.neighbour().begin(), mesh.neighbour().end());
    f.cellFaceStart.push_back(0);
    forAll(mesh.cells(), c)
    {
        for (label fc : mesh.cells()[c]) f.cellFaces.push_back(fc);
        f.cellFaceStart.push_back(f.cellFaces.size());
    }
    forAll(mesh.boundary(), p)                                   // WENOUpwindFit.C:433,:478,:493
    {
        const fvPatch& fp = mesh.boundary()[p];
        wenoext_patch q{WENOEXT_PATCH_GENERIC, int32_t(fp.start()), int32_t(fp.size()), -1};
        if (isA<processorFvPatch>(fp))
            { q.type = WENOEXT_PATCH_PROCESSOR; q.nbr = refCast<const processorFvPatch>(fp).neighbProcNo(); }
        else if (isA<cyclicFvPatch>(fp))
        {
            q.type = WENOEXT_PATCH_CYCLIC;
            #ifdef FOAM_NEW_COUPLED_PATCHES
            q.nbr = refCast<const cyclicFvPatch>(fp).nbrPatchID();
            #else
            q.nbr = refCast<const cyclicFvPatch>(fp).neighbPatchID();
            #endif
        }
        else if (isA<emptyFvPatch>(fp)) q.type = WENOEXT_PATCH_EMPTY;
        f.patches.push_back(q);
    }

    // stencils, processor map and matrices in LIST ORDER (deleted slots kept: the
    // "first surviving stencil gets dm" rule of WENOCoeff.C:198 depends on it)
    std::map<const void*, int32_t> poolId;                        // matrixDB representatives
    f.cellStencilStart.push_back(0); f.stencilEntryStart.push_back(0); f.matrixStart.push_back(0);
    forAll(wb.stencilsID(), c)
    {
        forAll(wb.stencilsID()[c], s)
        {
            const labelList& ids = wb.stencilsID()[c][s];
            for (label j = 0; j < ids.size(); ++j)
            {
                f.stencilsID.push_back(ids[j]);
                f.cellToProcMap.push_back(wb.cellToProcMap()[c][s][j]);
            }
            f.stencilEntryStart.push_back(f.stencilsID.size());
            int32_t m = -1;
            if (ids[0] != int(WENOBase::Cell::deleted) && ids[0] != int(WENOBase::Cell::empty))
            {
                const auto& A = wb.LSmatrix()[c][s]();            // blaze row-major, padded rows
                auto it = poolId.find(&A);
                if (it == poolId.end())
                {
                    m = poolId.size(); poolId[&A] = m;
                    for (size_t r = 0; r < A.rows(); ++r)         // strip Blaze's row padding
                        for (size_t j = 0; j < A.columns(); ++j) f.matrixData.push_back(A(r, j));
                    f.matrixCols.push_back(A.columns());
                    f.matrixStart.push_back(f.matrixData.size());
                }
                else m = it->second;
            }
            f.stencilMatrix.push_back(m);
        }
        f.cellStencilStart.push_back(f.stencilMatrix.size());
        const auto& Bc = wb.B()[c];
        for (label r = 0; r < K; ++r) for (label q = 0; q < K; ++q) f.B.push_back(Bc(r, q));
    }
    const label cube = (polOrder + 1)*(polOrder + 1)*(polOrder + 1);
    f.intBasTrans.resize(size_t(nFaces)*2*cube);
    forAll(wb.intBasTrans(), fc)                                  // List3D.H:238 index order
        for (int side = 0; side < 2; ++side)
        {
            const auto& L = side == 0 ? wb.intBasTrans()[fc].first() : wb.intBasTrans()[fc].second();
            for (label n = 0; n <= polOrder; ++n) for (label m = 0; m <= polOrder; ++m)
                for (label l = 0; l <= polOrder; ++l)
                    f.intBasTrans[(size_t(fc)*2 + side)*cube + (n*(polOrder+1) + m)*(polOrder+1) + l] = L(n, m, l);
        }
    f.refFacAr.assign(wb.refFacAr().begin(), wb.refFacAr().end());

    // halo lists, indexed by processor id (WENOBase.H:122-138)
    f.sendProc.assign(wb.sendProcList().begin(), wb.sendProcList().end());
    f.recvProc.assign(wb.receiveProcList().begin(), wb.receiveProcList().end());
    f.recvSize.assign(wb.receiveHaloSize().begin(), wb.receiveHaloSize().end());
    f.sendHaloStart.push_back(0);
    forAll(wb.sendHaloCellIDList(), p)
    {
        for (label id : wb.sendHaloCellIDList()[p]) f.sendHaloCellID.push_back(id);
        f.sendHaloStart.push_back(f.sendHaloCellID.size());
    }

    wenoext_plan_desc& d = f.d;
    d.polOrder = polOrder; d.nDvt = K;
    for (int i = 0; i < 3; ++i) d.dim[i] = wb.dimList()[0][i];    // same for all cells, WENOBase.C:1147-1154
    d.nCells = nCells; d.nFaces = nFaces; d.nInternalFaces = nInt;
    d.owner = f.owner.data(); d.neighbour = f.neighbour.data();
    d.cellFaceStart = f.cellFaceStart.data(); d.cellFaces = f.cellFaces.data();
    d.nPatches = f.patches.size(); d.patches = f.patches.data();
    d.nStencils = f.stencilMatrix.size();
    d.cellStencilStart = f.cellStencilStart.data(); d.stencilEntryStart = f.stencilEntryStart.data();
    d.stencilsID = f.stencilsID.data(); d.cellToProcMap = f.cellToProcMap.data();
    d.stencilMatrix = f.stencilMatrix.data(); d.nUniqueMatrices = f.matrixCols.size();
    d.matrixStart = f.matrixStart.data(); d.matrixCols = f.matrixCols.data(); d.matrixData = f.matrixData.data();
    d.B = f.B.data(); d.intBasTrans = f.intBasTrans.data(); d.refFacAr = f.refFacAr.data();
    d.nProcs = Pstream::nProcs(); d.myProc = Pstream::myProcNo();
    d.sendProcList = f.sendProc.data(); d.receiveProcList = f.recvProc.data();
    d.receiveHaloSize = f.recvSize.data();
    d.sendHaloStart = f.sendHaloStart.data(); d.sendHaloCellID = f.sendHaloCellID.data();
    d.patchNbrCell = nullptr;                // the shim passes patchNeighbourField() per call instead

    wenoext_plan* plan = nullptr;
    if (wenoext_plan_create(&d, device, &plan) != WENOEXT_OK)
        FatalErrorInFunction << "wenoext_b200: " << wenoext_last_error() << exit(FatalError);
    return plan;
}

// Second way to the same plan: parse the cache the stock library wrote to
// constant/WENOBase<order>/ (WENOBase::writeList, WENOBase.C:1265-1312) instead of walking the
// accessors; only the mesh addressing comes from the fvMesh. See INTEGRATION.md section 4.
wenoext_plan* planFromCache(const fvMesh& mesh, const label polOrder, const int device)
{
    std::vector<wenoext_patch> patches;
    forAll(mesh.boundary(), patchI)
    {
        const fvPatch& p = mesh.boundary()[patchI];
        wenoext_patch q{WENOEXT_PATCH_GENERIC, int32_t(p.start()), int32_t(p.size()), -1};
        if (isA<processorFvPatch>(p))
        {
            q.type = WENOEXT_PATCH_PROCESSOR;
            q.nbr = refCast<const processorFvPatch>(p).neighbProcNo();
        }
        else if (isA<cyclicFvPatch>(p))
        {
            q.type = WENOEXT_PATCH_CYCLIC;
            q.nbr = refCast<const cyclicFvPatch>(p).neighbPatchID();
        }
        else if (isA<emptyFvPatch>(p)) q.type = WENOEXT_PATCH_EMPTY;
        patches.push_back(q);
    }
    // labels are 32-bit here (WM_LABEL_SIZE=32); with 64-bit labels copy owner/neighbour to int32 first
    wenoext_mesh_desc m{};
    m.nCells = mesh.nCells(); m.nFaces = mesh.nFaces(); m.nInternalFaces = mesh.nInternalFaces();
    m.owner = mesh.faceOwner().cdata(); m.neighbour = mesh.neighbour().cdata();
    m.nPatches = patches.size(); m.patches = patches.data();

    const fileName dir = mesh.time().path()/"constant"/("WENOBase" + Foam::name(polOrder));
    wenoext_setup* s = nullptr;
    if (wenoext_cache_read(dir.c_str(), &m, Pstream::myProcNo(), 0, &s) != WENOEXT_OK)
        FatalErrorInFunction << "wenoext_b200: " << wenoext_last_error() << exit(FatalError);
    wenoext_plan_desc d;
    wenoext_setup_desc(s, &d);
    wenoext_plan* plan = nullptr;
    const int rc = wenoext_plan_create(&d, device, &plan);
    wenoext_setup_destroy(s);
    if (rc != WENOEXT_OK)
        FatalErrorInFunction << "wenoext_b200: " << wenoext_last_error() << exit(FatalError);
    return plan;
}
}}
