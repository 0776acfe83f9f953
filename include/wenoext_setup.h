/* wenoext_setup — host-side producers of a wenoext_plan_desc (libwenoext_setup.so, plain C++, no CUDA):
 *   - a stand-in for the reference's WENOBase constructor on generated box meshes (OpenFOAM is not
 *     available to run the reference's own setup),
 *   - reader / writer of the reference's on-disk cache constant/WENOBase<order>/ and reader of
 *     constant/polyMesh addressing, so that a case preprocessed by the stock library can drive the
 *     GPU path without OpenFOAM.
 * Kept out of the product library libwenoext_b200.so (include/wenoext_b200.h): nothing here runs per
 * time step. All citations are relative to the reference tree (WENO-OF/WENOEXT @ 78907ff).
 * Every function returns 0 on success; wenoext_setup_last_error() gives the message.
 */
#ifndef WENOEXT_SETUP_H
#define WENOEXT_SETUP_H

#include "wenoext_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

const char* wenoext_setup_last_error(void);

/* ---------------------------------------------------------------------------
 * Host setup stand-in (WENOBase::WENOBase, WENOBase.C:674-938) on generated box
 * meshes, because OpenFOAM is not available to run the reference's own setup.
 * Produces a wenoext_plan_desc. Runs once, on the host.
 * ------------------------------------------------------------------------- */
typedef struct wenoext_box_spec {
    int32_t n[3];
    double lo[3], hi[3];
    int32_t sideType[6]; /* xlo,xhi,ylo,yhi,zlo,zhi: WENOEXT_PATCH_GENERIC/EMPTY/CYCLIC */
    int32_t tetSplit;    /* 0 hexes, 6 = every hex split into 6 tets */
    int32_t nProc[3];    /* `decomposePar simple` block decomposition */
} wenoext_box_spec;

typedef struct wenoext_setup_params {
    int32_t polOrder;
    double extendRatio;      /* WENODict, default 2.5 */
    int32_t bestConditioned; /* default 0 */
    double maxCondition;     /* default 1e-5 */
    int32_t checkCondition;  /* default 1 */
    int32_t nThreads;        /* 0 = all */
    int32_t useClassCache;   /* 1: translation-invariant cell classes computed once */
    int32_t gpuPinv;         /* 1: pseudo-inverses of all stencils in one batch on the GPU
                                (wenoext_pinv_batched) instead of per stencil on the host;
                                generic path without bestConditioned only */
    int32_t device;          /* CUDA device of that batch */
} wenoext_setup_params;

typedef struct wenoext_setup wenoext_setup;

int wenoext_setup_box(const wenoext_box_spec* box, const wenoext_setup_params* prm,
                      int32_t myProc, wenoext_setup** out);
int wenoext_setup_destroy(wenoext_setup* s);
/* Pointers in `desc` stay valid until wenoext_setup_destroy. */
int wenoext_setup_desc(wenoext_setup* s, wenoext_plan_desc* desc);
/* distributeStencils (WENOBase.C:334-407): what this rank requests from `proc`
 * (processor-local cell ids there), and the owner-side setter. */
int wenoext_setup_halo_request(wenoext_setup* s, int32_t proc, const int32_t** ids, int64_t* n);
int wenoext_setup_set_send_halo(wenoext_setup* s, int32_t proc, const int32_t* ids, int64_t n);
/* mesh geometry for building synthetic fields: C() [nCells][3], Sf() [nFaces][3] */
int wenoext_setup_geometry(wenoext_setup* s, const double** cellCentres,
                           const double** faceAreaVec);

/* ---------------------------------------------------------------------------
 * The reference's on-disk cache constant/WENOBase<order>/ (SURVEY.md 8f-2): the
 * eleven files of WENOBase::writeList / readList (WENOBase.C:1196-1312; matrixDB.C:
 * 276-415; BlazeIO.C:30-134; List3D.H:295-351) in OpenFOAM's BINARY stream framing.
 * A cache written by the reference (one per processor directory) can drive the GPU
 * path without re-running the preprocessing, and a cache written here can be read
 * by the reference's readList.
 * ------------------------------------------------------------------------- */
/* labelBytes 4 or 8 (WM_LABEL_SIZE/8); simdDoubles: Blaze's row padding of the host
 * that will read the files (8 AVX-512, 4 AVX/AVX2, 2 SSE2, 1 none; 0 = 8). */
int wenoext_cache_write(wenoext_setup* s, const char* dir, int labelBytes, int simdDoubles);

/* The mesh addressing the cache does not hold; the reference takes it from the
 * fvMesh (mesh.owner(), mesh.neighbour(), mesh.cells(), mesh.boundary()). */
typedef struct wenoext_mesh_desc {
    int32_t nCells, nFaces, nInternalFaces;
    const int32_t* owner;         /* [nFaces] */
    const int32_t* neighbour;     /* [nInternalFaces] */
    const int32_t* cellFaceStart; /* mesh.cells() CSR [nCells+1]; NULL: derived from owner/
                                     neighbour as primitiveMesh::calcCells does (owned faces
                                     in ascending order, then neighboured faces) */
    const int32_t* cellFaces;
    int32_t nPatches;
    const wenoext_patch* patches;
    const int32_t* patchNbrCell; /* optional, see wenoext_plan_desc; may be NULL */
} wenoext_mesh_desc;

/* labelBytes 0 = detect. The result is a wenoext_setup like wenoext_setup_box's
 * (wenoext_setup_desc gives the plan descriptor); its geometry arrays are empty. */
int wenoext_cache_read(const char* dir, const wenoext_mesh_desc* mesh, int32_t myProc,
                       int labelBytes, wenoext_setup** out);
/* The same from an OpenFOAM case (or processor<N>) directory: constant/WENOBase<polOrder>/ plus
 * the mesh addressing from constant/polyMesh/{owner, neighbour, boundary} (FoamFile header,
 * ascii or binary; processor / cyclic / empty patches recognised, everything else is a patch
 * with zero correction). No OpenFOAM needed. */
int wenoext_case_read(const char* caseDir, int32_t polOrder, int32_t myProc, int labelBytes,
                      wenoext_setup** out);
/* Mesh addressing of a setup as a wenoext_mesh_desc (pointers valid until destroy). */
int wenoext_setup_mesh(wenoext_setup* s, wenoext_mesh_desc* mesh);

#ifdef __cplusplus
}
#endif
#endif /* WENOEXT_SETUP_H */
