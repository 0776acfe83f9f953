// Reader and writer of the reference's on-disk cache `constant/WENOBase<order>/`
// (SURVEY.md 8f-2): the eleven files WENOBase::writeList produces and
// WENOBase::readList consumes (libWENOEXT/WENOBase/WENOBase.C:1196-1312), in
// OpenFOAM's BINARY stream framing, without OpenFOAM.
//
// What the files hold, in the reference's write order (WENOBase.C:1273-1311):
//   sendProcList, receiveProcList   labelList        [nProcs]
//   DimLists                        labelListList    [nCells][3]
//   StencilIDs, CellToProcMap       List<labelListList> [nCells][nStencils][entries]
//   Pseudoinverses                  matrixDB         (matrixDB.C:276-357)
//   sendHaloCellIDList              labelListList    [nProcs][...]
//   receiveHaloSize                 labelList        [nProcs]
//   B                               List<blaze::DynamicMatrix<double>> [nCells]
//   intBasTrans                     count + per face two List3D<scalar> (List3D.H:295-321)
//   refFacAr                        List<scalar>     [nFaces]
//
// Stream framing (OpenFOAM OSstream/ISstream, not part of the reference tree):
// numbers are written as text even in BINARY mode; a raw block is `(` + bytes + `)`;
// a list of a contiguous type is `N` then, if N > 0, one raw block of N*sizeof(T)
// bytes; a list of anything else is `N ( elem elem ... )`. Blaze matrices are
// `rows cols spacing (raw)` with rows*spacing doubles for a row-major matrix
// (BlazeIO/BlazeIO.C:103-134); List3D is `sx sy sz (raw)`.
//
// The mesh addressing (owner, neighbour, cells, patches) is NOT in the cache: the
// reference takes it from the fvMesh it is constructed with, so the caller
// supplies it (HostPlan members owner/neighbour/cellFaces/patches are left alone).
#pragma once
#include <string>

#include "weno_setup.h"

namespace wenoext {

struct CacheFormat {
    int labelBytes = 4;   // WM_LABEL_SIZE/8 (CMakeLists.txt:216); 0 on read = detect
    int simdDoubles = 8;  // Blaze pads a row to a multiple of the SIMD width (8 with AVX-512,
                          // 4 with AVX/AVX2, 2 with SSE2): 9x21 -> spacing 24, matrixDB.H:69
};

// Writes the eleven files into `dir` (created if missing).
void writeWenoBaseCache(const HostPlan& hp, const std::string& dir, const CacheFormat& fmt);

// Fills the WENOBase members of `hp` from `dir`: polOrder, nDvt, dim, nCells (from
// StencilIDs), nFaces (from refFacAr), the stencil CSR, the matrix pool (database order),
// B, intBasTrans, refFacAr and the halo lists. `myProc` is not stored in the cache.
// On return fmt holds the label size found and the row padding of the first matrix.
void readWenoBaseCache(const std::string& dir, int32_t myProc, HostPlan& hp, CacheFormat& fmt);

// constant/polyMesh/{owner, neighbour, boundary} of an OpenFOAM case (or processor) directory:
// the mesh addressing the cache does not hold. FoamFile header, `format ascii` or `binary`
// (label size from the header's `arch` entry, else detected); `boundary` is a list of
// dictionaries with type / nFaces / startFace (+ neighbProcNo, neighbourPatch). Patch types
// map to the kinds the hot path distinguishes (WENOUpwindFit.C:433, :478, :493): processor
// and processorCyclic -> PATCH_PROCESSOR, cyclic -> PATCH_CYCLIC, empty -> PATCH_EMPTY, anything
// else (wall, patch, symmetry, cyclicAMI ...) -> PATCH_GENERIC (zero correction, as the
// reference leaves them).
void readPolyMeshAddressing(const std::string& polyMeshDir, HostPlan& hp);

// matrixDB::hashMatrix (matrixDB.C:171-199), shared with the setup's MatrixPool.
int32_t matrixKey(const double* A, size_t n);

}  // namespace wenoext
