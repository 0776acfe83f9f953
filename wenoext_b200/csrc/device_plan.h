// Device-side plan: the reference's nested WENOBase containers (a11 in
// SURVEY.md 8a) flattened into one self-contained, 16-byte aligned RECORD per
// cell so that a single TMA bulk copy brings everything static a cell needs
// into shared memory.
//
// Record layout (byte offsets from the record start; K = nDvt):
//   RecHeader                                   16 B
//   double A[sum_s K*N_s]    LSmatrix()[cell][s], surviving stencils in list order
//   double B[K*K]            B()[cell]
//   double faceD[nFc][K+1]   own-side intBasTrans()[f][side] compacted in sumFlux
//                            enumeration order (WENOUpwindFit.C:304-319), refFacAr()[f]
//   double quirk[K]          intBasTrans()[cellI][1] compacted (WENOUpwindFit01.C:125)
//   int32  N[nSt]            columns of each surviving stencil
//   int32  idx[sum_s N_s]    flat gather index into the value buffer
//                            (local id, or nCells + haloOffset(proc) + haloId)
//   int32  faceI[nFc][3]     face id, flat index of the cell across the face (-1 none), kind
//   pad to 16 B
// Only internal and coupled-patch faces are listed (others get zero correction).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace wenoext {

struct RecHeader {
    int32_t nSt;   // surviving stencils (0 for Cell::empty)
    int32_t nFc;   // listed faces
    int32_t totN;  // sum of N_s
    int32_t flags; // bit0: some stencil entry is a halo cell
};

// kind bits of a listed face
enum : int32_t {
    FACE_SIDE_NEI = 1,      // this cell is the face's neighbour (side 1)
    FACE_TYPE_SHIFT = 1,    // (kind >> 1): 0 internal, 1 processor, 2 cyclic
    FACE_INTERNAL = 0,
    FACE_PROCESSOR = 1,
    FACE_CYCLIC = 2
};

// Arguments of the fused cell kernel (by value, __grid_constant__).
struct CellKernelArgs {
    const unsigned char* records;
    const int64_t* recOff;     // [nCells+1] byte offsets
    const int32_t* cellList;   // nullable: work item -> cell
    int32_t nWork;
    int32_t K;
    int32_t nComp;             // stride of corr / faceValue / thetaPatch
    int32_t c0;                // first component of this batch
    int32_t VS;                // stride (doubles) of the packed value buffer
    const double* vals;        // [nCells+nHalo][VS]
    const double* flux;        // [nFaces]
    double* corr;              // [nFaces][nComp]
    double* faceValue;         // nullable [nInternalFaces][nComp]
    double* thetaPatch;        // nullable [nFaces-nInternalFaces][nComp] (FIT01)
    int32_t nInternalFaces;
    int32_t scheme, bounded, bugCompat;
    uint32_t pExp;
    double dm, eps;
    // shared-memory geometry
    int32_t nSlots, slotBytes;
    int32_t maxTotN, maxSt, maxFc;
    int32_t nConsumerWarps;
};

struct PatchKernelArgs {
    int32_t nBnd;              // boundary faces
    int32_t nInternalFaces;
    int32_t nComp;
    const int32_t* bfPatchType;   // [nBnd] FACE_* or -1 (not coupled)
    const int32_t* bfPartner;     // [nBnd] cyclic: partner FACE index; processor: unused
    const int32_t* bfNbrCell;     // [nBnd] flat value index of the cell across, or -1
    const int32_t* owner;         // [nFaces]
    const double* flux;
    const double* vf;             // [nCells][nComp] (limiter psi)
    const double* vals;           // packed values incl. halo (VS stride) for psiN, may be null
    int32_t VS, c0, CB;
    const double* patchNbrField;  // nullable [nBnd][nComp]
    const double* recv;           // [nBnd][nComp] neighbour ranks' btsfUD
    const double* thetaPatch;     // FIT01
    double* corr;
    int32_t scheme, bounded;
};

void launchCellKernel(const CellKernelArgs& a, int CB, int gridBlocks, size_t smemBytes,
                      cudaStream_t st);
size_t cellKernelSmemBytes(int K, int CB, int nSlots, int slotBytes, int maxTotN, int maxSt,
                           int maxFc, int nConsumerWarps);
void launchPatchKernel(const PatchKernelArgs& a, cudaStream_t st);
void launchPackValues(const double* vf, int nCells, int nComp, int c0, int CB, int VS,
                      double* vals, cudaStream_t st);
void launchPackHalo(const double* vals, int VS, const int32_t* sendIds, int64_t nSend,
                    double* sendBuf, cudaStream_t st);

}  // namespace wenoext
