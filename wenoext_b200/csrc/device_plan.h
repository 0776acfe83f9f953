// Device-side plan: the reference's nested WENOBase containers (a11 in
// SURVEY.md 8a) flattened into one self-contained, 16-byte aligned RECORD per
// cell, so that two TMA bulk copies (meta part, matrix part) bring everything
// static a cell needs into shared memory.
//
// Record layout (byte offsets from the record start; K = nDvt, L = lanes of the
// warp group that processes one cell, fixed per plan):
//   META PART
//   RecHeader                                   16 B
//   int32  N[nSt]            columns of each surviving stencil (list order, so
//                            stencil 0 is the "first surviving" one that gets dm,
//                            WENOCoeff.C:198)
//   int32  stOff[nSt]        prefix sum of N (entry offset of the stencil's b block)
//   int32  rnd[nRounds][2]   per round of L rows: {jmax, offset of its A block in doubles}
//   int32  idx[sum_s N_s]    flat gather index into the value buffer
//                            (local id, or nCells + haloOffset(proc) + haloId)
//   int32  faceI[nFc][3]     face id, flat index of the cell across the face (-1 none), kind
//   pad to 8 B
//   double B[K*K]            B()[cell]
//   double faceD[nFc][K+1]   own-side intBasTrans()[f][side] compacted in sumFlux
//                            enumeration order (WENOUpwindFit.C:304-319), refFacAr()[f]
//   double quirk[K]          intBasTrans()[cellI][1] compacted (WENOUpwindFit01.C:125)
//   pad to 16 B                                          -> metaBytes
//   MATRIX PART
//   double A2[...]           LSmatrix()[cell][s] of the surviving stencils, permuted
//                            for the kernel's lane-per-row mapping: row p = s*K + k
//                            belongs to round r = p / L; a round with R rows is stored
//                            column-interleaved, element (j, p) at j*R + (p - r*L), so
//                            the R lanes read consecutive doubles (no bank conflicts)
//                            for every j. Rows shorter than the round's jmax are
//                            zero-filled (only when stencil sizes differ inside a
//                            round, i.e. bestConditioned).
//   pad to 16 B                                          -> recBytes
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

// per cell: where its record lives (one 16-byte load for the TMA producer)
struct RecInfo {
    int64_t off;        // byte offset of the record
    int32_t totBytes;   // whole record
    int32_t metaBytes;  // meta part (multiple of 16)
};

// kind bits of a listed face
enum : int32_t {
    FACE_SIDE_NEI = 1,      // this cell is the face's neighbour (side 1)
    FACE_TYPE_SHIFT = 1,    // (kind >> 1): 0 internal, 1 processor, 2 cyclic
    FACE_INTERNAL = 0,
    FACE_PROCESSOR = 1,
    FACE_CYCLIC = 2
};

__host__ __device__ inline int recRounds(int nSt, int K, int L) { return (nSt * K + L - 1) / L; }

// byte offsets of the meta sections
struct RecLayout {
    int32_t oN, oStOff, oRnd, oIdx, oFaceI, oB, oFaceD, oQuirk, metaBytes;
};
__host__ __device__ inline RecLayout recLayout(int K, int L, int nSt, int totN, int nFc) {
    RecLayout l;
    int o = (int)sizeof(RecHeader);
    l.oN = o;
    o += 4 * nSt;
    l.oStOff = o;
    o += 4 * nSt;
    l.oRnd = o;
    o += 8 * recRounds(nSt, K, L);
    l.oIdx = o;
    o += 4 * totN;
    l.oFaceI = o;
    o += 12 * nFc;
    o = (o + 7) & ~7;
    l.oB = o;
    o += 8 * K * K;
    l.oFaceD = o;
    o += 8 * nFc * (K + 1);
    l.oQuirk = o;
    o += 8 * K;
    l.metaBytes = (o + 15) & ~15;
    return l;
}

// Arguments of the fused cell kernel (by value, __grid_constant__).
struct CellKernelArgs {
    const unsigned char* records;
    const RecInfo* recInfo;    // [nCells]
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
    // launch geometry
    int32_t G;                 // warps per cell group (L = 32*G, fixed by the plan)
    int32_t nGroups;           // cell groups per CTA
    int32_t nSlots, slotBytes; // shared-memory ring of records
    int32_t maxTotN, maxRows, maxSt, maxFc;
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

int launchCellKernel(const CellKernelArgs& a, int CB, int gridBlocks, size_t smemBytes,
                     cudaStream_t st);
// shared memory of one cell group's scratch / of the whole CTA
size_t groupScratchBytes(int K, int CB, int maxTotN, int maxRows, int maxSt, int maxFc);
size_t cellKernelSmemBytes(const CellKernelArgs& a, int CB);
void launchPatchKernel(const PatchKernelArgs& a, cudaStream_t st);
void launchPackValues(const double* vf, int nCells, int nComp, int c0, int CB, int VS,
                      double* vals, cudaStream_t st);
void launchPackHalo(const double* vals, int VS, const int32_t* sendIds, int64_t nSend,
                    double* sendBuf, cudaStream_t st);

}  // namespace wenoext
