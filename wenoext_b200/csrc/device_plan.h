// Device-side plan: the reference's nested WENOBase containers (a11 in
// SURVEY.md 8a) flattened into one self-contained, 16-byte aligned RECORD per
// cell. A record is a sequence of PARTS, each fetched into shared memory by one
// 1-D TMA bulk copy exactly when the kernel needs it:
//   head | tail | chunk 0 | chunk 1 | ...
//
// The fused kernel gives one LANE to every (cell, stencil) pair: the lane owns
// all K coefficients of its stencil for all components in registers. The record
// is laid out for that mapping (stencil index fastest), so that the lanes of a
// cell read consecutive shared-memory words.
//
// Layout (K = nDvt, nSt = surviving stencils in list order, so stencil 0 is the
// "first surviving" one that gets dm, WENOCoeff.C:198; Nmax = max_s N_s;
// D = chunkCols(K) matrix columns per chunk):
//   HEAD (needed from the start of the cell to its end)
//     RecHeader                                   32 B
//     int32  N[nSt]              columns of each stencil, pad to 16 B
//     int32  faceI[nFc][4]       face id, flat index of the cell across (-1 none), kind, 0
//     int32  idx[min(GA*D,Nmax)][nSt]  gather indices of the columns of chunks 0..GA-1
//                                (GA = GATHER_AHEAD)
//     pad to 16 B                                          -> headBytes
//   TAIL (needed after the matrix phase)
//     double Bs[K*(K+1)/2]       B()[cell] folded for the quadratic form: row k holds
//                                B[k][k], then B[k][q] + B[q][k] for q > k (symTriSize)
//     double faceD[nFc][K+1]     own-side intBasTrans()[f][side] compacted in sumFlux
//                                enumeration order (WENOUpwindFit.C:304-319), refFacAr()[f]
//     double quirk[K]            intBasTrans()[cellI][1] compacted (WENOUpwindFit01.C:125)
//     pad to 16 B                                          -> tailBytes
//   CHUNK ch: columns [ch*D, ch*D+cols) of LSmatrix()[cell][s] for all stencils
//   (bitwise copies, permuted), cols = min(D, Nmax - ch*D), followed by the gather
//   indices of the columns of chunk ch+GA (icols = clamp(Nmax - (ch+GA)*D, 0, D)):
//     double2 pairs[cols][K/2][nSt]   (A_s[2q][j], A_s[2q+1][j])
//     double  single[cols][nSt]       A_s[K-1][j]            (K odd only)
//     int32   idx[icols][nSt]
//     pad to 16 B
//   Matrix entries are zero where j >= N_s (only when stencil sizes differ, i.e.
//   bestConditioned). A gather index is the flat index of entry j of stencil s into
//   the value buffer (local id, or nCells + haloOffset(proc) + haloId); the cell
//   itself where j >= N_s (so that the value difference is exactly 0).
// Only internal and coupled-patch faces are listed (others get zero correction).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace wenoext {

struct RecHeader {
    int32_t nSt;    // surviving stencils (0 for Cell::empty)
    int32_t nFc;    // listed faces
    int32_t Nmax;   // max columns over the stencils
    int32_t flags;  // bit0: some stencil entry is a halo cell
    int32_t oFace, oIdx;           // byte offsets inside the head
    int32_t headBytes, tailBytes;  // part sizes
};

// per cell: where its record lives and how big its parts are (one 16-byte load
// gives the issuing lane everything it needs)
struct RecInfo {
    int64_t off;                   // byte offset of the record
    uint16_t headBytes, tailBytes;
    int16_t nSt, Nmax;             // chunk sizes follow from these
};

// columns per matrix chunk (also the unroll depth of the kernel's column loop)
__host__ __device__ constexpr int chunkCols(int K) { return K <= 9 ? 3 : 2; }
// c^T B c = sum_k c_k (B_kk c_k + sum_{q>k} (B_kq + B_qk) c_q): K(K+1)/2 folded entries
__host__ __device__ constexpr int symTriSize(int K) { return K * (K + 1) / 2; }
// the gather indices of chunk ch + GATHER_AHEAD travel with chunk ch (the kernel
// issues the gathers of a column GATHER_AHEAD chunk steps before it needs the value)
constexpr int GATHER_AHEAD = 2;
__host__ __device__ inline int clampCols(int n, int D) { return n < 0 ? 0 : (n > D ? D : n); }
// bytes of chunk ch of a cell with nSt stencils and Nmax columns
__host__ __device__ inline int chunkBytes(int K, int nSt, int Nmax, int ch) {
    const int D = chunkCols(K);
    const int cols = clampCols(Nmax - ch * D, D);
    const int icols = clampCols(Nmax - (ch + GATHER_AHEAD) * D, D);
    return (cols * nSt * K * 8 + icols * nSt * 4 + 15) & ~15;
}
// byte offset of the travelling gather indices inside a chunk holding `cols` columns
__host__ __device__ inline int chunkIdxOffset(int K, int nSt, int cols) { return cols * nSt * K * 8; }
// byte offset of chunk ch from the first chunk: all chunks up to cf1 are full
__host__ __device__ inline int64_t chunkOffset(int K, int nSt, int Nmax, int ch) {
    const int D = chunkCols(K);
    int cf1 = Nmax / D - GATHER_AHEAD;  // first chunk that may not be full
    if (cf1 < 0) cf1 = 0;
    const int full = (D * nSt * K * 8 + D * nSt * 4 + 15) & ~15;
    if (ch <= cf1) return (int64_t)ch * full;
    int64_t o = (int64_t)cf1 * full;
    for (int c = cf1; c < ch; ++c) o += chunkBytes(K, nSt, Nmax, c);
    return o;
}

// kind bits of a listed face
enum : int32_t {
    FACE_SIDE_NEI = 1,      // this cell is the face's neighbour (side 1)
    FACE_TYPE_SHIFT = 1,    // (kind >> 1): 0 internal, 1 processor, 2 cyclic
    FACE_INTERNAL = 0,
    FACE_PROCESSOR = 1,
    FACE_CYCLIC = 2
};

__host__ __device__ inline RecHeader recLayout(int K, int nSt, int Nmax, int nFc) {
    const int D = chunkCols(K);
    RecHeader h;
    h.nSt = nSt;
    h.nFc = nFc;
    h.Nmax = Nmax;
    h.flags = 0;
    int o = ((int)sizeof(RecHeader) + 4 * nSt + 15) & ~15;
    h.oFace = o;
    o += 16 * nFc;
    h.oIdx = o;
    o += 4 * clampCols(Nmax, GATHER_AHEAD * D) * nSt;
    h.headBytes = (o + 15) & ~15;
    h.tailBytes = (8 * (symTriSize(K) + nFc * (K + 1) + K) + 15) & ~15;
    return h;
}
__host__ __device__ inline int64_t recMatrixBytes(int K, int nSt, int Nmax) {
    const int D = chunkCols(K);
    int64_t b = 0;
    for (int ch = 0; ch * D < Nmax; ++ch) b += chunkBytes(K, nSt, Nmax, ch);
    return b;
}

// Arguments of the fused cell kernel (by value, __grid_constant__).
struct CellKernelArgs {
    const unsigned char* records;
    const RecInfo* recInfo;    // [nCells]
    const int32_t* cellList;   // nullable: work item -> cell
    int32_t workBase;          // first work item of this launch (slabs of one call)
    int32_t* tileCounter;      // [2], zeroed before the launch: next tile of the two ranges
    int32_t nWork;
    // Halo gate (parallel plans, interior and boundary cells in ONE launch): work items
    // [0, gateItem) are interior cells, [gateItem, nWork) boundary cells, which read halo
    // values. A warp that has run out of interior tiles goes on with boundary tiles
    // (tileCounter[1]) if gateFlags[q] >= gateEpoch for all nGate senders (raised by their
    // halo_push_kernel once exchange #1 has landed in `vals`) and stops otherwise; a
    // follow-up launch with the same counters and nGate == 0 takes what is left.
    // gateItem < 0: one range, no gate.
    // 2-D bulk copies (plan.cu: buildTmaMaps): CUtensorMap[2 + chunks] in global memory or
    // nullptr; part p (0 head, 1 tail, 2 + ch chunk step ch) of the layout (tmaZ, tmaW) has
    // tmaBytes[p] bytes per cell and starts tmaElemOff[p] 8-byte elements into the record
    const void* tmaMaps;
    uint32_t tmaZ;
    int32_t tmaW;
    uint32_t tmaBytes[18], tmaElemOff[18];
    const int32_t* gateFlags;
    int32_t nGate, gateEpoch, gateItem;
    const int32_t* abortFlag;  // nullable, DEVICE memory: non-zero = the halo of this step never arrived, do nothing
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
    int32_t cellsPerWarp;      // 4 or 1: 32 / cellsPerWarp stencil lanes per cell
    int32_t nWarps;            // warps per CTA (each runs its own TMA pipeline)
    int32_t ringSteps;         // chunk steps in flight per warp (ring depth, <= nChunks)
    int32_t headSlotBytes, tailSlotBytes, chunkSlotBytes;  // per-cell shared-memory slots
    int32_t warpBytes;         // shared memory per warp: 2 heads + tail + ring, x cellsPerWarp
    int32_t nChunks;           // chunk steps per cell: ceil(maxN / chunkCols(K))
    int32_t maxN;              // max stencil columns in the plan
    // developer aid (only read when built with -DWENOEXT_KDEBUG): first violated
    // bounds check is recorded in dbg[0..15]
    int32_t nVals, nFaces;
    int32_t* dbg;
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

// fvc::surfaceIntegrate(flux * face value) per cell (wenoext_div)
struct DivKernelArgs {
    int32_t nCells, nInternalFaces, nComp;  // nCells: how many cells this launch integrates
    int32_t cell0;                 // ... starting at this one
    const int32_t* cellFaceStart;  // [nCells+1]
    const int32_t* cellFaces;      // faces of a cell in ascending face index
    const int32_t* owner;          // [nFaces]
    const int32_t* neighbour;      // [nInternalFaces]
    const int32_t* bfPatchType;    // [nBnd] FACE_* or -1 (not coupled)
    const int32_t* bfNbrCell;      // [nBnd] flat value index across a coupled face, or null
    const double* vf;              // [nCells][nComp]
    const double* vals;            // packed values incl. halo (stride VS) of the LAST component batch
    int32_t VS, c0, CB;
    const double* flux;            // [nFaces]
    const double* corr;            // [nFaces][nComp]
    const double* patchNbrField;   // nullable [nBnd][nComp]
    const double* bndValue;        // [nBnd][nComp]
    const double* cellVolumes;     // [nCells]
    double* div;                   // [nCells][nComp]
};
void launchDivKernel(const DivKernelArgs& a, cudaStream_t st);
void launchGatherRows(const double* src, const int32_t* rows, int64_t n, int nComp, double* dst,
                      cudaStream_t st);

void launchScatterRows(const double* src, const int32_t* rows, int64_t n, int nComp, double* dst,
                       cudaStream_t st);

int launchCellKernel(const CellKernelArgs& a, int CB, int gridBlocks, size_t smemBytes,
                     cudaStream_t st);
size_t cellKernelSmemBytes(const CellKernelArgs& a);
bool cellKernelSupportsK(int K);
int cellKernelMaxCB(int K);  // components batched per matrix read
int cellKernelMaxWarps();  // warps per CTA the kernel is compiled for
void launchPatchKernel(const PatchKernelArgs& a, cudaStream_t st);
void launchPackValues(const double* vf, int nCells, int nComp, int c0, int CB, int VS,
                      double* vals, cudaStream_t st);
void launchPackHalo(const double* vals, int VS, const int32_t* sendIds, int64_t nSend,
                    double* sendBuf, cudaStream_t st);
struct HaloPushArgs {
    int32_t nPeers, VS, epoch;
    const double* vals;        // this rank's packed values (exchange #2: corr, VS = nComp)
    const int32_t* sendIds;    // sendHaloCellIDList flattened, peer after peer; nullptr: rows src0[p] + j
    int64_t off[33];           // [nPeers+1] running count of cells pushed to peers 0..i-1
    int64_t src0[32];          // first entry of peer i in sendIds
    double* dst[32];           // receiver's value buffer at nCells + haloOffset(this rank)
    int32_t* flag[32];         // receiver's flag for this rank
    int32_t* done;             // block counter, zero between launches
};
void launchHaloPush(const HaloPushArgs& a, cudaStream_t st);
void launchHaloWait(const int32_t* flags, int n, int32_t epoch, cudaStream_t st, int32_t* status,
                    int32_t* statusDev);

}  // namespace wenoext
