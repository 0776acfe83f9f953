// Device plan creation (flatten + upload + on-device record expansion) and the
// per-call orchestration of correction(). See include/wenoext_b200.h.
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>
#include <climits>
#include <dlfcn.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/wenoext_b200.h"
#include "device_plan.h"
#include "plan_internal.h"

namespace wenoext {

thread_local std::string g_lastError;

void setError(const std::string& s) { g_lastError = s; }

#define CUDA_TRY(expr)                                                                   \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess) {                                                         \
            setError(std::string(#expr) + ": " + cudaGetErrorString(_e));                \
            cudaGetLastError(); /* reported: do not leave it for the next call */        \
            return WENOEXT_ERR_CUDA;                                                     \
        }                                                                                \
    } while (0)

namespace {

template <class T>
int upload(T** dptr, const T* h, size_t n) {
    *dptr = nullptr;
    if (n == 0) return 0;
    CUDA_TRY(cudaMalloc((void**)dptr, n * sizeof(T)));
    CUDA_TRY(cudaMemcpy(*dptr, h, n * sizeof(T), cudaMemcpyHostToDevice));
    return 0;
}

struct ExpandArgs {
    int32_t nCells, nFaces, nInternalFaces, K, cube;
    const RecInfo* recInfo;
    const int32_t *cellNSt, *cellNmax, *cellNFc;
    const int64_t *cellStencilStart, *stencilEntryStart;
    const int32_t *stencilsID, *cellToProcMap, *stencilMatrix;
    const int64_t* matrixStart;
    const int32_t* matrixCols;
    const double* matrixData;
    const double *B, *intBasTrans, *refFacAr;
    const int32_t *owner, *neighbour, *cellFaceStart, *cellFaces;
    const int32_t* bfPatchType;  // [nBnd]
    const int32_t* bfNbrCell;    // [nBnd] or null
    const int64_t* haloOff;      // [nProcs]
    const int32_t* enumIdx;      // [K] cube index of coefficient k
    unsigned char* records;
    uint8_t* cellHasHalo;
    int32_t* cellExt;  // [nCells][3]: max local value index read, min / max listed face
};

// One warp per cell: copies the de-duplicated matrices and the per-cell / per-face
// setup data into the cell's record (bitwise copies, so parity is unaffected). The
// record memory is zero-filled beforehand (columns beyond a stencil's size).
__global__ void expand_records_kernel(const ExpandArgs a) {
    const int cell = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (cell >= a.nCells) return;
    const int K = a.K, KP = K / 2;
    unsigned char* rec = a.records + a.recInfo[cell].off;
    const int nSt = a.cellNSt[cell], Nmax = a.cellNmax[cell], nFc = a.cellNFc[cell];
    const RecHeader lay = recLayout(K, nSt, Nmax, nFc);
    int32_t* Ns = reinterpret_cast<int32_t*>(rec + sizeof(RecHeader));
    int32_t* faceI = reinterpret_cast<int32_t*>(rec + lay.oFace);
    double* Bm = reinterpret_cast<double*>(rec + lay.headBytes);
    double* faceD = Bm + symTriSize(K);
    double* quirk = faceD + (size_t)nFc * (K + 1);
    const int D = chunkCols(K);
    unsigned char* chunks = rec + lay.headBytes + lay.tailBytes;
    auto chunkAt = [&](int ch) { return chunks + chunkOffset(K, nSt, Nmax, ch); };
    // gather-index row of column j: the rows of chunks 0..GA-1 are in the head, the
    // others in chunk ch-GA (which always holds D columns)
    auto idxRow = [&](int j) {
        const int ch = j / D, jc = j - ch * D;
        unsigned char* base = ch < GATHER_AHEAD
                                  ? rec + lay.oIdx + (size_t)ch * D * nSt * 4
                                  : chunkAt(ch - GATHER_AHEAD) + chunkIdxOffset(K, nSt, D);
        return reinterpret_cast<int32_t*>(base) + (size_t)jc * nSt;
    };
    for (int i = lane; i < Nmax * nSt; i += 32) {
        const int j = i / nSt;
        idxRow(j)[i - j * nSt] = cell;  // unused entries: a valid index, difference 0
    }
    __syncwarp();
    bool halo = false;
    int maxId = cell, minFace = INT32_MAX, maxFace = -1;  // what the cell reads (extents)
    if (nSt > 0) {
        int live = 0;
        for (int64_t s = a.cellStencilStart[cell]; s < a.cellStencilStart[cell + 1]; ++s) {
            const int m = a.stencilMatrix[s];
            if (m < 0) continue;
            const int N = a.matrixCols[m];
            const double* src = a.matrixData + a.matrixStart[m];
            if (lane == 0) Ns[live] = N;
            for (int i = lane; i < K * N; i += 32) {
                const int k = i / N, j = i - k * N;
                const int ch = j / D, jc = j - ch * D;
                const int cols = min(D, Nmax - ch * D);
                double* chunk = reinterpret_cast<double*>(chunkAt(ch));
                if (k < 2 * KP)
                    chunk[((size_t)(jc * KP + (k >> 1)) * nSt + live) * 2 + (k & 1)] = src[i];
                else
                    chunk[(size_t)cols * KP * nSt * 2 + (size_t)jc * nSt + live] = src[i];
            }
            const int64_t e0 = a.stencilEntryStart[s];
            for (int j = lane; j < N; j += 32) {
                const int32_t id = a.stencilsID[e0 + 1 + j];
                const int32_t pm = a.cellToProcMap ? a.cellToProcMap[e0 + 1 + j] : -1;
                int32_t flat = id;
                if (pm >= 0) {
                    flat = a.nCells + (int32_t)a.haloOff[pm] + id;
                    halo = true;
                } else {
                    maxId = max(maxId, id);
                }
                idxRow(j)[live] = flat;
            }
            live++;
        }
    }
    // B folded for the quadratic form (device_plan.h): diagonal, then B[k][q] + B[q][k]
    for (int i = lane; i < K * K; i += 32) {
        const int k = i / K, q = i - k * K;
        if (q < k) continue;
        const double* Bc = a.B + (size_t)cell * K * K;
        Bm[k * K - k * (k - 1) / 2 + (q - k)] = q == k ? Bc[i] : Bc[i] + Bc[q * K + k];
    }
    for (int k = lane; k < K; k += 32)
        quirk[k] = cell < a.nFaces ? a.intBasTrans[((size_t)cell * 2 + 1) * a.cube + a.enumIdx[k]] : 0.0;
    int fOut = 0;
    for (int i = a.cellFaceStart[cell]; i < a.cellFaceStart[cell + 1]; ++i) {
        const int f = a.cellFaces[i];
        int side = 0, other = -1, type = FACE_INTERNAL;
        if (f < a.nInternalFaces) {
            if (a.owner[f] == cell) {
                other = a.neighbour[f];
            } else {
                side = 1;
                other = a.owner[f];
            }
        } else {
            type = a.bfPatchType[f - a.nInternalFaces];
            if (type < 0) continue;
            other = a.bfNbrCell ? a.bfNbrCell[f - a.nInternalFaces] : -1;
        }
        if (other >= 0 && other < a.nCells) maxId = max(maxId, other);
        minFace = min(minFace, f);
        maxFace = max(maxFace, f);
        double* fd = faceD + (size_t)fOut * (K + 1);
        for (int k = lane; k < K; k += 32)
            fd[k] = a.intBasTrans[((size_t)f * 2 + side) * a.cube + a.enumIdx[k]];
        if (lane == 0) {
            fd[K] = a.refFacAr[f];
            faceI[fOut * 4 + 0] = f;
            faceI[fOut * 4 + 1] = other;
            faceI[fOut * 4 + 2] = side | (type << FACE_TYPE_SHIFT);
            faceI[fOut * 4 + 3] = 0;
        }
        fOut++;
    }
    halo = __any_sync(0xffffffffu, halo);
    maxId = __reduce_max_sync(0xffffffffu, maxId);
    if (lane == 0) {
        a.cellExt[(size_t)cell * 3 + 0] = maxId;
        a.cellExt[(size_t)cell * 3 + 1] = minFace;
        a.cellExt[(size_t)cell * 3 + 2] = maxFace;
        RecHeader h = lay;
        h.flags = halo ? 1 : 0;
        *reinterpret_cast<RecHeader*>(rec) = h;
        a.cellHasHalo[cell] = halo ? 1 : 0;
    }
}

}  // namespace
}  // namespace wenoext

using namespace wenoext;

// Tensor maps for the 2-D bulk copies of the cell kernel: the records as a matrix of
// nCells rows of recStride bytes; the box of part p (head, tail, chunk step ch) of the
// dominant layout is [bytes_p / 8 elements] x [4 cells]. One copy then serves the four cells
// of a pass, instead of four per-lane copies behind a serialising loop. Failure of any step
// just leaves the per-lane path in charge.
static void buildTmaMaps(wenoext_plan* P) {
    const int K = P->K, D = chunkCols(K);
    const int nSt = P->tmaW & 0xffff, Nmax = P->tmaW >> 16;
    const uint32_t headB = P->tmaZ & 0xffffu, tailB = P->tmaZ >> 16;
    const int nCh = (Nmax + D - 1) / D;
    if (nSt <= 0 || Nmax <= 0 || nCh + 2 > 18 || P->recStride % 16) return;
    uint32_t bytes[18] = {0}, off[18] = {0};
    bytes[0] = headB;
    bytes[1] = tailB;
    off[1] = headB;
    uint32_t run = headB + tailB;
    for (int ch = 0; ch < nCh; ++ch) {
        bytes[2 + ch] = (uint32_t)chunkBytes(K, nSt, Nmax, ch);
        off[2 + ch] = run;
        run += bytes[2 + ch];
    }
    for (int p = 0; p < nCh + 2; ++p)
        if (bytes[p] == 0 || bytes[p] % 16 || bytes[p] / 8 > 256 || off[p] % 16) return;
    PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !encode) {
        cudaGetLastError();
        return;
    }
    std::vector<CUtensorMap> maps(nCh + 2);
    for (int p = 0; p < nCh + 2; ++p) {
        const cuuint64_t gdim[2] = {(cuuint64_t)(P->recStride / 8), (cuuint64_t)P->nCells};
        const cuuint64_t gstr[1] = {(cuuint64_t)P->recStride};
        const cuuint32_t box[2] = {bytes[p] / 8, 4};
        const cuuint32_t estr[2] = {1, 1};
        if (encode(&maps[p], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, P->records, gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return;
    }
    if (cudaMalloc((void**)&P->tmaMaps, maps.size() * sizeof(CUtensorMap)) != cudaSuccess) {
        cudaGetLastError();
        P->tmaMaps = nullptr;
        return;
    }
    cudaMemcpy(P->tmaMaps, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice);
    for (int p = 0; p < 18; ++p) {
        P->tmaBytes[p] = bytes[p];
        P->tmaElemOff[p] = off[p] / 8;
    }
}

static int validateDesc(const wenoext_plan_desc* d) {
    if (!d) {
        setError("desc is NULL");
        return WENOEXT_ERR_INVALID;
    }
    if (d->nDvt <= 0 || d->nCells <= 0 || d->nFaces < d->nInternalFaces || d->nInternalFaces < 0) {
        setError("desc: bad sizes");
        return WENOEXT_ERR_INVALID;
    }
    if (!d->owner || !d->cellFaceStart || !d->cellFaces || !d->cellStencilStart ||
        !d->stencilEntryStart || !d->stencilsID || !d->stencilMatrix || !d->matrixStart ||
        !d->matrixCols || !d->matrixData || !d->B || !d->intBasTrans || !d->refFacAr) {
        setError("desc: required array is NULL");
        return WENOEXT_ERR_INVALID;
    }
    if (d->nInternalFaces > 0 && !d->neighbour) {
        setError("desc: neighbour is NULL");
        return WENOEXT_ERR_INVALID;
    }
    if (d->nProcs > 1 && (!d->sendProcList || !d->receiveProcList || !d->receiveHaloSize ||
                          !d->sendHaloStart || !d->cellToProcMap)) {
        setError("desc: parallel plan needs halo lists and cellToProcMap");
        return WENOEXT_ERR_INVALID;
    }
    return 0;
}

// Every index the kernels will follow without a bounds check: cell ids of the stencils (local, or
// halo ids against receiveHaloSize of their processor), matrix ids, face and cell ids of the mesh
// addressing, patches and their partners, the send lists. A stale constant/WENOBase* cache with matching
// counts or a bad descriptor fails here with a message instead of an illegal address in a kernel.
static int validateIds(const wenoext_plan_desc* d) {
    auto bad = [](const std::string& what) {
        setError("desc: " + what);
        return WENOEXT_ERR_INVALID;
    };
    const int nC = d->nCells, nF = d->nFaces, nI = d->nInternalFaces, R = d->nProcs > 1 ? d->nProcs : 0;
    if (d->nStencils < 0 || d->cellStencilStart[0] != 0 || d->cellStencilStart[nC] != d->nStencils)
        return bad("cellStencilStart does not span [0, nStencils]");
    for (int c = 0; c < nC; ++c)
        if (d->cellStencilStart[c + 1] < d->cellStencilStart[c]) return bad("cellStencilStart is not monotone");
    if (d->stencilEntryStart[0] != 0) return bad("stencilEntryStart does not start at 0");
    if (d->nUniqueMatrices < 0 || d->matrixStart[0] != 0) return bad("matrix pool offsets");
    for (int m = 0; m < d->nUniqueMatrices; ++m)
        if (d->matrixCols[m] < 0 || d->matrixStart[m + 1] - d->matrixStart[m] != (int64_t)d->nDvt * d->matrixCols[m])
            return bad("matrixStart does not match nDvt x matrixCols of matrix " + std::to_string(m));
    if (R) {
        if (d->myProc < 0 || d->myProc >= R) return bad("myProc out of range");
        if (d->sendHaloStart[0] != 0) return bad("sendHaloStart does not start at 0");
        for (int p = 0; p < R; ++p) {
            if (d->receiveHaloSize[p] < 0 || d->sendHaloStart[p + 1] < d->sendHaloStart[p])
                return bad("halo sizes of processor " + std::to_string(p));
            for (int64_t i = d->sendHaloStart[p]; i < d->sendHaloStart[p + 1]; ++i)
                if (d->sendHaloCellID[i] < 0 || d->sendHaloCellID[i] >= nC)
                    return bad("sendHaloCellID out of range for processor " + std::to_string(p));
        }
    }
    for (int64_t s = 0; s < d->nStencils; ++s) {
        const int64_t e0 = d->stencilEntryStart[s], e1 = d->stencilEntryStart[s + 1];
        if (e1 <= e0) return bad("stencil " + std::to_string(s) + " has no entries (deleted stencils keep one)");
        if (d->stencilMatrix[s] >= d->nUniqueMatrices) return bad("stencilMatrix out of range");
        const int32_t first = d->stencilsID[e0];
        if (d->stencilMatrix[s] < 0 || first == WENOEXT_CELL_DELETED || first == WENOEXT_CELL_EMPTY) continue;
        for (int64_t e = e0 + 1; e < e1; ++e) {
            const int32_t id = d->stencilsID[e];
            const int32_t pm = d->cellToProcMap ? d->cellToProcMap[e] : WENOEXT_CELL_LOCAL;
            if (pm == WENOEXT_CELL_LOCAL) {
                if (id < 0 || id >= nC) return bad("stencil cell id out of range in stencil " + std::to_string(s));
            } else if (pm != WENOEXT_CELL_DELETED) {
                if (pm < 0 || pm >= R) return bad("cellToProcMap names a processor out of range");
                if (id < 0 || id >= d->receiveHaloSize[pm])
                    return bad("halo id beyond receiveHaloSize[" + std::to_string(pm) + "] in stencil " + std::to_string(s));
            }
        }
    }
    for (int f = 0; f < nF; ++f)
        if (d->owner[f] < 0 || d->owner[f] >= nC) return bad("owner out of range");
    for (int f = 0; f < nI; ++f)
        if (d->neighbour[f] < 0 || d->neighbour[f] >= nC) return bad("neighbour out of range");
    if (d->cellFaceStart[0] != 0) return bad("cellFaceStart does not start at 0");
    for (int c = 0; c < nC; ++c) {
        if (d->cellFaceStart[c + 1] < d->cellFaceStart[c]) return bad("cellFaceStart is not monotone");
        for (int i = d->cellFaceStart[c]; i < d->cellFaceStart[c + 1]; ++i)
            if (d->cellFaces[i] < 0 || d->cellFaces[i] >= nF) return bad("cellFaces out of range");
    }
    if (d->nPatches < 0 || (d->nPatches > 0 && !d->patches)) return bad("patches");
    for (int p = 0; p < d->nPatches; ++p) {
        const wenoext_patch& pt = d->patches[p];
        if (pt.size < 0 || pt.start < nI || (int64_t)pt.start + pt.size > nF)
            return bad("patch " + std::to_string(p) + " lies outside the boundary faces");
        if (pt.type == WENOEXT_PATCH_CYCLIC) {
            if (pt.nbr < 0 || pt.nbr >= d->nPatches || d->patches[pt.nbr].type != WENOEXT_PATCH_CYCLIC ||
                d->patches[pt.nbr].size != pt.size)
                return bad("cyclic patch " + std::to_string(p) + " has no matching neighbour patch");
        } else if (pt.type == WENOEXT_PATCH_PROCESSOR) {
            if (R == 0 || pt.nbr < 0 || pt.nbr >= R || pt.nbr == d->myProc)
                return bad("processor patch " + std::to_string(p) + " names a rank out of range");
        }
    }
    if (d->patchNbrCell) {
        int64_t nHalo = 0;
        for (int p = 0; p < R; ++p) nHalo += d->receiveHaloSize[p];
        for (int f = 0; f < nF - nI; ++f)
            if (d->patchNbrCell[f] < -1 || d->patchNbrCell[f] >= nC + nHalo) return bad("patchNbrCell out of range");
    }
    return 0;
}

extern "C" int wenoext_plan_create(const wenoext_plan_desc* d, int device, wenoext_plan** out) {
    if (!out) {
        setError("out is NULL");
        return WENOEXT_ERR_INVALID;
    }
    *out = nullptr;
    if (int rc = validateDesc(d)) return rc;
    if (int rc = validateIds(d)) return rc;
    int nDev = 0;
    if (cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0) {
        setError("no CUDA device: wenoext_b200 has no CPU fallback");
        return WENOEXT_ERR_CUDA;
    }
    CUDA_TRY(cudaSetDevice(device));
    noAllocResolve(device);  // address of the static scratch of wenoext_correction_noalloc (no allocation)
    wenoext_plan* P = new wenoext_plan();
    P->device = device;
    CUDA_TRY(cudaDeviceGetAttribute(&P->nSM, cudaDevAttrMultiProcessorCount, device));
    CUDA_TRY(cudaDeviceGetAttribute(&P->smemLimit, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    const int K = d->nDvt;
    P->polOrder = d->polOrder;
    P->K = K;
    P->nCells = d->nCells;
    P->nFaces = d->nFaces;
    P->nInt = d->nInternalFaces;
    P->nBnd = d->nFaces - d->nInternalFaces;
    P->nProcs = d->nProcs > 0 ? d->nProcs : 1;
    P->myProc = d->myProc;
    P->patches.assign(d->patches, d->patches + d->nPatches);
    if (const char* e = getenv("WENOEXT_WARPS")) P->cfgWarps = atoi(e);
    if (const char* e = getenv("WENOEXT_CELLS")) P->cfgCells = atoi(e);
    if (const char* e = getenv("WENOEXT_SLOTS")) P->cfgSlots = atoi(e);
    if (const char* e = getenv("WENOEXT_GRID")) P->cfgGrid = atoi(e);
    if (const char* e = getenv("WENOEXT_SMEM_KB")) P->cfgSmemKB = atoi(e);
    if (const char* e = getenv("WENOEXT_MERGE")) P->cfgMerge = atoi(e);
    if (const char* e = getenv("WENOEXT_P2P")) P->cfgP2P = atoi(e);
    if (const char* e = getenv("WENOEXT_PATCH_PUSH")) P->cfgPatchPush = atoi(e);

    // coefficient enumeration: n-outer, m, l-inner (WENOUpwindFit.C:304-319)
    const int s1 = d->polOrder + 1;
    std::vector<int32_t> enumIdx;
    for (int n = 0; n <= d->dim[0]; ++n)
        for (int m = 0; m <= d->dim[1]; ++m)
            for (int l = 0; l <= d->dim[2]; ++l)
                if ((n + m + l) <= d->polOrder && (n + m + l) > 0)
                    enumIdx.push_back(n * s1 * s1 + m * s1 + l);
    if ((int)enumIdx.size() != K) {
        setError("desc: nDvt does not match polOrder/dim");
        delete P;
        return WENOEXT_ERR_INVALID;
    }

    // boundary-face tables
    std::vector<int32_t> bfType(std::max(P->nBnd, 1), -1), bfPartner(std::max(P->nBnd, 1), -1);
    int64_t nCoupled = 0;
    for (const auto& pt : P->patches) {
        if (pt.type != WENOEXT_PATCH_PROCESSOR && pt.type != WENOEXT_PATCH_CYCLIC) continue;
        for (int i = 0; i < pt.size; ++i) {
            const int bf = pt.start + i - P->nInt;
            if (pt.type == WENOEXT_PATCH_PROCESSOR) {
                bfType[bf] = FACE_PROCESSOR;
                P->hasProcessor = true;
            } else {
                bfType[bf] = FACE_CYCLIC;
                bfPartner[bf] = P->patches[pt.nbr].start + i;
            }
            nCoupled++;
        }
    }
    P->hasCoupled = nCoupled > 0;
    P->hasNbrCell = d->patchNbrCell != nullptr;

    // halo offsets
    P->haloOff.assign(P->nProcs + 1, 0);
    P->sendOff.assign(P->nProcs + 1, 0);
    P->sendProcList.assign(P->nProcs, -1);
    P->receiveProcList.assign(P->nProcs, -1);
    if (P->nProcs > 1) {
        for (int p = 0; p < P->nProcs; ++p) {
            P->haloOff[p + 1] = P->haloOff[p] + d->receiveHaloSize[p];
            P->sendOff[p + 1] = d->sendHaloStart[p + 1];
            P->sendProcList[p] = d->sendProcList[p];
            P->receiveProcList[p] = d->receiveProcList[p];
        }
    }
    P->nHaloTotal = P->haloOff[P->nProcs];
    P->nSendTotal = P->sendOff[P->nProcs];

    // per-cell record sizes
    if (!cellKernelSupportsK(K)) {
        setError("unsupported polynomial space (nDvt must be 2, 3, 5, 9, 14, 19, 20 or 34: orders 1-4 in 3-D, 1-5 in 2-D)");
        delete P;
        return WENOEXT_ERR_INVALID;
    }
    const int nC = d->nCells;
    std::vector<int32_t> cellNSt(nC), cellNmax(nC), cellNFc(nC);
    std::vector<RecInfo> recInfo(nC);
    int64_t algFixed = 0, flopsPerComp = 0, nLive = 0, recOff = 0;
    // ONE liveness rule for host sizing and device expansion: a stencil is live if it has a matrix, its
    // first entry is not Cell::deleted and its cell is not Cell::empty (WENOCoeff.C:120, :333); everything
    // else gets matrix id -1 in the array the expansion kernel reads
    std::vector<int32_t> liveMatrix(d->stencilMatrix, d->stencilMatrix + d->nStencils);
    for (int c = 0; c < nC; ++c) {
        const int64_t s0 = d->cellStencilStart[c], sE = d->cellStencilStart[c + 1];
        const bool empty = s0 == sE || d->stencilsID[d->stencilEntryStart[s0]] == WENOEXT_CELL_EMPTY;
        for (int64_t s = s0; s < sE; ++s)
            if (empty || d->stencilsID[d->stencilEntryStart[s]] == WENOEXT_CELL_DELETED) liveMatrix[s] = -1;
    }
    for (int c = 0; c < nC; ++c) {
        int nSt = 0, totN = 0, Nmax = 0;
        const int64_t s0 = d->cellStencilStart[c], sE = d->cellStencilStart[c + 1];
        const bool empty = false;  // (folded into liveMatrix)
        if (!empty)
            for (int64_t s = s0; s < sE; ++s) {
                const int m = liveMatrix[s];
                if (m < 0) continue;
                const int N = d->matrixCols[m];
                if (d->stencilEntryStart[s + 1] - d->stencilEntryStart[s] != N + 1) {
                    setError("desc: stencil size does not match its matrix columns");
                    delete P;
                    return WENOEXT_ERR_INVALID;
                }
                nSt++;
                totN += N;
                Nmax = std::max(Nmax, N);
            }
        int nFc = 0;
        for (int i = d->cellFaceStart[c]; i < d->cellFaceStart[c + 1]; ++i) {
            const int f = d->cellFaces[i];
            if (f < P->nInt || bfType[f - P->nInt] >= 0) nFc++;
        }
        if (nSt > 32) {
            setError("more than 32 stencils in a cell is not supported");
            delete P;
            return WENOEXT_ERR_INVALID;
        }
        cellNSt[c] = nSt;
        cellNmax[c] = Nmax;
        cellNFc[c] = nFc;
        const RecHeader lay = recLayout(K, nSt, Nmax, nFc);
        const int64_t tot = (int64_t)lay.headBytes + lay.tailBytes + recMatrixBytes(K, nSt, Nmax);
        recInfo[c].off = recOff;
        recInfo[c].headBytes = (uint16_t)lay.headBytes;
        recInfo[c].tailBytes = (uint16_t)lay.tailBytes;
        recInfo[c].nSt = (int16_t)nSt;
        recInfo[c].Nmax = (int16_t)Nmax;
        if (lay.headBytes > 65520 || lay.tailBytes > 65520 || Nmax > 32767) {
            setError("cell record parts too large (polynomial order too high)");
            delete P;
            return WENOEXT_ERR_INVALID;
        }
        P->maxHeadBytes = std::max(P->maxHeadBytes, (int64_t)lay.headBytes);
        P->maxTailBytes = std::max(P->maxTailBytes, (int64_t)lay.tailBytes);
        recOff += tot;
        P->maxRecordBytes = std::max(P->maxRecordBytes, tot);
        P->maxN = std::max(P->maxN, Nmax);
        P->maxSt = std::max(P->maxSt, nSt);
        P->maxFc = std::max(P->maxFc, nFc);
        nLive += nSt;
        if (nSt > 0) algFixed += (int64_t)8 * K * totN + 8 * K * K + 4 * totN;
        flopsPerComp += (int64_t)2 * K * totN + (int64_t)nSt * (2 * K * K + 2 * K) + 2 * nSt * K;
    }
    P->maxSt = std::max(P->maxSt, 1);
    P->recordBytesTotal = recOff;
    // Uniform record stride (2-D bulk copies, see buildTmaMaps): when padding every record to the
    // size of the largest costs next to nothing in footprint (nothing in traffic: parts are
    // fetched with their own sizes), cell c's record starts at c * stride, and the same part
    // of four consecutive cells is ONE strided box for the copy engine.
    P->recStride = 0;
    {
        const int64_t stride = (P->maxRecordBytes + 15) / 16 * 16;
        const char* e = getenv("WENOEXT_TMA2D");
        // WENOEXT_TMA2D=0: off; =2: on whatever the padding costs (tests on small meshes)
        const bool force = e && atoi(e) == 2;
        if (nC > 0 && !(e && atoi(e) == 0) && (force || (double)stride * nC <= 1.05 * (double)recOff)) {
            for (int c = 0; c < nC; ++c) recInfo[c].off = (int64_t)c * stride;
            P->recStride = stride;
            P->recordBytesTotal = stride * nC;
        }
    }
    // the most frequent layout (interior cells): the one the 2-D copies are set up for
    {
        std::map<std::pair<uint32_t, int32_t>, int> freq;
        for (int c = 0; c < nC; ++c)
            if (recInfo[c].nSt > 0)
                ++freq[{(uint32_t)recInfo[c].headBytes | ((uint32_t)recInfo[c].tailBytes << 16),
                        (int32_t)recInfo[c].nSt | ((int32_t)recInfo[c].Nmax << 16)}];
        int best = 0;
        for (const auto& kv : freq)
            if (kv.second > best) {
                best = kv.second;
                P->tmaZ = kv.first.first;
                P->tmaW = kv.first.second;
            }
    }
    const int64_t nWritten = (int64_t)P->nInt + nCoupled;
    algFixed += nWritten * (8 * K + 24);
    flopsPerComp += nWritten * 2 * K;

    // upload the flat setup arrays (temporary) and expand on the device
    ExpandArgs ea{};
    ea.nCells = nC;
    ea.nFaces = d->nFaces;
    ea.nInternalFaces = d->nInternalFaces;
    ea.K = K;
    ea.cube = s1 * s1 * s1;
    std::vector<void*> temps;
    auto up = [&](auto** dp, const auto* h, size_t n) {
        int rc = upload(dp, h, n);
        if (!rc && *dp) temps.push_back((void*)*dp);
        return rc;
    };
    int rc = 0;
    RecInfo* dRecInfo = nullptr;
    int32_t *dNSt, *dNmax, *dNFc, *dSID, *dC2P = nullptr, *dSM, *dMC, *dNei = nullptr, *dCFS, *dCF,
                                  *dEnum, *dBfNbr = nullptr;
    int64_t *dCSS, *dSES, *dMS, *dHaloOff;
    double *dMD, *dB, *dIBT, *dRFA;
    const int64_t nEntries = d->stencilEntryStart[d->nStencils];
    rc = rc ? rc : upload(&dRecInfo, recInfo.data(), recInfo.size());
    P->recInfo = dRecInfo;
    rc = rc ? rc : up(&dNSt, cellNSt.data(), (size_t)nC);
    rc = rc ? rc : up(&dNmax, cellNmax.data(), (size_t)nC);
    rc = rc ? rc : up(&dNFc, cellNFc.data(), (size_t)nC);
    rc = rc ? rc : up(&dCSS, d->cellStencilStart, (size_t)nC + 1);
    rc = rc ? rc : up(&dSES, d->stencilEntryStart, (size_t)d->nStencils + 1);
    rc = rc ? rc : up(&dSID, d->stencilsID, (size_t)nEntries);
    if (d->cellToProcMap) rc = rc ? rc : up(&dC2P, d->cellToProcMap, (size_t)nEntries);
    rc = rc ? rc : up(&dSM, liveMatrix.data(), (size_t)d->nStencils);
    rc = rc ? rc : up(&dMS, d->matrixStart, (size_t)d->nUniqueMatrices + 1);
    rc = rc ? rc : up(&dMC, d->matrixCols, (size_t)d->nUniqueMatrices);
    rc = rc ? rc : up(&dMD, d->matrixData, (size_t)d->matrixStart[d->nUniqueMatrices]);
    rc = rc ? rc : up(&dB, d->B, (size_t)nC * K * K);
    rc = rc ? rc : up(&dIBT, d->intBasTrans, (size_t)d->nFaces * 2 * ea.cube);
    rc = rc ? rc : up(&dRFA, d->refFacAr, (size_t)d->nFaces);
    rc = rc ? rc : upload(&P->owner, d->owner, (size_t)d->nFaces);
    if (d->nInternalFaces) rc = rc ? rc : up(&dNei, d->neighbour, (size_t)d->nInternalFaces);
    rc = rc ? rc : up(&dCFS, d->cellFaceStart, (size_t)nC + 1);
    rc = rc ? rc : up(&dCF, d->cellFaces, (size_t)d->cellFaceStart[nC]);
    {   // wenoext_div sums a cell's faces in ascending face index (OpenFOAM's loop order)
        std::vector<int32_t> sorted(d->cellFaces, d->cellFaces + d->cellFaceStart[nC]);
        for (int c = 0; c < nC; ++c)
            std::sort(sorted.begin() + d->cellFaceStart[c], sorted.begin() + d->cellFaceStart[c + 1]);
        rc = rc ? rc : upload(&P->cellFaceStartD, d->cellFaceStart, (size_t)nC + 1);
        rc = rc ? rc : upload(&P->cellFacesSortedD, sorted.data(), sorted.size());
        if (d->nInternalFaces) rc = rc ? rc : upload(&P->neighbourD, d->neighbour, (size_t)d->nInternalFaces);
    }
    rc = rc ? rc : upload(&P->bfPatchType, bfType.data(), bfType.size());
    rc = rc ? rc : upload(&P->bfPartner, bfPartner.data(), bfPartner.size());
    if (d->patchNbrCell && P->nBnd > 0)
        rc = rc ? rc : upload(&P->bfNbrCell, d->patchNbrCell, (size_t)P->nBnd);
    dBfNbr = P->bfNbrCell;
    rc = rc ? rc : up(&dHaloOff, P->haloOff.data(), P->haloOff.size());
    rc = rc ? rc : up(&dEnum, enumIdx.data(), enumIdx.size());
    if (P->nSendTotal > 0) rc = rc ? rc : upload(&P->sendIds, d->sendHaloCellID, (size_t)P->nSendTotal);
    uint8_t* dHasHalo = nullptr;
    int32_t* dCellExt = nullptr;
    if (!rc) {
        cudaError_t e = cudaMalloc((void**)&P->records, (size_t)P->recordBytesTotal);
        if (e == cudaSuccess) e = cudaMemset(P->records, 0, (size_t)P->recordBytesTotal);
        if (e == cudaSuccess) e = cudaMalloc((void**)&dHasHalo, (size_t)nC);
        if (e == cudaSuccess) e = cudaMalloc((void**)&dCellExt, (size_t)nC * 3 * sizeof(int32_t));
        if (e != cudaSuccess) {
            setError(std::string("cudaMalloc(records): ") + cudaGetErrorString(e));
            rc = WENOEXT_ERR_CUDA;
        }
    }
    if (!rc) {
        ea.recInfo = dRecInfo;
        ea.cellNSt = dNSt;
        ea.cellNmax = dNmax;
        ea.cellNFc = dNFc;
        ea.cellStencilStart = dCSS;
        ea.stencilEntryStart = dSES;
        ea.stencilsID = dSID;
        ea.cellToProcMap = dC2P;
        ea.stencilMatrix = dSM;
        ea.matrixStart = dMS;
        ea.matrixCols = dMC;
        ea.matrixData = dMD;
        ea.B = dB;
        ea.intBasTrans = dIBT;
        ea.refFacAr = dRFA;
        ea.owner = P->owner;
        ea.neighbour = dNei;
        ea.cellFaceStart = dCFS;
        ea.cellFaces = dCF;
        ea.bfPatchType = P->bfPatchType;
        ea.bfNbrCell = dBfNbr;
        ea.haloOff = dHaloOff;
        ea.enumIdx = dEnum;
        ea.records = P->records;
        ea.cellHasHalo = dHasHalo;
        ea.cellExt = dCellExt;
        const int64_t threads = (int64_t)nC * 32;
        expand_records_kernel<<<(unsigned)((threads + 255) / 256), 256>>>(ea);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            setError(std::string("expand_records_kernel: ") + cudaGetErrorString(e));
            rc = WENOEXT_ERR_CUDA;
        }
    }
    if (!rc && P->recStride > 0) buildTmaMaps(P);
    // interior / boundary cell lists (overlap of the halo exchange)
    std::vector<int32_t> li, lb;
    if (!rc) {
        std::vector<uint8_t> hasHalo(nC);
        cudaMemcpy(hasHalo.data(), dHasHalo, (size_t)nC, cudaMemcpyDeviceToHost);
        for (int c = 0; c < nC; ++c) (hasHalo[c] ? lb : li).push_back(c);
        P->nInterior = (int32_t)li.size();
        P->nBoundaryCells = (int32_t)lb.size();
        if (P->nBoundaryCells > 0) {
            // one allocation, boundary list right behind the interior list, so that a single
            // launch can run over both (halo gate of the cell kernel)
            std::vector<int32_t> all(li);
            all.insert(all.end(), lb.begin(), lb.end());
            rc = rc ? rc : upload(&P->cellListInterior, all.data(), all.size());
            P->cellListBoundary = P->cellListInterior + li.size();
        }
    }
    // Slabs for the pipelined host-buffer path (serial plans): what a prefix of the cells
    // reads (cell values, fluxes) and which faces are final once it has been processed.
    if (!rc && P->nProcs == 1) {
        // (24 slabs: 128^3 end to end 4.30-4.34 ms per step against 4.40-4.45 with 32 and 4.38-4.45 with 16, two
        // boxes, round 2; ends on whole waves of tiles made no difference)
        int S = 24;
        if (const char* e = getenv("WENOEXT_E2E_SLABS")) S = std::max(1, atoi(e));
        S = std::min(S, 64);  // (tile counter slots of overlapping launches)
        if (!getenv("WENOEXT_E2E_SLABS"))
            while (S > 1 && nC < S * 32768) S /= 2;  // a slab should be worth its launch
        else if (nC < S * 8192)
            S = 1;
        if (S > 1) {
            std::vector<int32_t> ext((size_t)nC * 3);
            cudaMemcpy(ext.data(), dCellExt, ext.size() * sizeof(int32_t), cudaMemcpyDeviceToHost);
            // Tapered slabs: what cannot overlap is the H2D copy in front of the first slab and
            // the D2H copy behind the last one, so those two are small (weights 1, 2, 4, 8 ... 8,
            // 4, 2, 1); WENOEXT_E2E_TAPER=0: equal slabs
            std::vector<double> wgt(S, 8.0);
            const char* tp = getenv("WENOEXT_E2E_TAPER");
            if (S >= 6 && !(tp && atoi(tp) == 0))
                for (int k = 0; k < 3; ++k) wgt[k] = wgt[S - 1 - k] = (double)(1 << k);
            double tot = 0, run = 0;
            for (double w : wgt) tot += w;
            P->slabEnd.clear();
            for (int k = 0; k < S; ++k) {
                run += wgt[k];
                const int64_t e = (int64_t)((double)nC * run / tot);
                P->slabEnd.push_back(k == S - 1 ? nC : (int32_t)std::min<int64_t>(nC, (e + 31) / 32 * 32));
            }
            while (P->slabEnd.size() > 1 && P->slabEnd[P->slabEnd.size() - 2] >= nC) P->slabEnd.pop_back();
            S = (int)P->slabEnd.size();
            P->needVf.assign(S, 0);
            P->needFlux.assign(S, 0);
            P->doneFaces.assign(S, 0);
            int c = 0, mv = 0, mf = 0;
            for (int k = 0; k < S; ++k) {
                for (; c < P->slabEnd[k]; ++c) {
                    mv = std::max(mv, ext[(size_t)c * 3] + 1);
                    mf = std::max(mf, ext[(size_t)c * 3 + 2] + 1);
                }
                P->needVf[k] = std::min(nC, std::max(mv, P->slabEnd[k]));
                P->needFlux[k] = mf;
            }
            int lo = P->nInt;
            c = nC;
            for (int k = S - 1; k >= 0; --k) {
                for (; c > P->slabEnd[k]; --c) lo = std::min(lo, ext[(size_t)(c - 1) * 3 + 1]);
                P->doneFaces[k] = std::min(lo, P->nInt);
            }
            // cells whose internal faces are all final after slab k (a prefix of the cells): wenoext_div
            P->doneCells.assign(S, 0);
            int dc = 0;
            for (int k = 0; k < S; ++k) {
                const int lim = k == S - 1 ? P->nInt : P->doneFaces[k];
                for (; dc < nC; ++dc) {
                    bool ok = true;
                    for (int i = d->cellFaceStart[dc]; i < d->cellFaceStart[dc + 1] && ok; ++i)
                        ok = d->cellFaces[i] >= P->nInt || d->cellFaces[i] < lim;
                    if (!ok) break;
                }
                P->doneCells[k] = k == S - 1 ? nC : dc;
            }
        }
    }
    // Parallel plans: slabs over the interior cell list; faces of boundary cells are fixed up.
    if (!rc && P->nProcs > 1 && P->nBoundaryCells > 0) {
        int S = 8;
        if (const char* e = getenv("WENOEXT_E2E_SLABS")) S = std::max(1, atoi(e));
        const int nI = P->nInterior;
        if (nI < S * 8192) S = 1;
        if (S > 1) {
            std::vector<int32_t> ext((size_t)nC * 3);
            cudaMemcpy(ext.data(), dCellExt, ext.size() * sizeof(int32_t), cudaMemcpyDeviceToHost);
            const int per = ((nI + S - 1) / S + 31) / 32 * 32;
            P->slabEnd.clear();
            for (int k = 0; k < S; ++k) P->slabEnd.push_back(std::min(nI, (k + 1) * per));
            while (P->slabEnd.size() > 1 && P->slabEnd[P->slabEnd.size() - 2] >= nI) P->slabEnd.pop_back();
            S = (int)P->slabEnd.size();
            // what the items of slabs <= k read (cell values, fluxes): inputs can then arrive
            // slab by slab as in the serial path
            P->needVf.assign(S, 0);
            P->needFlux.assign(S, 0);
            {
                int jj = 0, mv = 0, mf = 0;
                for (int k = 0; k < S; ++k) {
                    for (; jj < P->slabEnd[k]; ++jj) {
                        const int cc = li[jj];
                        mv = std::max(mv, std::max(ext[(size_t)cc * 3], cc) + 1);
                        mf = std::max(mf, ext[(size_t)cc * 3 + 2] + 1);
                    }
                    P->needVf[k] = std::min(nC, mv);
                    P->needFlux[k] = mf;
                }
            }
            P->doneFaces.assign(S, 0);
            int lo = P->nInt, j = nI;
            for (int k = S - 1; k >= 0; --k) {
                for (; j > P->slabEnd[k]; --j) lo = std::min(lo, ext[(size_t)li[j - 1] * 3 + 1]);
                P->doneFaces[k] = std::min(lo, P->nInt);
            }
            for (int c : lb)
                for (int i = d->cellFaceStart[c]; i < d->cellFaceStart[c + 1]; ++i)
                    if (d->cellFaces[i] < P->nInt) P->fixFacesH.push_back(d->cellFaces[i]);
            std::sort(P->fixFacesH.begin(), P->fixFacesH.end());
            P->fixFacesH.erase(std::unique(P->fixFacesH.begin(), P->fixFacesH.end()), P->fixFacesH.end());
            if (!P->fixFacesH.empty())
                rc = rc ? rc : upload(&P->fixFacesD, P->fixFacesH.data(), P->fixFacesH.size());
        }
    }
    for (void* t : temps) cudaFree(t);
    if (dHasHalo) cudaFree(dHasHalo);
    if (dCellExt) cudaFree(dCellExt);
    if (rc) {
        wenoext_plan_destroy(P);
        return rc;
    }
    cudaEventCreateWithFlags(&P->evA, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&P->evB, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&P->evC, cudaEventDisableTiming);
    wenoext_plan_info& I = P->info;
    I.nCells = nC;
    I.nFaces = d->nFaces;
    I.nInternalFaces = d->nInternalFaces;
    I.nCoupledFaces = nCoupled;
    I.nLiveStencils = nLive;
    I.nHaloCells = P->nHaloTotal;
    I.nInteriorCells = P->nInterior;
    I.recordBytes = P->recordBytesTotal;
    I.maxRecordBytes = P->maxRecordBytes;
    I.algBytesFixed = algFixed;
    I.algBytesPerComp = (int64_t)8 * nC + 8 * nWritten;
    I.algFlopsPerComp = flopsPerComp;
    *out = P;
    return 0;
}

// 64-bit labels (WM_LABEL_SIZE=64 builds): narrow every label array into temporaries, refusing values
// that do not fit, and create the plan from the 32-bit descriptor. The lengths come from the CSR offsets
// of the descriptor itself, so the offsets are looked at before the arrays they index.
extern "C" int wenoext_plan_create_labels64(const wenoext_plan_desc64* d, int device, wenoext_plan** out) {
    if (!d || !out) {
        setError("desc64 or out is NULL");
        return WENOEXT_ERR_INVALID;
    }
    const char* bad = nullptr;
    auto fits = [](int64_t v) { return v >= INT32_MIN && v <= INT32_MAX; };
    auto scalar = [&](int64_t v, const char* name) {
        if (!fits(v) && !bad) bad = name;
        return (int32_t)v;
    };
    auto narrow = [&](const int64_t* src, int64_t n, const char* name, std::vector<int32_t>& dst) -> const int32_t* {
        if (!src || n <= 0) return nullptr;
        dst.resize((size_t)n);
        for (int64_t i = 0; i < n; ++i) {
            if (!fits(src[i])) {
                if (!bad) bad = name;
                return nullptr;
            }
            dst[(size_t)i] = (int32_t)src[i];
        }
        return dst.data();
    };
    wenoext_plan_desc n{};
    n.polOrder = scalar(d->polOrder, "polOrder");
    n.nDvt = scalar(d->nDvt, "nDvt");
    for (int k = 0; k < 3; ++k) n.dim[k] = scalar(d->dim[k], "dim");
    n.nCells = scalar(d->nCells, "nCells");
    n.nFaces = scalar(d->nFaces, "nFaces");
    n.nInternalFaces = scalar(d->nInternalFaces, "nInternalFaces");
    n.nPatches = scalar(d->nPatches, "nPatches");
    n.nUniqueMatrices = scalar(d->nUniqueMatrices, "nUniqueMatrices");
    n.nProcs = scalar(d->nProcs, "nProcs");
    n.myProc = scalar(d->myProc, "myProc");
    n.nStencils = d->nStencils;
    if (!bad && (n.nCells <= 0 || n.nFaces < n.nInternalFaces || n.nInternalFaces < 0 || n.nPatches < 0 ||
                 n.nStencils < 0 || n.nUniqueMatrices < 0 || n.nProcs < 0))
        bad = "sizes";
    if (!bad && (!d->cellFaceStart || !d->cellStencilStart || !d->stencilEntryStart)) bad = "CSR offsets (NULL)";
    if (bad) {
        setError(std::string("desc64: ") + bad + " out of range");
        return WENOEXT_ERR_INVALID;
    }
    std::vector<int32_t> owner, nei, cfs, cf, sid, c2p, sm, mc, spl, rpl, rhs, shc, pnc;
    std::vector<wenoext_patch> patches;
    n.owner = narrow(d->owner, n.nFaces, "owner", owner);
    n.neighbour = narrow(d->neighbour, n.nInternalFaces, "neighbour", nei);
    n.cellFaceStart = narrow(d->cellFaceStart, (int64_t)n.nCells + 1, "cellFaceStart", cfs);
    const int64_t nCF = n.cellFaceStart ? d->cellFaceStart[n.nCells] : 0;
    n.cellFaces = narrow(d->cellFaces, nCF, "cellFaces", cf);
    if (d->patches && n.nPatches > 0) {
        patches.resize((size_t)n.nPatches);
        for (int i = 0; i < n.nPatches; ++i) {
            patches[i].type = scalar(d->patches[i].type, "patches.type");
            patches[i].start = scalar(d->patches[i].start, "patches.start");
            patches[i].size = scalar(d->patches[i].size, "patches.size");
            patches[i].nbr = scalar(d->patches[i].nbr, "patches.nbr");
        }
        n.patches = patches.data();
    }
    n.cellStencilStart = d->cellStencilStart;
    n.stencilEntryStart = d->stencilEntryStart;
    const int64_t nEntries = d->stencilEntryStart[n.nStencils];
    if (nEntries < 0) bad = "stencilEntryStart";
    n.stencilsID = narrow(d->stencilsID, nEntries, "stencilsID", sid);
    n.cellToProcMap = narrow(d->cellToProcMap, nEntries, "cellToProcMap", c2p);
    n.stencilMatrix = narrow(d->stencilMatrix, n.nStencils, "stencilMatrix", sm);
    n.matrixStart = d->matrixStart;
    n.matrixCols = narrow(d->matrixCols, n.nUniqueMatrices, "matrixCols", mc);
    n.matrixData = d->matrixData;
    n.B = d->B;
    n.intBasTrans = d->intBasTrans;
    n.refFacAr = d->refFacAr;
    n.sendProcList = narrow(d->sendProcList, n.nProcs, "sendProcList", spl);
    n.receiveProcList = narrow(d->receiveProcList, n.nProcs, "receiveProcList", rpl);
    n.receiveHaloSize = narrow(d->receiveHaloSize, n.nProcs, "receiveHaloSize", rhs);
    n.sendHaloStart = d->sendHaloStart;
    if (d->sendHaloStart && d->sendHaloCellID && n.nProcs > 0)
        n.sendHaloCellID = narrow(d->sendHaloCellID, d->sendHaloStart[n.nProcs], "sendHaloCellID", shc);
    n.patchNbrCell = narrow(d->patchNbrCell, (int64_t)n.nFaces - n.nInternalFaces, "patchNbrCell", pnc);
    if (bad) {
        setError(std::string("desc64: ") + bad + " holds a value that does not fit 32 bits");
        return WENOEXT_ERR_INVALID;
    }
    return wenoext_plan_create(&n, device, out);
}

extern "C" int wenoext_plan_destroy(wenoext_plan* P) {
    if (!P) return 0;
    cudaSetDevice(P->device);
    if (P->h_fix) cudaFreeHost(P->h_fix);
    if (P->haloStatusH) cudaFreeHost(P->haloStatusH);
    if (P->haloStatusDev) cudaFree(P->haloStatusDev);
    if (P->stepFluxOwned) cudaFree(P->stepFluxOwned);
    void* ptrs[] = {P->neighbourD, P->fixFacesD, P->d_fix, P->cellFaceStartD, P->cellFacesSortedD, P->d_corr, P->d_bnd, P->d_vol, P->d_div,
                    P->tileCounter, P->haloFlags, P->pushDone, P->tmaMaps, P->records, P->recInfo, P->cellListInterior, P->owner,
                    P->bfPatchType, P->bfPartner, P->bfNbrCell, P->sendIds, P->vals, P->sendBuf,
                    P->recvPatch, P->thetaPatch, P->sendPatch, P->s_vf, P->s_flux, P->s_pnf, P->s_corr, P->s_fv};
    for (void* p : ptrs)
        if (p && p != P->valsIpc[0] && p != P->valsIpc[1]) cudaFree(p);
    for (auto& pp : P->pushPeers) {
        for (double* v : pp.vals)
            if (v) cudaIpcCloseMemHandle(v);
        if (pp.flags) cudaIpcCloseMemHandle(pp.flags);
        if (pp.recvPatch) cudaIpcCloseMemHandle(pp.recvPatch);
    }
    for (double* v : P->valsIpc)
        if (v) cudaFree(v);
    if (P->comm) {
        if (NcclApi* n = ncclApi()) n->CommDestroy(P->comm);
    }
    if (P->commStream) cudaStreamDestroy(P->commStream);
    if (P->sIn) cudaStreamDestroy(P->sIn);
    if (P->sAlt) cudaStreamDestroy(P->sAlt);
    if (P->sPack) cudaStreamDestroy(P->sPack);
    if (P->sDiv) cudaStreamDestroy(P->sDiv);
    for (cudaEvent_t e : P->evCopy) cudaEventDestroy(e);
    for (cudaEvent_t e : P->evDiv) cudaEventDestroy(e);
    if (P->sOut) cudaStreamDestroy(P->sOut);
    for (cudaEvent_t e : P->evSlab) cudaEventDestroy(e);
    cudaEvent_t evs[] = {P->evA, P->evB, P->evC};
    for (cudaEvent_t e : evs)
        if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : P->timeEvents) cudaEventDestroy(e);
    delete P;
    return 0;
}

// ---------------------------------------------------------------------------
// NCCL (dlopen)
// ---------------------------------------------------------------------------
namespace wenoext {
NcclApi* ncclApi() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.handle ? &api : nullptr;
    tried = true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
        api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) {
        setError(std::string("cannot dlopen libnccl: ") + dlerror());
        return nullptr;
    }
    auto sym = [&](const char* n) { return dlsym(api.handle, n); };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.Send = (decltype(api.Send))sym("ncclSend");
    api.Recv = (decltype(api.Recv))sym("ncclRecv");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    if (!api.GetUniqueId || !api.CommInitRank || !api.Send || !api.Recv || !api.GroupStart ||
        !api.GroupEnd) {
        setError("libnccl is missing required symbols");
        dlclose(api.handle);
        api.handle = nullptr;
        return nullptr;
    }
    return &api;
}
}  // namespace wenoext

#define NCCL_TRY(api, expr)                                                               \
    do {                                                                                  \
        int _r = (expr);                                                                  \
        if (_r != 0) {                                                                    \
            setError(std::string(#expr) + ": " +                                          \
                     ((api)->GetErrorString ? (api)->GetErrorString(_r) : "nccl error")); \
            return WENOEXT_ERR_NCCL;                                                      \
        }                                                                                 \
    } while (0)

extern "C" int wenoext_nccl_unique_id(char id[128]) {
    NcclApi* n = ncclApi();
    if (!n) return WENOEXT_ERR_NCCL;
    NCCL_TRY(n, n->GetUniqueId(id));
    return 0;
}

// One-sided halo push (plan_internal.h: p2p). Every rank exports its two value buffers and
// its flag array with CUDA IPC; handles and the tables a sender needs (where its values go in
// the receiver's buffer, which flag is its own) travel in one ncclAllGather. All ranks take
// the same decision: if any rank cannot take part, everybody stays with NCCL send/recv.
namespace {
struct P2PBlobHead {
    cudaIpcMemHandle_t hVals[2], hFlags, hRecvPatch;
    int32_t ok, nCells;
    int32_t patchOk;  // exchange #2 can be pushed too: at most one processor patch per neighbour
};
}  // namespace

static int setupHaloPush(wenoext_plan* P, NcclApi* n) {
    if (!n->AllGather) return 0;
    const int R = P->nProcs, me = P->myProc;
    // senders and receivers of this rank; the double buffering relies on a symmetric
    // relation (whoever sends to me also receives from me)
    std::vector<int32_t> slotOf(R, -1);
    int32_t ok = 1, nRecv = 0, nSend = 0;
    for (int p = 0; p < R; ++p) {
        const bool snd = P->sendProcList[p] != -1 && P->sendOff[p + 1] > P->sendOff[p];
        const bool rcv = P->receiveProcList[p] != -1 && P->haloOff[p + 1] > P->haloOff[p];
        if (snd != rcv || (snd && p == me)) ok = 0;
        if (rcv) slotOf[p] = nRecv++;
        if (snd) ++nSend;
    }
    if (nRecv > 32 || nSend > 32) ok = 0;
    const size_t valsN = ((size_t)P->nCells + P->nHaloTotal) * 4;
    bool allocOk = true;
    for (int i = 0; i < 2; ++i)
        allocOk = allocOk && cudaMalloc((void**)&P->valsIpc[i], std::max<size_t>(valsN, 1) * 8) == cudaSuccess;
    // flags: [0, 32) one per halo sender, [32, 64) one per processor-patch sender
    allocOk = allocOk && cudaMalloc((void**)&P->haloFlags, 64 * sizeof(int32_t)) == cudaSuccess;
    allocOk = allocOk && cudaMalloc((void**)&P->pushDone, 2 * sizeof(int32_t)) == cudaSuccess;
    // exchange #2: the receive buffer of the processor-patch values must be mappable by the peers, so
    // it is allocated here for the widest field type instead of lazily per call
    std::vector<int32_t> patchSlotOf(R, -1), patchStartOf(R, -1), patchSizeOf(R, 0);
    int32_t patchOk = P->cfgPatchPush ? 1 : 0, nPatchSend = 0;
    for (const auto& pt : P->patches) {
        if (pt.type != WENOEXT_PATCH_PROCESSOR || pt.size == 0) continue;
        if (pt.nbr < 0 || pt.nbr >= R || pt.nbr == me || patchSlotOf[pt.nbr] >= 0) {
            patchOk = 0;  // two patches towards one neighbour: keep the NCCL exchange
            continue;
        }
        patchSlotOf[pt.nbr] = nPatchSend++;
        patchStartOf[pt.nbr] = pt.start;
        patchSizeOf[pt.nbr] = pt.size;
    }
    if (nPatchSend > 32) patchOk = 0;
    const size_t patchN = (size_t)std::max(P->nBnd, 1) * wenoext_plan::kPatchMaxComp;
    double* recvPatchIpc = nullptr;
    if (allocOk && P->hasProcessor)
        allocOk = cudaMalloc((void**)&recvPatchIpc, patchN * 8) == cudaSuccess &&
                  cudaMemset(recvPatchIpc, 0, patchN * 8) == cudaSuccess;
    if (allocOk && !P->haloStatusDev) {
        allocOk = cudaMalloc((void**)&P->haloStatusDev, sizeof(int32_t)) == cudaSuccess &&
                  cudaMemset(P->haloStatusDev, 0, sizeof(int32_t)) == cudaSuccess;
    }
    if (allocOk && !P->haloStatusH) {
        allocOk = cudaHostAlloc((void**)&P->haloStatusH, sizeof(int32_t), cudaHostAllocMapped) == cudaSuccess;
        if (allocOk) {
            *P->haloStatusH = 0;
            allocOk = cudaHostGetDevicePointer((void**)&P->haloStatusD, P->haloStatusH, 0) == cudaSuccess;
        }
    }
    P2PBlobHead head{};
    if (allocOk) {
        cudaMemset(P->haloFlags, 0, 64 * sizeof(int32_t));
        cudaMemset(P->pushDone, 0, 2 * sizeof(int32_t));
        allocOk = cudaIpcGetMemHandle(&head.hVals[0], P->valsIpc[0]) == cudaSuccess &&
                  cudaIpcGetMemHandle(&head.hVals[1], P->valsIpc[1]) == cudaSuccess &&
                  cudaIpcGetMemHandle(&head.hFlags, P->haloFlags) == cudaSuccess;
        if (allocOk && recvPatchIpc)
            allocOk = cudaIpcGetMemHandle(&head.hRecvPatch, recvPatchIpc) == cudaSuccess;
    }
    if (!allocOk) {
        cudaGetLastError();
        ok = 0;
    }
    head.ok = ok;
    head.nCells = P->nCells;
    head.patchOk = patchOk;
    // blob: head | haloBase[R] (int64: where sender q's values start in my buffer) | slot[R]
    //       | patchOff[R] (int32: boundary-face index where sender q's patch values go, -1 none)
    //       | patchSize[R] | patchSlot[R]
    const size_t blob = sizeof(P2PBlobHead) + (size_t)R * 8 + (size_t)R * 4 * 4;
    const size_t blobPad = (blob + 15) / 16 * 16;
    std::vector<unsigned char> mine(blobPad, 0), all(blobPad * R, 0);
    memcpy(mine.data(), &head, sizeof(head));
    for (int q = 0; q < R; ++q) {
        const int64_t base = (int64_t)P->nCells + P->haloOff[q];
        memcpy(mine.data() + sizeof(head) + (size_t)q * 8, &base, 8);
        memcpy(mine.data() + sizeof(head) + (size_t)R * 8 + (size_t)q * 4, &slotOf[q], 4);
        const int32_t poff = patchStartOf[q] >= 0 ? patchStartOf[q] - P->nInt : -1;
        memcpy(mine.data() + sizeof(head) + (size_t)R * 12 + (size_t)q * 4, &poff, 4);
        memcpy(mine.data() + sizeof(head) + (size_t)R * 16 + (size_t)q * 4, &patchSizeOf[q], 4);
        memcpy(mine.data() + sizeof(head) + (size_t)R * 20 + (size_t)q * 4, &patchSlotOf[q], 4);
    }
    unsigned char *dMine = nullptr, *dAll = nullptr;
    CUDA_TRY(cudaMalloc((void**)&dMine, blobPad));
    CUDA_TRY(cudaMalloc((void**)&dAll, blobPad * R));
    auto gather = [&](const void* src, void* dst, size_t bytes) -> int {
        CUDA_TRY(cudaMemcpy(dMine, src, bytes, cudaMemcpyHostToDevice));
        NCCL_TRY(n, n->AllGather(dMine, dAll, bytes, /*ncclChar*/ 0, P->comm, P->commStream));
        CUDA_TRY(cudaStreamSynchronize(P->commStream));
        CUDA_TRY(cudaMemcpy(dst, dAll, bytes * R, cudaMemcpyDeviceToHost));
        return 0;
    };
    int rc = gather(mine.data(), all.data(), blobPad);
    bool use = rc == 0;
    auto headOf = [&](int p) {
        P2PBlobHead h;
        memcpy(&h, all.data() + (size_t)p * blobPad, sizeof(h));
        return h;
    };
    for (int p = 0; use && p < R; ++p) use = headOf(p).ok != 0;
    // exchange #2 by push only if every rank can, and the two sides of every patch agree on its size
    bool usePatch = use;
    for (int p = 0; usePatch && p < R; ++p) usePatch = headOf(p).patchOk != 0;
    for (int p = 0; usePatch && p < R; ++p) {
        if (patchSlotOf[p] < 0) continue;
        const unsigned char* tab = all.data() + (size_t)p * blobPad + sizeof(P2PBlobHead);
        int32_t theirOff, theirSize;
        memcpy(&theirOff, tab + (size_t)R * 12 + (size_t)me * 4, 4);
        memcpy(&theirSize, tab + (size_t)R * 16 + (size_t)me * 4, 4);
        if (theirOff < 0 || theirSize != patchSizeOf[p]) usePatch = false;
    }
    // map the receivers' buffers
    int32_t opened = use ? 1 : 0;
    if (use) {
        for (int p = 0; p < R && opened; ++p) {
            if (!(P->sendProcList[p] != -1 && P->sendOff[p + 1] > P->sendOff[p])) continue;
            const P2PBlobHead h = headOf(p);
            wenoext_plan::PushPeer pp;
            pp.proc = p;
            const unsigned char* tab = all.data() + (size_t)p * blobPad + sizeof(P2PBlobHead);
            memcpy(&pp.haloBase, tab + (size_t)me * 8, 8);
            memcpy(&pp.slot, tab + (size_t)R * 8 + (size_t)me * 4, 4);
            bool o = pp.slot >= 0;
            for (int i = 0; i < 2 && o; ++i)
                o = cudaIpcOpenMemHandle((void**)&pp.vals[i], h.hVals[i], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
            o = o && cudaIpcOpenMemHandle((void**)&pp.flags, h.hFlags, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
            if (o && usePatch && patchSlotOf[p] >= 0) {
                int32_t poff, pslot;
                memcpy(&poff, tab + (size_t)R * 12 + (size_t)me * 4, 4);
                memcpy(&pslot, tab + (size_t)R * 20 + (size_t)me * 4, 4);
                pp.patchOff = poff;
                pp.patchSlot = pslot;
                pp.patchStart = patchStartOf[p];
                pp.patchSize = patchSizeOf[p];
                o = cudaIpcOpenMemHandle((void**)&pp.recvPatch, h.hRecvPatch, cudaIpcMemLazyEnablePeerAccess) ==
                    cudaSuccess;
            }
            P->pushPeers.push_back(pp);  // (closed in plan_destroy, also when partly opened)
            if (!o) {
                cudaGetLastError();
                opened = 0;
            }
        }
    }
    // second agreement: every rank could map what it needs
    if (rc == 0) {
        std::vector<int32_t> st(4, opened), stAll((size_t)4 * R, 0);
        rc = gather(st.data(), stAll.data(), 16);
        for (int p = 0; p < R; ++p) use = use && stAll[(size_t)4 * p] != 0;
    }
    cudaFree(dMine);
    cudaFree(dAll);
    if (rc) return rc;
    if (!use) {
        for (auto& pp : P->pushPeers) {
            for (double*& v : pp.vals)
                if (v) cudaIpcCloseMemHandle(v), v = nullptr;
            if (pp.flags) cudaIpcCloseMemHandle(pp.flags), pp.flags = nullptr;
            if (pp.recvPatch) cudaIpcCloseMemHandle(pp.recvPatch), pp.recvPatch = nullptr;
        }
        P->pushPeers.clear();
        if (recvPatchIpc) cudaFree(recvPatchIpc);
        for (double*& v : P->valsIpc)
            if (v) cudaFree(v), v = nullptr;
        if (getenv("WENOEXT_VERBOSE")) fprintf(stderr, "[wenoext rank %d] halo push unavailable, using NCCL send/recv\n", me);
        return 0;
    }
    // every processor neighbour must also be a halo peer (it always is: face neighbours are in the
    // central stencil), otherwise its patch would have nobody to push to
    if (usePatch) {
        int mapped = 0;
        for (const auto& pp : P->pushPeers) mapped += pp.recvPatch != nullptr;
        if (mapped != nPatchSend) usePatch = false;
    }
    if (recvPatchIpc) {
        // the plan's patch buffers, sized once for the widest field type (prepareCall keeps them)
        if (P->recvPatch) cudaFree(P->recvPatch);
        if (P->thetaPatch) cudaFree(P->thetaPatch);
        if (P->sendPatch) cudaFree(P->sendPatch);
        P->recvPatch = recvPatchIpc;
        P->thetaPatch = P->sendPatch = nullptr;
        CUDA_TRY(cudaMalloc((void**)&P->thetaPatch, patchN * 8));
        CUDA_TRY(cudaMalloc((void**)&P->sendPatch, patchN * 8));
        CUDA_TRY(cudaMemset(P->thetaPatch, 0, patchN * 8));
        P->patchCap = patchN;
    }
    P->patchPush = usePatch && P->hasProcessor;
    P->nPatchSenders = P->patchPush ? nPatchSend : 0;
    P->nRecvPeers = nRecv;
    if (P->vals && P->vals != P->valsIpc[0] && P->vals != P->valsIpc[1]) cudaFree(P->vals);
    P->vals = P->valsIpc[0];
    P->valsCap = valsN;
    P->p2p = true;
    if (getenv("WENOEXT_VERBOSE"))
        fprintf(stderr, "[wenoext rank %d] halo push over peer memory: %d senders, %zu receivers; patch values %s\n",
                me, nRecv, P->pushPeers.size(), P->patchPush ? "pushed too" : "by ncclSend/ncclRecv");
    return 0;
}

extern "C" int wenoext_attach_nccl(wenoext_plan* P, const char id[128], int rank, int nRanks) {
    if (!P || !id) {
        setError("attach_nccl: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    if (rank != P->myProc || nRanks != P->nProcs) {
        setError("attach_nccl: rank/nRanks do not match the plan's myProc/nProcs");
        return WENOEXT_ERR_INVALID;
    }
    NcclApi* n = ncclApi();
    if (!n) return WENOEXT_ERR_NCCL;
    CUDA_TRY(cudaSetDevice(P->device));
    NcclId nid;
    memcpy(nid.internal, id, 128);
    NCCL_TRY(n, n->CommInitRank(&P->comm, nRanks, nid, rank));
    {
        // highest priority: when the exchange kernel and the persistent cell kernel become
        // ready together, the exchange gets its SMs first (the cell kernel leaves no room for
        // a second resident CTA, and late cell CTAs just take fewer tiles)
        int prLow = 0, prHigh = 0;
        CUDA_TRY(cudaDeviceGetStreamPriorityRange(&prLow, &prHigh));
        if (const char* e = getenv("WENOEXT_COMM_PRIO"))
            if (atoi(e) == 0) prHigh = prLow;
        CUDA_TRY(cudaStreamCreateWithPriority(&P->commStream, cudaStreamNonBlocking, prHigh));
    }
    if (P->cfgP2P) return setupHaloPush(P, n);
    return 0;
}

// ---------------------------------------------------------------------------
// small accessors
// ---------------------------------------------------------------------------
extern "C" const char* wenoext_last_error(void) { return g_lastError.c_str(); }

extern "C" int wenoext_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int wenoext_plan_get_info(const wenoext_plan* P, wenoext_plan_info* info) {
    if (!P || !info) {
        setError("plan_get_info: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    *info = P->info;
    return 0;
}

extern "C" int wenoext_set_params(wenoext_plan* P, double p, double dm, double epsilon) {
    if (!P || !(p >= 0)) {
        setError("set_params: bad argument");
        return WENOEXT_ERR_INVALID;
    }
    P->p = p;
    P->dm = dm;
    P->eps = epsilon;
    return 0;
}

extern "C" int wenoext_set_fit01_bug_compat(wenoext_plan* P, int on) {
    if (!P) return WENOEXT_ERR_INVALID;
    P->bugCompat = on ? 1 : 0;
    return 0;
}

extern "C" int wenoext_last_launch_count(const wenoext_plan* P) { return P ? P->lastLaunches : 0; }

extern "C" int wenoext_enable_kernel_timing(wenoext_plan* P, int on) {
    if (!P) return WENOEXT_ERR_INVALID;
    P->timing = on != 0;
    P->timeUsed = 0;
    return 0;
}

extern "C" int wenoext_kernel_time(wenoext_plan* P, double* msSum, int* nLaunches, int reset) {
    if (!P || !msSum) {
        setError("kernel_time: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    *msSum = 0.0;
    if (nLaunches) *nLaunches = (int)(P->timeUsed / 2);
    CUDA_TRY(cudaSetDevice(P->device));
    for (size_t i = 0; i + 1 < P->timeUsed; i += 2) {
        float ms = 0;
        CUDA_TRY(cudaEventSynchronize(P->timeEvents[i + 1]));
        CUDA_TRY(cudaEventElapsedTime(&ms, P->timeEvents[i], P->timeEvents[i + 1]));
        *msSum += (double)ms;
        if (getenv("WENOEXT_TIMING_VERBOSE") && i < 200) {  // developer aid
            float gap = 0;
            if (i >= 2) cudaEventElapsedTime(&gap, P->timeEvents[i - 1], P->timeEvents[i]);
            fprintf(stderr, "[wenoext rank %d] cell-kernel launch %zu: %.4f ms (gap before: %.4f ms)\n",
                    P->myProc, i / 2, ms, gap);
        }
    }
    if (reset) P->timeUsed = 0;
    return 0;
}

// ---------------------------------------------------------------------------
// correction(): three stream-ordered phases. With NCCL (one process per GPU)
// a rank runs them back to back; with in-process peers (wenoext_attach_local)
// the caller runs phase k on every rank before phase k+1.
//   phase 0  pack values [+ pack halo, post exchange #1, interior cells]
//   phase 1  [halo arrives] boundary cells (or all cells when serial)
//            [post exchange #2: coupled-patch values, swapData]
//   phase 2  coupled patches: received values, cyclic copy, patch limiter
// ---------------------------------------------------------------------------
// a halo_wait_kernel of an earlier or of this call gave up (WENOEXT_HALO_TIMEOUT_S): sticky
static int haloStatus(wenoext_plan* P) {
    if (!P->haloStatusH) return 0;
    const int32_t v = *reinterpret_cast<volatile int32_t*>(P->haloStatusH);
    if (v == 0) return 0;
    setError("halo exchange: the values of sender #" + std::to_string(v - 1) +
             " of this rank's sender list did not arrive within WENOEXT_HALO_TIMEOUT_S; results of this and "
             "later calls are invalid");
    return WENOEXT_ERR_TIMEOUT;
}

static int ensureBuf(double** p, size_t* cap, size_t n) {
    if (*cap >= n && *p) return 0;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    CUDA_TRY(cudaMalloc((void**)p, std::max<size_t>(n, 1) * sizeof(double)));
    *cap = n;
    return 0;
}

// gateItem >= 0: work items from gateItem on are boundary cells behind the halo gate
// (device_plan.h); followUp: second launch of such a pair, counters kept, gate known open
// `st`/`counterSlot`: launches of one call that may overlap (slabs on alternating streams) use
// their own stream and their own pair of tile counters
int wenoext::launchCells(wenoext_plan* P, const int32_t* list, int nWork, int workBase, int gateItem,
                         bool followUp, cudaStream_t st, int counterSlot) {
    if (nWork <= 0) return 0;
    wenoext_call c = P->call;
    if (st) c.st = st;
    CellKernelArgs a{};
    a.records = P->records;
    a.recInfo = P->recInfo;
    a.cellList = list;
    constexpr int kCounterSlots = 64;
    if (!P->tileCounter) CUDA_TRY(cudaMalloc((void**)&P->tileCounter, kCounterSlots * 2 * sizeof(int32_t)));
    int32_t* counter = P->tileCounter + 2 * (counterSlot % kCounterSlots);
    if (!followUp) CUDA_TRY(cudaMemsetAsync(counter, 0, 2 * sizeof(int32_t), c.st));
    a.tileCounter = counter;
    a.workBase = workBase;
    a.nWork = nWork;
    a.tmaMaps = P->tmaMaps;
    a.tmaZ = P->tmaZ;
    a.tmaW = P->tmaW;
    for (int p = 0; p < 18; ++p) {
        a.tmaBytes[p] = P->tmaBytes[p];
        a.tmaElemOff[p] = P->tmaElemOff[p];
    }
    a.gateItem = gateItem;
    a.abortFlag = P->haloStatusDev;  // set only by a timed-out halo_wait_kernel
    if (gateItem >= 0 && !followUp) {
        a.gateFlags = P->haloFlags;
        a.nGate = P->nRecvPeers;
        a.gateEpoch = P->haloEpoch;
    }
    a.K = P->K;
    a.nComp = c.nComp;
    a.c0 = c.c0;
    a.VS = c.VS;
    a.vals = P->vals;
    a.flux = c.flux;
    a.corr = c.corr;
    a.faceValue = c.fv;
    a.thetaPatch = (c.scheme == WENOEXT_SCHEME_UPWINDFIT01 && P->hasProcessor) ? P->thetaPatch : nullptr;
    a.nInternalFaces = P->nInt;
    a.scheme = c.scheme;
    a.bounded = c.bounded;
    a.bugCompat = P->bugCompat;
    a.pExp = (uint32_t)P->p;  // WENOCoeff.H:149: double -> unsigned
    a.dm = P->dm;
    a.eps = P->eps;
    a.maxN = P->maxN;
    a.nVals = (int32_t)(P->nCells + P->nHaloTotal);
    a.nFaces = P->nFaces;
#ifdef WENOEXT_KDEBUG
    if (!P->dbg) {
        CUDA_TRY(cudaMalloc((void**)&P->dbg, 64));
    }
    CUDA_TRY(cudaMemsetAsync(P->dbg, 0, 64, c.st));
    a.dbg = P->dbg;
#endif
    const int D = chunkCols(P->K);
    a.nChunks = std::max(1, (P->maxN + D - 1) / D);
    // Launch geometry. Every warp of the CTA runs its own TMA pipeline: two head
    // slots (this pass and the next), one tail slot and a ring of `ringSteps` chunk
    // steps, each for cellsPerWarp cells. Slot strides are 32 (mod 128) bytes so that
    // the cells of one pass, which read the same field at the same time, do not fall
    // into the same shared-memory banks. No scratch shared memory.
    auto bankStride = [](int64_t bytes) { return (int)((bytes + 95) / 128 * 128 + 32); };
    a.headSlotBytes = bankStride(P->maxHeadBytes);
    a.tailSlotBytes = bankStride(P->maxTailBytes);
    a.chunkSlotBytes = bankStride(chunkBytes(P->K, P->maxSt, (GATHER_AHEAD + 1) * D, 0));
    const size_t budget =
        (size_t)std::min(P->smemLimit, P->cfgSmemKB > 0 ? P->cfgSmemKB * 1024 : P->smemLimit) - 64;
    // lanes per cell: 6 (five cells per warp, tets and 2-D meshes), 8 (hexes) or 32
    int cpw = P->maxSt <= 6 ? 5 : (P->maxSt <= 8 ? 4 : 1);
    if (P->cfgCells > 0) cpw = P->cfgCells;
    if ((cpw != 5 && cpw != 4 && cpw != 1) || P->maxSt > 32 / cpw) {
        setError("unsupported stencil shape (more than 32 stencils per cell)");
        return WENOEXT_ERR_INVALID;
    }
    auto warpBytesFor = [&](int ring) {
        return (size_t)cpw * (2 * (size_t)a.headSlotBytes + a.tailSlotBytes +
                              (size_t)ring * a.chunkSlotBytes);
    };
    auto perWarp = [&](int ring) { return warpBytesFor(ring) + (size_t)(3 + ring) * 8; };
    // ring depth 3 (one step being consumed, two in flight), 2 or 1, never deeper than the number of steps
    // of a cell: the geometry with the most steps that can be in flight while their warp computes, i.e.
    // warps x min(ring, 2) (a ring of 1 refills only after it has consumed: nothing of the warp's own
    // stream overlaps its arithmetic; the third slot adds little once there are two). Ties go to the
    // deeper ring. Order 3 (K = 19): 7 warps x 2 instead of 8 x 1, 0.886 -> 0.983 of the roofline at 64^3
    // (profiles/r02_kernel_experiments.md, experiment 10).
    int ring = 0, nw = 0, best = 0;
    for (int r = std::min(3, a.nChunks); r >= 1; --r) {
        const int w = (int)std::min<size_t>(cellKernelMaxWarps(), budget / perWarp(r));
        if (w * std::min(r, 2) > best) {
            best = w * std::min(r, 2);
            nw = w;
            ring = r;
        }
    }
    if (P->cfgSlots > 0) {
        ring = std::min(P->cfgSlots, a.nChunks);
        nw = (int)std::min<size_t>(cellKernelMaxWarps(), budget / perWarp(ring));
    }
    if (P->cfgWarps > 0) nw = std::min(nw, P->cfgWarps);
    if (nw < 1) {
        setError("cell records do not fit the shared memory (polynomial order too high)");
        return WENOEXT_ERR_INVALID;
    }
    a.cellsPerWarp = cpw;
    a.nWarps = nw;
    a.ringSteps = ring;
    a.warpBytes = (int32_t)warpBytesFor(ring);
    const size_t smem = cellKernelSmemBytes(a);
    const int nTiles = ((followUp ? nWork - gateItem : nWork) + 31) / 32;
    int grid = std::min(P->cfgGrid > 0 ? P->cfgGrid : P->nSM, nTiles);
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    if (P->timing) {
        // one event pair per timed launch, summed by wenoext_kernel_time()
        if (P->timeUsed + 2 > P->timeEvents.size()) {
            cudaEvent_t e0, e1;
            CUDA_TRY(cudaEventCreate(&e0));
            CUDA_TRY(cudaEventCreate(&e1));
            P->timeEvents.push_back(e0);
            P->timeEvents.push_back(e1);
        }
        t0 = P->timeEvents[P->timeUsed];
        t1 = P->timeEvents[P->timeUsed + 1];
        P->timeUsed += 2;
        cudaEventRecord(t0, c.st);
    }
    if (launchCellKernel(a, c.CB, grid, smem, c.st)) {
        setError("cell kernel launch configuration rejected");
        return WENOEXT_ERR_CUDA;
    }
    if (t1) cudaEventRecord(t1, c.st);
#ifdef WENOEXT_KDEBUG
    {
        int32_t h[16];
        cudaError_t e = cudaMemcpyAsync(h, P->dbg, 64, cudaMemcpyDeviceToHost, c.st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c.st);
        fprintf(stderr, "[kdebug] sync=%s code=%d :", cudaGetErrorString(e), h[0]);
        for (int i = 1; i < 16; ++i) fprintf(stderr, " %d", h[i]);
        fprintf(stderr, "\n");
    }
#endif
    P->lastLaunches++;
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// exchange #1 as a one-sided push of this step's halo values into the receivers' buffers
static void pushHalo(wenoext_plan* P, cudaStream_t st) {
    if (P->pushPeers.empty()) return;
    const wenoext_call& c = P->call;
    HaloPushArgs a{};
    a.nPeers = (int32_t)P->pushPeers.size();
    a.VS = c.VS;
    a.epoch = P->haloEpoch;
    a.vals = P->vals;
    a.sendIds = P->sendIds;
    a.done = P->pushDone;
    for (int i = 0; i < a.nPeers; ++i) {
        const auto& pp = P->pushPeers[i];
        a.src0[i] = P->sendOff[pp.proc];
        a.off[i + 1] = a.off[i] + (P->sendOff[pp.proc + 1] - P->sendOff[pp.proc]);
        a.dst[i] = pp.vals[P->haloEpoch & 1] + pp.haloBase * c.VS;
        a.flag[i] = pp.flags + pp.slot;
    }
    launchHaloPush(a, st);
    P->lastLaunches++;
}

// exchange #2 (swapData, WENOUpwindFit.C:326-385) as a one-sided push: this rank's processor-patch
// values of `corr` (btsfUD: written by the owner cells where flux > 0, zero elsewhere) go straight into
// the neighbours' receive buffers over peer memory; the last block raises their patch flags. No
// rendezvous and no communication kernel on the compute stream.
static void pushPatches(wenoext_plan* P, cudaStream_t st) {
    const wenoext_call& c = P->call;
    HaloPushArgs a{};
    a.VS = c.nComp;
    a.epoch = P->haloEpoch;
    a.vals = c.corr;
    a.sendIds = nullptr;
    a.done = P->pushDone + 1;
    int n = 0;
    for (const auto& pp : P->pushPeers) {
        if (!pp.recvPatch || pp.patchSize == 0) continue;
        a.src0[n] = pp.patchStart;
        a.off[n + 1] = a.off[n] + pp.patchSize;
        a.dst[n] = pp.recvPatch + (size_t)pp.patchOff * c.nComp;
        a.flag[n] = pp.flags + 32 + pp.patchSlot;
        ++n;
    }
    a.nPeers = n;
    if (n == 0) return;
    launchHaloPush(a, st);
    P->lastLaunches++;
}

static int phase0(wenoext_plan* P, bool launchInterior = true) {
    const wenoext_call& c = P->call;
    const bool parallel = P->nProcs > 1;
    if (P->p2p) {
        // values of step k live in buffer k & 1: a sender may already push step k+1 while this
        // rank still reads step k (it cannot be further ahead: it needs this rank's step-k+1
        // halo, pushed after this rank's step k, to finish its own step k+1)
        ++P->haloEpoch;
        P->vals = P->valsIpc[P->haloEpoch & 1];
    }
    launchPackValues(c.vf, P->nCells, c.nComp, c.c0, c.CB, c.VS, P->vals, c.st);
    P->lastLaunches++;
    if (!parallel) return 0;
    P->mergedLaunch = false;
    if (P->p2p) {
        pushHalo(P, c.st);
        if (!launchInterior) return 0;
        // One launch for interior and boundary cells. The gathers are coherent loads ordered behind
        // the ld.acquire.sys on the halo flags; the launch is additionally kept to the case where the
        // halo region of `vals` starts on a 32-byte sector, so that no sector an interior gather
        // touched holds halo values at all.
        P->mergedLaunch = P->cfgMerge && P->nBoundaryCells > 0 && ((int64_t)P->nCells * c.VS) % 4 == 0;
        if (P->mergedLaunch)
            return launchCells(P, P->cellListInterior, P->nInterior + P->nBoundaryCells, 0, P->nInterior);
        return launchCells(P, P->cellListInterior, P->nInterior);
    }
    if (P->nSendTotal > 0) {
        launchPackHalo(P->vals, c.VS, P->sendIds, P->nSendTotal, P->sendBuf, c.st);
        P->lastLaunches++;
    }
    CUDA_TRY(cudaEventRecord(P->evA, c.st));
    if (P->comm) {
        NcclApi* n = ncclApi();
        CUDA_TRY(cudaStreamWaitEvent(P->commStream, P->evA, 0));
        NCCL_TRY(n, n->GroupStart());
        for (int p = 0; p < P->nProcs; ++p) {
            const int64_t ns = P->sendOff[p + 1] - P->sendOff[p];
            const int64_t nr = P->haloOff[p + 1] - P->haloOff[p];
            if (P->sendProcList[p] != -1 && ns > 0)
                NCCL_TRY(n, n->Send(P->sendBuf + P->sendOff[p] * c.VS, (size_t)ns * c.VS, 8, p,
                                    P->comm, P->commStream));
            if (P->receiveProcList[p] != -1 && nr > 0)
                NCCL_TRY(n, n->Recv(P->vals + ((size_t)P->nCells + P->haloOff[p]) * c.VS,
                                    (size_t)nr * c.VS, 8, p, P->comm, P->commStream));
        }
        NCCL_TRY(n, n->GroupEnd());
        CUDA_TRY(cudaEventRecord(P->evB, P->commStream));
    }
    // interior cells overlap the exchange (the reference waits first, WENOCoeff.C:295)
    if (!launchInterior) return 0;
    return launchCells(P, P->cellListInterior, P->nInterior);
}

static int phase1(wenoext_plan* P) {
    const wenoext_call& c = P->call;
    const bool parallel = P->nProcs > 1;
    if (!parallel) {
        if (int rc = launchCells(P, nullptr, P->nCells)) return rc;
    } else {
        if (P->p2p) {
            launchHaloWait(P->haloFlags, P->nRecvPeers, P->haloEpoch, c.st, P->haloStatusD, P->haloStatusDev);
            if (P->nRecvPeers > 0) P->lastLaunches++;
        } else if (P->comm) {
            CUDA_TRY(cudaStreamWaitEvent(c.st, P->evB, 0));
        } else {
            // in-process peers: pull my halo segments out of the owners' send buffers
            for (int p = 0; p < P->nProcs; ++p) {
                const int64_t nr = P->haloOff[p + 1] - P->haloOff[p];
                if (P->receiveProcList[p] == -1 || nr == 0) continue;
                wenoext_plan* Q = P->localPeers[p];
                CUDA_TRY(cudaStreamWaitEvent(c.st, Q->evA, 0));
                CUDA_TRY(cudaMemcpyAsync(P->vals + ((size_t)P->nCells + P->haloOff[p]) * c.VS,
                                         Q->sendBuf + Q->sendOff[P->myProc] * c.VS,
                                         (size_t)nr * c.VS * 8, cudaMemcpyDeviceToDevice, c.st));
            }
        }
        if (P->mergedLaunch) {
            // whatever the first launch left of the boundary cells (normally nothing: the
            // exchange lands long before the interior cells are done)
            if (int rc = launchCells(P, P->cellListInterior, P->nInterior + P->nBoundaryCells, 0,
                                     P->nInterior, true))
                return rc;
        } else if (int rc = launchCells(P, P->cellListBoundary, P->nBoundaryCells)) {
            return rc;
        }
        P->mergedLaunch = false;
    }
    if (P->hasProcessor && P->patchPush) {
        pushPatches(P, c.st);
    } else if (P->hasProcessor) {
        CUDA_TRY(cudaEventRecord(P->evC, c.st));
        if (P->comm) {
            // swapData (WENOUpwindFit.C:326-385): whole patch arrays, both directions
            NcclApi* n = ncclApi();
            NCCL_TRY(n, n->GroupStart());
            for (const auto& pt : P->patches) {
                if (pt.type != WENOEXT_PATCH_PROCESSOR || pt.size == 0) continue;
                NCCL_TRY(n, n->Send(c.corr + (size_t)pt.start * c.nComp, (size_t)pt.size * c.nComp, 8,
                                    pt.nbr, P->comm, c.st));
                NCCL_TRY(n, n->Recv(P->recvPatch + (size_t)(pt.start - P->nInt) * c.nComp,
                                    (size_t)pt.size * c.nComp, 8, pt.nbr, P->comm, c.st));
            }
            NCCL_TRY(n, n->GroupEnd());
        } else {
            // in-process peers: snapshot what swapData would send (btsfUD) before
            // any rank's phase 2 rewrites its patch values in place
            CUDA_TRY(cudaMemcpyAsync(P->sendPatch, c.corr + (size_t)P->nInt * c.nComp,
                                     (size_t)P->nBnd * c.nComp * 8, cudaMemcpyDeviceToDevice, c.st));
            CUDA_TRY(cudaEventRecord(P->evC, c.st));
        }
    }
    return 0;
}

int wenoext::phase2(wenoext_plan* P) {
    const wenoext_call& c = P->call;
    if (!P->hasCoupled) return 0;
    if (P->hasProcessor && !P->comm) {
        for (const auto& pt : P->patches) {
            if (pt.type != WENOEXT_PATCH_PROCESSOR || pt.size == 0) continue;
            wenoext_plan* Q = P->localPeers[pt.nbr];
            const wenoext_patch* qp = nullptr;
            for (const auto& x : Q->patches)
                if (x.type == WENOEXT_PATCH_PROCESSOR && x.nbr == P->myProc) qp = &x;
            if (!qp || qp->size != pt.size) {
                setError("local exchange: processor patches do not match");
                return WENOEXT_ERR_INVALID;
            }
            CUDA_TRY(cudaStreamWaitEvent(c.st, Q->evC, 0));
            CUDA_TRY(cudaMemcpyAsync(P->recvPatch + (size_t)(pt.start - P->nInt) * c.nComp,
                                     Q->sendPatch + (size_t)(qp->start - Q->nInt) * c.nComp,
                                     (size_t)pt.size * c.nComp * 8, cudaMemcpyDeviceToDevice, c.st));
        }
    }
    if (P->patchPush && P->nPatchSenders > 0) {
        // the neighbours' patch values of this step (they push on their own: cannot deadlock)
        launchHaloWait(P->haloFlags + 32, P->nPatchSenders, P->haloEpoch, c.st, P->haloStatusD, P->haloStatusDev);
        P->lastLaunches++;
    }
    PatchKernelArgs a{};
    a.nBnd = P->nBnd;
    a.nInternalFaces = P->nInt;
    a.nComp = c.nComp;
    a.bfPatchType = P->bfPatchType;
    a.bfPartner = P->bfPartner;
    a.bfNbrCell = P->bfNbrCell;
    a.owner = P->owner;
    a.flux = c.flux;
    a.vf = c.vf;
    a.vals = P->vals;
    a.VS = c.VS;
    a.c0 = c.c0;
    a.CB = c.CB;
    a.patchNbrField = c.pnf;
    a.recv = P->recvPatch;
    a.thetaPatch = P->thetaPatch;
    a.corr = c.corr;
    a.scheme = c.scheme;
    a.bounded = c.bounded;
    launchPatchKernel(a, c.st);
    P->lastLaunches++;
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// public per-step entry points
// ---------------------------------------------------------------------------
static bool usePipeline(const wenoext_plan* P, int nComp, bool wantFaceValue) {
    if (P->slabEnd.size() <= 1 || nComp > cellKernelMaxCB(P->K)) return false;
    return P->nProcs == 1 || (P->comm != nullptr && !wantFaceValue);
}

static int prepareCall(wenoext_plan* P, int scheme, int bounded, int nComp, const double* vf,
                       const double* flux, const double* pnf, double* corr, double* fv,
                       int memKind, void* stream) {
    if (!P || !vf || !corr) {
        setError("correction: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    if (!flux && !P->stepFlux) {
        setError("correction: faceFlux is NULL and no field was given to wenoext_set_face_flux");
        return WENOEXT_ERR_INVALID;
    }
    if (int rc = haloStatus(P)) return rc;
    if (nComp < 1 || (scheme != WENOEXT_SCHEME_UPWINDFIT && scheme != WENOEXT_SCHEME_UPWINDFIT01)) {
        setError("correction: bad scheme or nComp");
        return WENOEXT_ERR_INVALID;
    }
    if (P->nProcs > 1 && !P->comm && P->localPeers.empty()) {
        setError("correction: parallel plan without a communicator (wenoext_attach_nccl)");
        return WENOEXT_ERR_INVALID;
    }
    if (P->patchPush && nComp > wenoext_plan::kPatchMaxComp) {
        setError("correction: more than 9 components on a plan that pushes its patch values over peer memory");
        return WENOEXT_ERR_INVALID;
    }
    if (scheme == WENOEXT_SCHEME_UPWINDFIT && bounded && P->hasProcessor && !pnf && !P->hasNbrCell) {
        setError("correction: bounded limiter on processor patches needs patchNbrField or desc.patchNbrCell");
        return WENOEXT_ERR_INVALID;
    }
    CUDA_TRY(cudaSetDevice(P->device));
    cudaStream_t st = (cudaStream_t)stream;
    wenoext_call& c = P->call;
    c = wenoext_call();
    c.scheme = scheme;
    c.bounded = bounded ? 1 : 0;
    c.nComp = nComp;
    c.st = st;
    const size_t nV = (size_t)P->nCells * nComp, nF = (size_t)P->nFaces;
    const size_t nB = (size_t)P->nBnd * nComp, nO = nF * nComp, nFV = (size_t)P->nInt * nComp;
    if (memKind == WENOEXT_MEM_HOST) {
        if (int rc = ensureBuf(&P->s_vf, &P->s_vfCap, nV)) return rc;
        if (flux)
            if (int rc = ensureBuf(&P->s_flux, &P->s_fluxCap, nF)) return rc;
        if (int rc = ensureBuf(&P->s_corr, &P->s_corrCap, nO)) return rc;
        // (the pipelined path copies slab by slab, see pipelinedHostCall)
        const bool pipelined = usePipeline(P, nComp, fv != nullptr);
        if (!pipelined) {
            CUDA_TRY(cudaMemcpyAsync(P->s_vf, vf, nV * 8, cudaMemcpyHostToDevice, st));
            if (flux) CUDA_TRY(cudaMemcpyAsync(P->s_flux, flux, nF * 8, cudaMemcpyHostToDevice, st));
        }
        c.vf = P->s_vf;
        c.flux = flux ? P->s_flux : P->stepFlux;
        c.corr = P->s_corr;
        if (pnf && nB) {
            if (int rc = ensureBuf(&P->s_pnf, &P->s_pnfCap, nB)) return rc;
            if (!pipelined) CUDA_TRY(cudaMemcpyAsync(P->s_pnf, pnf, nB * 8, cudaMemcpyHostToDevice, st));
            c.pnf = P->s_pnf;
        }
        if (fv) {
            if (int rc = ensureBuf(&P->s_fv, &P->s_fvCap, nFV)) return rc;
            c.fv = P->s_fv;
        }
    } else {
        c.vf = vf;
        c.flux = flux ? flux : P->stepFlux;
        c.pnf = pnf;
        c.corr = corr;
        c.fv = fv;
    }
    if (int rc = ensureBuf(&P->vals, &P->valsCap, ((size_t)P->nCells + P->nHaloTotal) * 4)) return rc;
    if (P->nSendTotal)
        if (int rc = ensureBuf(&P->sendBuf, &P->sendCap, (size_t)P->nSendTotal * 4)) return rc;
    if (P->hasCoupled && P->patchCap < nB) {
        if (P->recvPatch) cudaFree(P->recvPatch);
        if (P->thetaPatch) cudaFree(P->thetaPatch);
        if (P->sendPatch) cudaFree(P->sendPatch);
        P->recvPatch = P->thetaPatch = P->sendPatch = nullptr;
        CUDA_TRY(cudaMalloc((void**)&P->sendPatch, nB * 8));
        CUDA_TRY(cudaMalloc((void**)&P->recvPatch, nB * 8));
        CUDA_TRY(cudaMalloc((void**)&P->thetaPatch, nB * 8));
        CUDA_TRY(cudaMemset(P->recvPatch, 0, nB * 8));
        CUDA_TRY(cudaMemset(P->thetaPatch, 0, nB * 8));
        P->patchCap = nB;
    }
    P->lastLaunches = 0;
    // boundaryField starts as pTraits::zero (WENOUpwindFit.C:230-246)
    if (nB) CUDA_TRY(cudaMemsetAsync(c.corr + (size_t)P->nInt * nComp, 0, nB * 8, st));
    return 0;
}

static void setBatch(wenoext_plan* P, int c0) {
    wenoext_call& c = P->call;
    c.c0 = c0;
    c.CB = std::min(cellKernelMaxCB(P->K), c.nComp - c0);
    c.VS = c.CB == 3 ? 4 : c.CB;
}

static int finishCall(wenoext_plan* P, double* corr, double* fv, int memKind) {
    if (memKind != WENOEXT_MEM_HOST || P->skipD2H) return 0;
    const wenoext_call& c = P->call;
    CUDA_TRY(cudaMemcpyAsync(corr, c.corr, (size_t)P->nFaces * c.nComp * 8, cudaMemcpyDeviceToHost, c.st));
    if (fv) CUDA_TRY(cudaMemcpyAsync(fv, c.fv, (size_t)P->nInt * c.nComp * 8, cudaMemcpyDeviceToHost, c.st));
    CUDA_TRY(cudaStreamSynchronize(c.st));
    return haloStatus(P);
}

// Host-buffer call on a serial plan, pipelined over slabs of cells: the H2D copy of
// what slab k+1 reads, the kernel of slab k and the D2H copy of the faces slab k-1
// finalised run concurrently on three streams (copy engines in both directions).
static int pipelinedHostCall(wenoext_plan* P, const double* vf, const double* flux, const double* pnf,
                             double* corr, double* fv) {
    wenoext_call& c = P->call;
    const int S = (int)P->slabEnd.size();
    const int nComp = c.nComp;
    if (!P->sIn) {
        CUDA_TRY(cudaStreamCreateWithFlags(&P->sIn, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&P->sOut, cudaStreamNonBlocking));
    }
    if (!P->sAlt) CUDA_TRY(cudaStreamCreateWithFlags(&P->sAlt, cudaStreamNonBlocking));
    // The pack of a slab's cell values is a kernel: while a slab kernel fills the SMs it gets a slot only now
    // and then. On the input stream it held up the NEXT slab's copies (inputs arrived every 0.205 ms per slab
    // of 106 K cells, the kernels need one every 0.197 ms: the kernels waited for their inputs). The packs
    // therefore have a stream of their own, of the highest priority, and the copy engine runs ahead.
    if (!P->sPack) {
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        CUDA_TRY(cudaStreamCreateWithPriority(&P->sPack, cudaStreamNonBlocking, hi));
    }
    while ((int)P->evSlab.size() < 2 * S + 2) {
        cudaEvent_t e;
        CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        P->evSlab.push_back(e);
    }
    while ((int)P->evCopy.size() < S) {
        cudaEvent_t e;
        CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        P->evCopy.push_back(e);
    }
    // the staging buffers may still be in use by the caller's stream
    cudaEvent_t evStart = P->evSlab[2 * S], evEnd = P->evSlab[2 * S + 1];
    CUDA_TRY(cudaEventRecord(evStart, c.st));
    CUDA_TRY(cudaStreamWaitEvent(P->sIn, evStart, 0));
    CUDA_TRY(cudaStreamWaitEvent(P->sOut, evStart, 0));
    CUDA_TRY(cudaStreamWaitEvent(P->sAlt, evStart, 0));
    double* dvf = const_cast<double*>(c.vf);
    double* dflux = const_cast<double*>(c.flux);
    if (pnf && P->nBnd)
        CUDA_TRY(cudaMemcpyAsync(const_cast<double*>(c.pnf), pnf, (size_t)P->nBnd * nComp * 8,
                                 cudaMemcpyHostToDevice, P->sIn));
    // Without coupled patches the boundary part of the result is final as prepareCall left it (zero,
    // WENOUpwindFit.C:230-246): it goes out first, under the kernels, instead of behind the last one
    int tailEnd = P->nFaces;
    if (!P->hasCoupled && !P->skipD2H && P->nFaces > P->nInt) {
        CUDA_TRY(cudaMemcpyAsync(corr + (size_t)P->nInt * nComp, c.corr + (size_t)P->nInt * nComp,
                                 (size_t)(P->nFaces - P->nInt) * nComp * 8, cudaMemcpyDeviceToHost, P->sOut));
        tailEnd = P->nInt;
    }
    // wenoext_div (see plan_internal.h: doneCells)
    const bool divPipe = P->skipD2H && P->divPipeOn && !P->hasCoupled && (int)P->doneCells.size() == S &&
                         !getenv("WENOEXT_DIV_PIPE_OFF");
    int divSent = 0;
    if (divPipe && !P->sDiv) {  // (not the pack stream: a pack queued behind a div kernel would wait for that slab's kernel)
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        CUDA_TRY(cudaStreamCreateWithPriority(&P->sDiv, cudaStreamNonBlocking, hi));
    }
    // the divergence of a boundary cell needs the flux on its boundary faces, which sit at the end of the flux
    // field and would otherwise arrive with the last slab only (the cell kernel never reads them on meshes
    // without coupled patches)
    if (divPipe && flux && P->nFaces > P->nInt)
        CUDA_TRY(cudaMemcpyAsync(const_cast<double*>(c.flux) + P->nInt, flux + P->nInt,
                                 (size_t)(P->nFaces - P->nInt) * 8, cudaMemcpyHostToDevice, P->sIn));
    while (divPipe && (int)P->evDiv.size() < S) {
        cudaEvent_t e;
        CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        P->evDiv.push_back(e);
    }
    int haveVf = 0, haveFlux = 0, sent = 0, cell0 = 0;
    // developer aid (WENOEXT_E2E_TRACE=1): device timeline of the call, per slab: inputs there, kernel done,
    // output copy done, in ms after the start of the call
    static const bool trace = getenv("WENOEXT_E2E_TRACE") != nullptr;
    std::vector<cudaEvent_t> tr;
    auto mark = [&](cudaStream_t st) {
        if (!trace) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, st);
        tr.push_back(e);
    };
    mark(c.st);
    for (int k = 0; k < S; ++k) {
        const int nv = P->needVf[k], nf = k == S - 1 ? P->nFaces : std::max(haveFlux, (int)P->needFlux[k]);
        if (nv > haveVf)
            CUDA_TRY(cudaMemcpyAsync(dvf + (size_t)haveVf * nComp, vf + (size_t)haveVf * nComp,
                                     (size_t)(nv - haveVf) * nComp * 8, cudaMemcpyHostToDevice, P->sIn));
        if (flux && nf > haveFlux)
            CUDA_TRY(cudaMemcpyAsync(dflux + haveFlux, flux + haveFlux, (size_t)(nf - haveFlux) * 8,
                                     cudaMemcpyHostToDevice, P->sIn));
        // the pack follows its copies on the pack stream (in slab order, so a slab's kernel depends on one
        // event only); the input stream goes straight on with the next slab's copies
        CUDA_TRY(cudaEventRecord(P->evCopy[k], P->sIn));
        CUDA_TRY(cudaStreamWaitEvent(P->sPack, P->evCopy[k], 0));
        if (nv > haveVf) {
            launchPackValues(c.vf + (size_t)haveVf * nComp, nv - haveVf, nComp, c.c0, c.CB, c.VS,
                             P->vals + (size_t)haveVf * c.VS, P->sPack);
            P->lastLaunches++;
        }
        CUDA_TRY(cudaEventRecord(P->evSlab[2 * k], P->sPack));
        mark(P->sPack);
        // slabs alternate between two streams: the CTAs of slab k+1 move onto the SMs as those
        // of slab k run out of tiles, instead of waiting for the whole grid to drain (slabs
        // write disjoint faces and only read the inputs)
        cudaStream_t ks = (k & 1) ? P->sAlt : c.st;
        CUDA_TRY(cudaStreamWaitEvent(ks, P->evSlab[2 * k], 0));
        haveVf = std::max(haveVf, nv);
        haveFlux = nf;
        if (int rc = launchCells(P, nullptr, P->slabEnd[k] - cell0, cell0, -1, false, ks, k)) return rc;
        cell0 = P->slabEnd[k];
        CUDA_TRY(cudaEventRecord(P->evSlab[2 * k + 1], ks));
        mark(ks);
        // (the output stream waits for every slab in order: doneFaces[k] counts on slabs <= k)
        CUDA_TRY(cudaStreamWaitEvent(P->sOut, P->evSlab[2 * k + 1], 0));
        if (k == S - 1 && S > 1) {  // join: what follows on the caller's stream needs all slabs
            CUDA_TRY(cudaStreamWaitEvent(c.st, P->evSlab[2 * k + 1], 0));
            CUDA_TRY(cudaStreamWaitEvent(c.st, P->evSlab[2 * (k - 1) + 1], 0));
        }
        const int done = k == S - 1 ? sent : P->doneFaces[k];  // the rest follows the patch kernel
        if (done > sent && !P->skipD2H) {
            CUDA_TRY(cudaMemcpyAsync(corr + (size_t)sent * nComp, c.corr + (size_t)sent * nComp,
                                     (size_t)(done - sent) * nComp * 8, cudaMemcpyDeviceToHost, P->sOut));
            if (fv)
                CUDA_TRY(cudaMemcpyAsync(fv + (size_t)sent * nComp, c.fv + (size_t)sent * nComp,
                                         (size_t)(done - sent) * nComp * 8, cudaMemcpyDeviceToHost,
                                         P->sOut));
            sent = done;
        }
        if (divPipe && P->doneCells[k] > divSent) {
            // the cells whose faces are final now: integrate them (pack stream: highest priority, a small
            // kernel that must not queue behind the slab kernels) and send the rows home
            wenoext::DivKernelArgs da = P->divPipeArgs;
            da.vf = c.vf;
            da.flux = c.flux;
            da.corr = c.corr;
            da.cell0 = divSent;
            da.nCells = P->doneCells[k] - divSent;
            CUDA_TRY(cudaStreamWaitEvent(P->sDiv, P->evSlab[2 * k], 0));      // inputs (and evStart behind them)
            CUDA_TRY(cudaStreamWaitEvent(P->sDiv, P->evSlab[2 * k + 1], 0));  // this slab's kernel
            if (k > 0) CUDA_TRY(cudaStreamWaitEvent(P->sDiv, P->evSlab[2 * (k - 1) + 1], 0));  // and the one on the other stream
            launchDivKernel(da, P->sDiv);
            P->lastLaunches++;
            CUDA_TRY(cudaEventRecord(P->evDiv[k], P->sDiv));
            CUDA_TRY(cudaStreamWaitEvent(P->sOut, P->evDiv[k], 0));
            CUDA_TRY(cudaMemcpyAsync(P->divPipeHost + (size_t)divSent * nComp, da.div + (size_t)divSent * nComp,
                                     (size_t)da.nCells * nComp * 8, cudaMemcpyDeviceToHost, P->sOut));
            divSent = P->doneCells[k];
        }
        mark(P->sOut);
    }
    if (int rc = phase2(P)) return rc;  // cyclic patches
    if (P->skipD2H) {  // wenoext_div: the face field stays on the device
        CUDA_TRY(cudaStreamSynchronize(P->sIn));
        if (divPipe) {  // ... and the divergence is already on its way home
            CUDA_TRY(cudaStreamSynchronize(P->sOut));
            CUDA_TRY(cudaStreamSynchronize(c.st));
            CUDA_TRY(cudaStreamSynchronize(P->sDiv));
            P->divPipeDone = true;
        }
        for (cudaEvent_t e : tr) cudaEventDestroy(e);
        return 0;
    }
    CUDA_TRY(cudaEventRecord(evEnd, c.st));
    CUDA_TRY(cudaStreamWaitEvent(P->sOut, evEnd, 0));
    if (tailEnd > sent)
        CUDA_TRY(cudaMemcpyAsync(corr + (size_t)sent * nComp, c.corr + (size_t)sent * nComp,
                                 (size_t)(tailEnd - sent) * nComp * 8, cudaMemcpyDeviceToHost, P->sOut));
    if (fv && P->nInt > sent)
        CUDA_TRY(cudaMemcpyAsync(fv + (size_t)sent * nComp, c.fv + (size_t)sent * nComp,
                                 (size_t)(P->nInt - sent) * nComp * 8, cudaMemcpyDeviceToHost, P->sOut));
    mark(P->sOut);
    CUDA_TRY(cudaStreamSynchronize(P->sOut));
    CUDA_TRY(cudaStreamSynchronize(c.st));
    if (trace) {
        auto at = [&](size_t i) {
            float ms = 0;
            cudaEventElapsedTime(&ms, tr[0], tr[i]);
            return ms;
        };
        fprintf(stderr, "[wenoext e2e trace] slab: cells, inputs there, kernel done, output copy done (ms)\n");
        for (int k = 0; k < S; ++k)
            fprintf(stderr, "  %2d %8d %7.3f %7.3f %7.3f\n", k, P->slabEnd[k] - (k ? P->slabEnd[k - 1] : 0),
                    at(1 + 3 * k), at(2 + 3 * k), at(3 + 3 * k));
        fprintf(stderr, "  last copy done %7.3f\n", at(tr.size() - 1));
        for (cudaEvent_t e : tr) cudaEventDestroy(e);
    }
    return 0;
}

// Tail of a pipelined host-buffer call on a parallel plan: the faces [sent, nFaces) follow in one
// copy; the internal faces below `sent` that boundary cells finalised after their range had been
// copied (fixFaces) are put in place last. If the caller's buffer is page-locked the device writes
// them straight into it (zero-copy stores ordered behind the bulk copies on the output stream);
// otherwise they travel compactly and the host scatters them.
static int finishParallelHostCall(wenoext_plan* P, double* corr, int sent, cudaEvent_t evEnd) {
    const wenoext_call& c = P->call;
    if (P->skipD2H) {  // wenoext_div: the face field stays on the device
        CUDA_TRY(cudaStreamSynchronize(P->sIn));
        return 0;
    }
    const int nComp = c.nComp;
    const size_t nFix = P->fixFacesH.size();
    double* mapped = nullptr;
    if (nFix && !getenv("WENOEXT_HOST_SCATTER")) {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, corr) == cudaSuccess && at.type == cudaMemoryTypeHost)
            mapped = static_cast<double*>(at.devicePointer);
        cudaGetLastError();
    }
    if (nFix && !mapped) {
        launchGatherRows(c.corr, P->fixFacesD, (int64_t)nFix, nComp, P->d_fix, c.st);
        P->lastLaunches++;
    }
    CUDA_TRY(cudaEventRecord(evEnd, c.st));
    CUDA_TRY(cudaStreamWaitEvent(P->sOut, evEnd, 0));
    if (nFix && !mapped)
        CUDA_TRY(cudaMemcpyAsync(P->h_fix, P->d_fix, nFix * nComp * 8, cudaMemcpyDeviceToHost, P->sOut));
    CUDA_TRY(cudaMemcpyAsync(corr + (size_t)sent * nComp, c.corr + (size_t)sent * nComp,
                             (size_t)(P->nFaces - sent) * nComp * 8, cudaMemcpyDeviceToHost, P->sOut));
    if (mapped) {
        launchScatterRows(c.corr, P->fixFacesD, (int64_t)nFix, nComp, mapped, P->sOut);
        P->lastLaunches++;
    }
    CUDA_TRY(cudaStreamSynchronize(P->sOut));
    CUDA_TRY(cudaStreamSynchronize(c.st));
    if (nFix && !mapped)
        for (size_t i = 0; i < nFix; ++i) {
            const size_t f = (size_t)P->fixFacesH[i];
            for (int k = 0; k < nComp; ++k) corr[f * nComp + k] = P->h_fix[i * nComp + k];
        }
    return 0;
}

// The same for a parallel plan (one process per GPU, NCCL): inputs are copied up front
// (the halo pack needs all cell values), the interior cell list is processed in slabs with
// the D2H copy of the faces they finalise overlapped, then boundary cells and coupled
// patches as usual. Internal faces written late by boundary cells travel in a compact
// buffer and are scattered into the result on the host.
static int pipelinedHostCallParallel(wenoext_plan* P, const double* vf, const double* flux,
                                     const double* pnf, double* corr) {
    wenoext_call& c = P->call;
    const int S = (int)P->slabEnd.size();
    const int nComp = c.nComp;
    if (!P->sIn) {
        CUDA_TRY(cudaStreamCreateWithFlags(&P->sIn, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&P->sOut, cudaStreamNonBlocking));
    }
    while ((int)P->evSlab.size() < 2 * S + 2) {
        cudaEvent_t e;
        CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        P->evSlab.push_back(e);
    }
    const size_t nFix = P->fixFacesH.size();
    if (nFix) {
        if (int rc = ensureBuf(&P->d_fix, &P->d_fixCap, nFix * nComp)) return rc;
        if (P->h_fixCap < nFix * nComp) {
            if (P->h_fix) cudaFreeHost(P->h_fix);
            P->h_fix = nullptr;
            CUDA_TRY(cudaMallocHost((void**)&P->h_fix, nFix * nComp * 8));
            P->h_fixCap = nFix * nComp;
        }
    }
    cudaEvent_t evStart = P->evSlab[2 * S], evEnd = P->evSlab[2 * S + 1];
    CUDA_TRY(cudaEventRecord(evStart, c.st));
    CUDA_TRY(cudaStreamWaitEvent(P->sIn, evStart, 0));
    CUDA_TRY(cudaStreamWaitEvent(P->sOut, evStart, 0));
    if (P->p2p && (int)P->needVf.size() == S && !getenv("WENOEXT_E2E_UPFRONT")) {
        // With the one-sided halo push nothing on this rank waits for a rendezvous, so the
        // inputs can arrive slab by slab as in the serial path: copy + pack on the input stream,
        // the slab kernels on two alternating streams, and the push right behind the piece
        // that completes the cell values (the peers need it only for their boundary cells).
        if (!P->sAlt) CUDA_TRY(cudaStreamCreateWithFlags(&P->sAlt, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamWaitEvent(P->sAlt, evStart, 0));
        // packs (and the halo push behind the last of them) on their own stream of the highest priority, so that
        // the input copies run ahead of the slab kernels (see pipelinedHostCall)
        if (!P->sPack) {
            int lo = 0, hi = 0;
            cudaDeviceGetStreamPriorityRange(&lo, &hi);
            CUDA_TRY(cudaStreamCreateWithPriority(&P->sPack, cudaStreamNonBlocking, hi));
        }
        while ((int)P->evCopy.size() < S) {
            cudaEvent_t e;
            CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            P->evCopy.push_back(e);
        }
        ++P->haloEpoch;
        P->vals = P->valsIpc[P->haloEpoch & 1];
        P->mergedLaunch = false;
        double* dvf = const_cast<double*>(c.vf);
        double* dflux = const_cast<double*>(c.flux);
        if (pnf && P->nBnd)
            CUDA_TRY(cudaMemcpyAsync(const_cast<double*>(c.pnf), pnf, (size_t)P->nBnd * nComp * 8,
                                     cudaMemcpyHostToDevice, P->sIn));
        int haveVf = 0, haveFlux = 0, sent = 0, item0 = 0;
        bool pushed = false;
        for (int k = 0; k < S; ++k) {
            const int nv = k == S - 1 ? P->nCells : P->needVf[k];
            const int nf = k == S - 1 ? P->nFaces : std::max(haveFlux, (int)P->needFlux[k]);
            const int packFrom = haveVf;
            if (nv > haveVf)
                CUDA_TRY(cudaMemcpyAsync(dvf + (size_t)haveVf * nComp, vf + (size_t)haveVf * nComp,
                                         (size_t)(nv - haveVf) * nComp * 8, cudaMemcpyHostToDevice, P->sIn));
            if (flux && nf > haveFlux)
                CUDA_TRY(cudaMemcpyAsync(dflux + haveFlux, flux + haveFlux, (size_t)(nf - haveFlux) * 8,
                                         cudaMemcpyHostToDevice, P->sIn));
            haveFlux = nf;
            CUDA_TRY(cudaEventRecord(P->evCopy[k], P->sIn));
            CUDA_TRY(cudaStreamWaitEvent(P->sPack, P->evCopy[k], 0));
            if (nv > packFrom) {
                launchPackValues(c.vf + (size_t)packFrom * nComp, nv - packFrom, nComp, c.c0, c.CB, c.VS,
                                 P->vals + (size_t)packFrom * c.VS, P->sPack);
                P->lastLaunches++;
                haveVf = nv;
            }
            if (!pushed && haveVf == P->nCells) {
                pushHalo(P, P->sPack);
                pushed = true;
            }
            CUDA_TRY(cudaEventRecord(P->evSlab[2 * k], P->sPack));
            cudaStream_t ks = (k & 1) ? P->sAlt : c.st;
            CUDA_TRY(cudaStreamWaitEvent(ks, P->evSlab[2 * k], 0));
            if (int rc = launchCells(P, P->cellListInterior, P->slabEnd[k] - item0, item0, -1, false, ks, k))
                return rc;
            item0 = P->slabEnd[k];
            CUDA_TRY(cudaEventRecord(P->evSlab[2 * k + 1], ks));
            CUDA_TRY(cudaStreamWaitEvent(P->sOut, P->evSlab[2 * k + 1], 0));
            if (k == S - 1) {  // join
                CUDA_TRY(cudaStreamWaitEvent(c.st, P->evSlab[2 * k + 1], 0));
                if (S > 1) CUDA_TRY(cudaStreamWaitEvent(c.st, P->evSlab[2 * (k - 1) + 1], 0));
            }
            const int done = P->doneFaces[k];
            if (k < S - 1 && done > sent && !P->skipD2H) {
                CUDA_TRY(cudaMemcpyAsync(corr + (size_t)sent * nComp, c.corr + (size_t)sent * nComp,
                                         (size_t)(done - sent) * nComp * 8, cudaMemcpyDeviceToHost, P->sOut));
                sent = done;
            }
        }
        if (int rc = phase1(P)) return rc;  // halo wait, boundary cells, exchange #2
        if (int rc = phase2(P)) return rc;  // coupled patches
        return finishParallelHostCall(P, corr, sent, evEnd);
    }
    CUDA_TRY(cudaMemcpyAsync(const_cast<double*>(c.vf), vf, (size_t)P->nCells * nComp * 8,
                             cudaMemcpyHostToDevice, P->sIn));
    if (flux)
        CUDA_TRY(cudaMemcpyAsync(const_cast<double*>(c.flux), flux, (size_t)P->nFaces * 8,
                                 cudaMemcpyHostToDevice, P->sIn));
    if (pnf && P->nBnd)
        CUDA_TRY(cudaMemcpyAsync(const_cast<double*>(c.pnf), pnf, (size_t)P->nBnd * nComp * 8,
                                 cudaMemcpyHostToDevice, P->sIn));
    CUDA_TRY(cudaEventRecord(P->evSlab[0], P->sIn));
    CUDA_TRY(cudaStreamWaitEvent(c.st, P->evSlab[0], 0));
    if (int rc = phase0(P, false)) return rc;  // pack, exchange #1 on the communication stream
    int sent = 0, item0 = 0;
    for (int k = 0; k < S; ++k) {
        if (int rc = launchCells(P, P->cellListInterior, P->slabEnd[k] - item0, item0)) return rc;
        item0 = P->slabEnd[k];
        CUDA_TRY(cudaEventRecord(P->evSlab[2 * k + 1], c.st));
        const int done = P->doneFaces[k];
        if (k < S - 1 && done > sent && !P->skipD2H) {
            CUDA_TRY(cudaStreamWaitEvent(P->sOut, P->evSlab[2 * k + 1], 0));
            CUDA_TRY(cudaMemcpyAsync(corr + (size_t)sent * nComp, c.corr + (size_t)sent * nComp,
                                     (size_t)(done - sent) * nComp * 8, cudaMemcpyDeviceToHost, P->sOut));
            sent = done;
        }
    }
    if (int rc = phase1(P)) return rc;  // boundary cells, exchange #2
    if (int rc = phase2(P)) return rc;  // coupled patches
    return finishParallelHostCall(P, corr, sent, evEnd);
}

extern "C" int wenoext_host_alloc(uint64_t bytes, void** out) {
    if (!out) {
        setError("host_alloc: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    *out = nullptr;
    CUDA_TRY(cudaHostAlloc(out, std::max<uint64_t>(bytes, 1), cudaHostAllocPortable));
    return 0;
}
extern "C" int wenoext_host_free(void* p) {
    if (p) CUDA_TRY(cudaFreeHost(p));
    return 0;
}

extern "C" int wenoext_set_face_flux(wenoext_plan* P, const double* faceFlux, int memKind, void* stream) {
    if (!P) {
        setError("set_face_flux: NULL plan");
        return WENOEXT_ERR_INVALID;
    }
    if (!faceFlux) {  // forget the field: later calls must pass faceFlux again
        P->stepFlux = nullptr;
        return 0;
    }
    CUDA_TRY(cudaSetDevice(P->device));
    if (memKind == WENOEXT_MEM_HOST) {
        if (int rc = ensureBuf(&P->stepFluxOwned, &P->stepFluxCap, (size_t)P->nFaces)) return rc;
        CUDA_TRY(cudaMemcpyAsync(P->stepFluxOwned, faceFlux, (size_t)P->nFaces * 8, cudaMemcpyHostToDevice,
                                 (cudaStream_t)stream));
        // host buffer: the caller may reuse it as soon as this returns
        CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
        P->stepFlux = P->stepFluxOwned;
    } else {
        P->stepFlux = faceFlux;
    }
    return 0;
}

extern "C" int wenoext_interpolate(wenoext_plan* P, int scheme, int bounded, int nComp,
                                   const double* vf, const double* faceFlux,
                                   const double* patchNbrField, double* corr, double* faceValue,
                                   int memKind, void* stream) {
    if (P && P->nProcs > 1 && !P->comm) {
        setError("correction: in-process peers need the split-phase calls (wenoext_correction_begin/...)");
        return WENOEXT_ERR_INVALID;
    }
    if (int rc = prepareCall(P, scheme, bounded, nComp, vf, faceFlux, patchNbrField, corr, faceValue,
                             memKind, stream))
        return rc;
    if (memKind == WENOEXT_MEM_HOST && usePipeline(P, nComp, faceValue != nullptr)) {
        setBatch(P, 0);
        if (P->nProcs == 1) return pipelinedHostCall(P, vf, faceFlux, patchNbrField, corr, faceValue);
        const int rc = pipelinedHostCallParallel(P, vf, faceFlux, patchNbrField, corr);
        return rc ? rc : haloStatus(P);
    }
    for (int c0 = 0; c0 < nComp; c0 += cellKernelMaxCB(P->K)) {
        setBatch(P, c0);
        if (int rc = phase0(P)) return rc;
        if (int rc = phase1(P)) return rc;
        if (int rc = phase2(P)) return rc;
    }
    return finishCall(P, corr, faceValue, memKind);
}

extern "C" int wenoext_correction(wenoext_plan* P, int scheme, int bounded, int nComp,
                                  const double* vf, const double* faceFlux,
                                  const double* patchNbrField, double* corr, int memKind,
                                  void* stream) {
    return wenoext_interpolate(P, scheme, bounded, nComp, vf, faceFlux, patchNbrField, corr, nullptr,
                               memKind, stream);
}

extern "C" int wenoext_div(wenoext_plan* P, int scheme, int bounded, int nComp, const double* vf,
                           const double* faceFlux, const double* patchNbrField, const double* bndValue,
                           const double* cellVolumes, double* div, int memKind, void* stream) {
    if (!P || !vf || !div || (P->nBnd > 0 && !bndValue)) {
        setError("div: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    if (!cellVolumes && !(memKind == WENOEXT_MEM_HOST && P->volValid)) {
        setError("div: cellVolumes is NULL and no earlier host-buffer call left the volumes on the device");
        return WENOEXT_ERR_INVALID;
    }
    if (!faceFlux && !P->stepFlux) {
        setError("div: faceFlux is NULL and no field was given to wenoext_set_face_flux");
        return WENOEXT_ERR_INVALID;
    }
    if (P->nProcs > 1 && !P->comm) {
        setError("div: in-process peers are not supported (attach NCCL)");
        return WENOEXT_ERR_INVALID;
    }
    if (P->hasProcessor && !patchNbrField && (!P->hasNbrCell || nComp > cellKernelMaxCB(P->K))) {
        setError("div: processor patches need patchNbrField (or desc.patchNbrCell with nComp <= 3)");
        return WENOEXT_ERR_INVALID;
    }
    CUDA_TRY(cudaSetDevice(P->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t nV = (size_t)P->nCells * nComp, nF = (size_t)P->nFaces, nB = (size_t)P->nBnd * nComp;
    const double *dvf = vf, *dflux = faceFlux ? faceFlux : P->stepFlux, *dpnf = patchNbrField, *dbnd = bndValue,
                 *dvol = cellVolumes, *dcorr = nullptr;
    double* ddiv = div;
    if (memKind == WENOEXT_MEM_HOST) {
        // The reconstruction is the ordinary host-buffer correction() (slab-pipelined where the plan
        // allows it: H2D of the next slab under the kernel of the current one) whose face field stays on
        // the device; only [nCells][nComp] comes back.
        if (int rc = ensureBuf(&P->d_vol, &P->d_volCap, (size_t)P->nCells)) return rc;
        if (int rc = ensureBuf(&P->d_div, &P->d_divCap, nV)) return rc;
        if (cellVolumes) {  // mesh.V() is constant: NULL on later calls reuses this copy
            CUDA_TRY(cudaMemcpyAsync(P->d_vol, cellVolumes, (size_t)P->nCells * 8, cudaMemcpyHostToDevice, st));
            P->volValid = true;
        }
        if (nB) {
            if (int rc = ensureBuf(&P->d_bnd, &P->d_bndCap, nB)) return rc;
            CUDA_TRY(cudaMemcpyAsync(P->d_bnd, bndValue, nB * 8, cudaMemcpyHostToDevice, st));
            dbnd = P->d_bnd;
        }
        {   // what the pipelined path needs to integrate cells as their faces become final
            DivKernelArgs& pa = P->divPipeArgs;
            pa = DivKernelArgs{};
            pa.nInternalFaces = P->nInt;
            pa.nComp = nComp;
            pa.cellFaceStart = P->cellFaceStartD;
            pa.cellFaces = P->cellFacesSortedD;
            pa.owner = P->owner;
            pa.neighbour = P->neighbourD;
            pa.bfPatchType = P->bfPatchType;
            pa.bfNbrCell = P->bfNbrCell;
            pa.bndValue = dbnd;
            pa.cellVolumes = P->d_vol;
            pa.div = P->d_div;
            P->divPipeHost = div;
            P->divPipeOn = P->nProcs == 1 && nComp <= cellKernelMaxCB(P->K);
            P->divPipeDone = false;
        }
        P->skipD2H = true;
        const int rc = wenoext_interpolate(P, scheme, bounded, nComp, vf, faceFlux, patchNbrField, div /*unused*/,
                                           nullptr, WENOEXT_MEM_HOST, stream);
        P->skipD2H = false;
        P->divPipeOn = false;
        if (rc) return rc;
        if (P->divPipeDone) return haloStatus(P);
        dvf = P->s_vf;
        dflux = faceFlux ? P->s_flux : P->stepFlux;
        dpnf = patchNbrField ? P->s_pnf : nullptr;
        dvol = P->d_vol;
        ddiv = P->d_div;
        dcorr = P->s_corr;
    } else {
        if (int rc = ensureBuf(&P->d_corr, &P->d_corrCap, nF * nComp)) return rc;
        if (int rc = wenoext_interpolate(P, scheme, bounded, nComp, dvf, dflux, dpnf, P->d_corr, nullptr,
                                         WENOEXT_MEM_DEVICE, stream))
            return rc;
        dcorr = P->d_corr;
    }
    DivKernelArgs a{};
    a.nCells = P->nCells;
    a.nInternalFaces = P->nInt;
    a.nComp = nComp;
    a.cellFaceStart = P->cellFaceStartD;
    a.cellFaces = P->cellFacesSortedD;
    a.owner = P->owner;
    a.neighbour = P->neighbourD;
    a.bfPatchType = P->bfPatchType;
    a.bfNbrCell = P->bfNbrCell;
    a.vf = dvf;
    a.vals = P->vals;
    a.VS = P->call.VS;
    a.c0 = P->call.c0;
    a.CB = P->call.CB;
    a.flux = dflux;
    a.corr = dcorr;
    a.patchNbrField = dpnf;
    a.bndValue = dbnd;
    a.cellVolumes = dvol;
    a.div = ddiv;
    launchDivKernel(a, st);
    P->lastLaunches++;
    CUDA_TRY(cudaGetLastError());
    if (memKind == WENOEXT_MEM_HOST) {
        CUDA_TRY(cudaMemcpyAsync(div, ddiv, nV * 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        return haloStatus(P);
    }
    return 0;
}

// In-process ranks (tests / single-process multi-GPU): plans[r]->myProc == r.
extern "C" int wenoext_attach_local(wenoext_plan** plans, int nRanks) {
    if (!plans || nRanks < 1) return WENOEXT_ERR_INVALID;
    for (int r = 0; r < nRanks; ++r)
        if (!plans[r] || plans[r]->myProc != r || plans[r]->nProcs != nRanks) {
            setError("attach_local: plans must be ordered by rank and share nProcs");
            return WENOEXT_ERR_INVALID;
        }
    for (int r = 0; r < nRanks; ++r) plans[r]->localPeers.assign(plans, plans + nRanks);
    return 0;
}

// Split-phase form of wenoext_correction for in-process ranks (device pointers,
// nComp <= 3): call _begin on every rank, then _exchange on every rank, then _end.
extern "C" int wenoext_correction_begin(wenoext_plan* P, int scheme, int bounded, int nComp,
                                        const double* vf, const double* faceFlux,
                                        const double* patchNbrField, double* corr, void* stream) {
    if (!P) {
        setError("correction_begin: NULL plan");
        return WENOEXT_ERR_INVALID;
    }
    if (nComp > cellKernelMaxCB(P->K)) {
        setError("split-phase correction supports one component batch (nComp <= 3, 1 at order 4 in 3-D)");
        return WENOEXT_ERR_INVALID;
    }
    if (int rc = prepareCall(P, scheme, bounded, nComp, vf, faceFlux, patchNbrField, corr, nullptr,
                             WENOEXT_MEM_DEVICE, stream))
        return rc;
    setBatch(P, 0);
    return phase0(P);
}
extern "C" int wenoext_correction_exchange(wenoext_plan* P) {
    if (!P) return WENOEXT_ERR_INVALID;
    CUDA_TRY(cudaSetDevice(P->device));
    return phase1(P);
}
extern "C" int wenoext_correction_end(wenoext_plan* P) {
    if (!P) return WENOEXT_ERR_INVALID;
    CUDA_TRY(cudaSetDevice(P->device));
    return phase2(P);
}
