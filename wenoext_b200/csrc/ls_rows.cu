// Device side of the WENOBase setup: the least-squares matrices of
// WENOBase::calcMatrix (WENOBase.C:410-597) assembled on the GPU and handed to the batched
// pseudo-inverse (pinv.cu) without leaving the device.
//
//   ls_rows_kernel   one thread per (target cell I, stencil cell J) pair: the volume moments of cell J
//                    in the reference frame of cell I (geometryWENO::transformIntegral,
//                    geometryWENO.C:222-325: divergence theorem over the triangulated surface of J,
//                    13-point Gauss rule per triangle, geometryWENO.C:346-460), then the row of the
//                    LS matrix (WENOBase::calcGeom / addCoeffs, WENOBase.C:600-670)
//   ls_gather_kernel A_s[r][:] = row of (I, stencil entry r+1): rows are shared by the stencils of a cell
//
// The arithmetic repeats the host restatement (setup/geometry_weno.cpp, setup/weno_setup.cpp) operation
// for operation, and this file is compiled with --fmad=false: the host library is built for baseline
// x86-64 (no FMA contraction), so the rows are BITWISE those of the host path
// (tests/test_gpu_pinv.py::test_ls_rows_on_device_equal_host_rows).
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../include/wenoext_b200.h"
#include "plan_internal.h"

namespace wenoext {

// pinv.cu: the batched pseudo-inverse on matrices that already sit on the device
int pinvBatchedOnDevice(int device, int32_t nMat, int32_t n, int maxRows, const int32_t* dRows,
                        const int64_t* dStart, const double* dA, double minCondition, double* dP,
                        int32_t* dZeros, double* dCond, int32_t* dConv, std::string& err);

namespace {

constexpr int LS_MAX_ORDER = 4;                       // cubes of (order+1)^3 <= 125 doubles per thread
constexpr int LS_MAX_CUBE = (LS_MAX_ORDER + 1) * (LS_MAX_ORDER + 1) * (LS_MAX_ORDER + 1);

// geometryWENO.H:198-213: 13-point degree-7 rule on the unit triangle (same table as the host)
__constant__ double c_gauss[13][3] = {
    {0.33333333333333, 0.33333333333333, -0.14957004446768},
    {0.26034596607904, 0.26034596607904, 0.17561525743321},
    {0.26034596607904, 0.47930806784192, 0.17561525743321},
    {0.47930806784192, 0.26034596607904, 0.17561525743321},
    {0.06513010290222, 0.06513010290222, 0.05334723560884},
    {0.06513010290222, 0.86973979419557, 0.05334723560884},
    {0.86973979419557, 0.06513010290222, 0.05334723560884},
    {0.31286549600487, 0.63844418856981, 0.07711376089026},
    {0.63844418856981, 0.04869031542532, 0.07711376089026},
    {0.04869031542532, 0.31286549600487, 0.07711376089026},
    {0.63844418856981, 0.31286549600487, 0.07711376089026},
    {0.31286549600487, 0.04869031542532, 0.07711376089026},
    {0.04869031542532, 0.63844418856981, 0.07711376089026}};

struct V3 {
    double x, y, z;
};
__device__ __forceinline__ V3 sub(const V3& a, const V3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 crossV(const V3& a, const V3& b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ double dotV(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ double magV(const V3& a) { return sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
__device__ __forceinline__ double ipow(double b, unsigned e) {
    double t = 1.0;
    for (unsigned i = 0; i < e; ++i) t *= b;
    return t;
}
__device__ __forceinline__ double fact(int n) {
    double f = 1;
    for (int i = 2; i <= n; ++i) f *= i;
    return f;
}
// geometryWENO::transformPoint (geometryWENO.C:327-343)
__device__ __forceinline__ V3 transformPoint(const double* J, const V3& p, const V3& x0) {
    const double v0 = p.x - x0.x, v1 = p.y - x0.y, v2 = p.z - x0.z;
    return {J[0] * v0 + J[1] * v1 + J[2] * v2, J[3] * v0 + J[4] * v1 + J[5] * v2,
            J[6] * v0 + J[7] * v1 + J[8] * v2};
}
// geometryWENO::gaussQuad, scalar branch (geometryWENO.C:346-460)
__device__ double gaussQuad(int n, int m, int l, const V3& xi0, const V3& v0, const V3& v1, const V3& v2) {
    double sum = 0.0;
    for (int j = 0; j < 13; ++j) {
        const double g0 = c_gauss[j][0], g1 = c_gauss[j][1];
        const double xi = v0.x * (1 - g0 - g1) + v1.x * g0 + v2.x * g1 - xi0.x;
        const double eta = v0.y * (1 - g0 - g1) + v1.y * g0 + v2.y * g1 - xi0.y;
        const double zeta = v0.z * (1 - g0 - g1) + v1.z * g0 + v2.z * g1 - xi0.z;
        sum += c_gauss[j][2] * ipow(xi, n) * ipow(eta, m) * ipow(zeta, l);
    }
    return sum;
}

struct LsArgs {
    int32_t polOrder, nDvt, dim[3];
    const double* points;       // [nPoints][3]
    const int64_t* triStart;    // [nSrcCells+1]
    const int32_t* triPoints;   // [nTri][3]
    const double* cellCentres;  // [nSrcCells][3]
    const double* cellVolumes;  // [nSrcCells]
    const double* JInv;         // [nFrames][9]
    const double* refPoint;     // [nFrames][3]
    const double* refDet;       // [nFrames]
    const int32_t* frameCell;   // [nFrames]
    const double* volI;         // [nFrames][(polOrder+1)^3]
    int64_t nPairs;
    const int32_t* pairFrame;
    const int32_t* pairCell;
    double* rows;               // [nPairs][nDvt]
};

__global__ void ls_rows_kernel(const LsArgs a) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.nPairs) return;
    const int f = a.pairFrame[p], cj = a.pairCell[p], ci = a.frameCell[f];
    const int P = a.polOrder, s1 = P + 1;
    double J[9];
    for (int k = 0; k < 9; ++k) J[k] = a.JInv[(size_t)f * 9 + k];
    const V3 x0{a.refPoint[(size_t)f * 3], a.refPoint[(size_t)f * 3 + 1], a.refPoint[(size_t)f * 3 + 2]};
    auto centreOf = [&](int c) {
        return V3{a.cellCentres[(size_t)c * 3], a.cellCentres[(size_t)c * 3 + 1], a.cellCentres[(size_t)c * 3 + 2]};
    };
    const V3 tI = transformPoint(J, centreOf(ci), x0);
    const V3 tJ = transformPoint(J, centreOf(cj), x0);

    // surfaceMoments of cell J about tJ in the frame of cell I (setup/geometry_weno.cpp: surfaceMoments)
    double mom[LS_MAX_CUBE];
    for (int k = 0; k < s1 * s1 * s1; ++k) mom[k] = 0.0;
    for (int64_t t = a.triStart[cj]; t < a.triStart[cj + 1]; ++t) {
        V3 v[3];
        for (int k = 0; k < 3; ++k) {
            const int32_t pt = a.triPoints[t * 3 + k];
            v[k] = transformPoint(J, V3{a.points[(size_t)pt * 3], a.points[(size_t)pt * 3 + 1], a.points[(size_t)pt * 3 + 2]}, x0);
        }
        V3 vn = crossV(sub(v[1], v[0]), sub(v[2], v[0]));
        const double area = 0.5 * magV(vn);
        if (area < 1.0e-150) continue;  // ROOTVSMALL
        const double sgn = dotV(vn, sub(v[0], tJ)) < 0.0 ? (-1.0 / magV(vn)) : (1.0 / magV(vn));
        vn = V3{sgn * vn.x, sgn * vn.y, sgn * vn.z};
        for (int n = 0; n <= P; ++n)
            for (int m = 0; m <= P; ++m)
                for (int l = 0; l <= P; ++l) {
                    if (n + m + l > P) continue;
                    double& o = mom[(n * s1 + m) * s1 + l];
                    if (n > 0)
                        o += 1.0 / (n + 1) * area * vn.x * gaussQuad(n + 1, m, l, tJ, v[0], v[1], v[2]);
                    else if (m > 0)
                        o += 1.0 / (m + 1) * area * vn.y * gaussQuad(n, m + 1, l, tJ, v[0], v[1], v[2]);
                    else
                        o += 1.0 / (l + 1) * area * vn.z * gaussQuad(n, m, l + 1, tJ, v[0], v[1], v[2]);
                }
    }
    // transformIntegral: scale by 1 / (|det JInv_I| * V_J)
    const double scale = 1.0 / (fabs(a.refDet[f]) * a.cellVolumes[cj]);
    for (int k = 0; k < s1 * s1 * s1; ++k) mom[k] *= scale;

    // the row: calcGeom for every (n, m, l) of the polynomial space, in the enumeration order of the matrix
    // columns (WENOBase.C:646-670), small entries flushed to zero
    const V3 x = sub(tJ, tI);
    const double* vI = a.volI + (size_t)f * s1 * s1 * s1;
    double* row = a.rows + (size_t)p * a.nDvt;
    int col = 0;
    for (int n = 0; n <= a.dim[0]; ++n)
        for (int m = 0; m <= a.dim[1]; ++m)
            for (int l = 0; l <= a.dim[2]; ++l) {
                if (n + m + l > P || n + m + l == 0) continue;
                double g = 0.0;
                for (int k = 0; k <= n; ++k)
                    for (int q = 0; q <= m; ++q)
                        for (int j = 0; j <= l; ++j)
                            g += fact(n) / (fact(k) * fact(n - k)) * fact(m) / (fact(q) * fact(m - q)) * fact(l) /
                                 (fact(j) * fact(l - j)) * ipow(x.x, k) * ipow(x.y, q) * ipow(x.z, j) *
                                 mom[((n - k) * s1 + (m - q)) * s1 + (l - j)];
                const double v = g - vI[(n * s1 + m) * s1 + l];
                if (col < a.nDvt) row[col] = fabs(v) < 1e-9 ? 0.0 : v;
                ++col;
            }
}

// A_s[r][k] = rows[rowPair[r0 + r]][k], r0 = start[s] / nDvt: one thread per matrix entry
__global__ void ls_gather_kernel(int64_t nEntries, int nDvt, const int64_t* __restrict__ rowPair,
                                 const double* __restrict__ rows, double* __restrict__ A) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nEntries) return;
    const int64_t r = i / nDvt;
    A[i] = rows[rowPair[r] * nDvt + (i - r * nDvt)];
}

template <class T>
struct Dev {
    T* p = nullptr;
    ~Dev() {
        if (p) cudaFree(p);
    }
    cudaError_t up(const T* h, size_t n) {
        cudaError_t e = cudaMalloc((void**)&p, std::max<size_t>(n, 1) * sizeof(T));
        if (e == cudaSuccess && h && n) e = cudaMemcpy(p, h, n * sizeof(T), cudaMemcpyHostToDevice);
        return e;
    }
};

}  // namespace
}  // namespace wenoext

extern "C" int wenoext_ls_pinv_batched(const wenoext_ls_batch* b, double minCondition, double* Ainv,
                                       int32_t* nZeros, double* condition, int32_t* converged, double* rowsOut,
                                       double* Aout, int device) {
    using namespace wenoext;
    auto bad = [](const char* what) {
        setError(std::string("ls_pinv_batched: ") + what);
        return WENOEXT_ERR_INVALID;
    };
    if (!b) return bad("batch is NULL");
    if (b->polOrder < 1 || b->polOrder > LS_MAX_ORDER) return bad("polynomial order out of range (1..4)");
    if (b->nDvt <= 0 || b->nPoints <= 0 || b->nSrcCells <= 0 || b->nFrames <= 0 || b->nPairs < 0 || b->nMat < 0)
        return bad("bad sizes");
    if (!b->points || !b->triStart || !b->triPoints || !b->cellCentres || !b->cellVolumes || !b->JInv ||
        !b->refPoint || !b->refDet || !b->frameCell || !b->volI || (b->nPairs && (!b->pairFrame || !b->pairCell)))
        return bad("required array is NULL");
    if (b->nMat > 0 && (!b->rows || !b->start || !b->rowPair || !Ainv || !nZeros || !converged))
        return bad("matrix arrays are NULL");
    // ids the kernels follow unchecked
    const int64_t nTri = b->triStart[b->nSrcCells];
    for (int64_t t = 0; t < nTri * 3; ++t)
        if (b->triPoints[t] < 0 || b->triPoints[t] >= b->nPoints) return bad("triPoints out of range");
    for (int32_t f = 0; f < b->nFrames; ++f)
        if (b->frameCell[f] < 0 || b->frameCell[f] >= b->nSrcCells) return bad("frameCell out of range");
    for (int64_t p = 0; p < b->nPairs; ++p)
        if (b->pairFrame[p] < 0 || b->pairFrame[p] >= b->nFrames || b->pairCell[p] < 0 ||
            b->pairCell[p] >= b->nSrcCells)
            return bad("pair out of range");
    int maxRows = 0;
    int64_t nRowsTotal = 0;
    for (int32_t i = 0; i < b->nMat; ++i) {
        if (b->rows[i] < 1 || b->start[i + 1] - b->start[i] != (int64_t)b->rows[i] * b->nDvt ||
            b->start[i] != nRowsTotal * b->nDvt)
            return bad("start[] does not match rows[] x nDvt");
        maxRows = std::max(maxRows, b->rows[i]);
        nRowsTotal += b->rows[i];
    }
    for (int64_t r = 0; r < nRowsTotal; ++r)
        if (b->rowPair[r] < 0 || b->rowPair[r] >= b->nPairs) return bad("rowPair out of range");

    auto fail = [](const char* what, cudaError_t e) {
        setError(std::string("ls_pinv_batched: ") + what + ": " + cudaGetErrorString(e));
        return WENOEXT_ERR_CUDA;
    };
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail("cudaSetDevice", e);
    const int cube = (b->polOrder + 1) * (b->polOrder + 1) * (b->polOrder + 1);
    Dev<double> dPts, dCen, dVol, dJ, dRef, dDet, dVolI, dRowsV, dA, dP, dCond;
    Dev<int64_t> dTriStart, dStart, dRowPair;
    Dev<int32_t> dTri, dFrameCell, dPF, dPC, dRowsN, dZ, dConv;
    const size_t total = (size_t)nRowsTotal * b->nDvt;
    if ((e = dPts.up(b->points, (size_t)b->nPoints * 3)) || (e = dTriStart.up(b->triStart, (size_t)b->nSrcCells + 1)) ||
        (e = dTri.up(b->triPoints, (size_t)nTri * 3)) || (e = dCen.up(b->cellCentres, (size_t)b->nSrcCells * 3)) ||
        (e = dVol.up(b->cellVolumes, (size_t)b->nSrcCells)) || (e = dJ.up(b->JInv, (size_t)b->nFrames * 9)) ||
        (e = dRef.up(b->refPoint, (size_t)b->nFrames * 3)) || (e = dDet.up(b->refDet, (size_t)b->nFrames)) ||
        (e = dFrameCell.up(b->frameCell, (size_t)b->nFrames)) || (e = dVolI.up(b->volI, (size_t)b->nFrames * cube)) ||
        (e = dPF.up(b->pairFrame, (size_t)b->nPairs)) || (e = dPC.up(b->pairCell, (size_t)b->nPairs)) ||
        (e = dRowsV.up(nullptr, (size_t)b->nPairs * b->nDvt)))
        return fail("upload", e);
    LsArgs a{};
    a.polOrder = b->polOrder;
    a.nDvt = b->nDvt;
    for (int k = 0; k < 3; ++k) a.dim[k] = b->dim[k];
    a.points = dPts.p;
    a.triStart = dTriStart.p;
    a.triPoints = dTri.p;
    a.cellCentres = dCen.p;
    a.cellVolumes = dVol.p;
    a.JInv = dJ.p;
    a.refPoint = dRef.p;
    a.refDet = dDet.p;
    a.frameCell = dFrameCell.p;
    a.volI = dVolI.p;
    a.nPairs = b->nPairs;
    a.pairFrame = dPF.p;
    a.pairCell = dPC.p;
    a.rows = dRowsV.p;
    cudaEvent_t t0, t1;
    cudaEventCreate(&t0);
    cudaEventCreate(&t1);
    cudaEventRecord(t0);
    if (b->nPairs > 0) ls_rows_kernel<<<(unsigned)((b->nPairs + 127) / 128), 128>>>(a);
    cudaEventRecord(t1);
    if ((e = cudaDeviceSynchronize()) != cudaSuccess) return fail("ls_rows_kernel", e);
    if (getenv("WENOEXT_SETUP_TIMING")) {
        float ms = 0;
        cudaEventElapsedTime(&ms, t0, t1);
        fprintf(stderr, "[wenoext setup] ls_rows_kernel: %lld rows of %d in %.3f ms\n", (long long)b->nPairs, b->nDvt, ms);
    }
    cudaEventDestroy(t0);
    cudaEventDestroy(t1);
    if (rowsOut && b->nPairs)
        cudaMemcpy(rowsOut, dRowsV.p, (size_t)b->nPairs * b->nDvt * 8, cudaMemcpyDeviceToHost);
    if (b->nMat == 0) return 0;
    if ((e = dRowPair.up(b->rowPair, (size_t)nRowsTotal)) || (e = dStart.up(b->start, (size_t)b->nMat + 1)) ||
        (e = dRowsN.up(b->rows, (size_t)b->nMat)) || (e = dA.up(nullptr, total)) || (e = dP.up(nullptr, total)) ||
        (e = dCond.up(nullptr, (size_t)b->nMat)) || (e = dZ.up(nullptr, (size_t)b->nMat)) ||
        (e = dConv.up(nullptr, (size_t)b->nMat)))
        return fail("matrix buffers", e);
    ls_gather_kernel<<<(unsigned)((total + 255) / 256), 256>>>((int64_t)total, b->nDvt, dRowPair.p, dRowsV.p, dA.p);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail("ls_gather_kernel", e);
    std::string err;
    if (int rc = pinvBatchedOnDevice(device, b->nMat, b->nDvt, maxRows, dRowsN.p, dStart.p, dA.p, minCondition,
                                     dP.p, dZ.p, dCond.p, dConv.p, err)) {
        setError(err);
        return rc;
    }
    if (Aout) cudaMemcpy(Aout, dA.p, total * 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(Ainv, dP.p, total * 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(nZeros, dZ.p, (size_t)b->nMat * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(converged, dConv.p, (size_t)b->nMat * 4, cudaMemcpyDeviceToHost);
    if (condition) cudaMemcpy(condition, dCond.p, (size_t)b->nMat * 8, cudaMemcpyDeviceToHost);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail("download", e);
    return 0;
}
