// Batched truncated pseudo-inverse on the GPU: the per-(cell, stencil)
// `SVD(A, minCondition)` + `VSinvUt()` of WENOBase::calcMatrix
// (libWENOEXT/WENOBase/WENOBase.C:527, :575; the SVD itself is OpenFOAM's, not in the
// reference tree) for all stencils of a mesh at once. Same algorithm as the host stand-in
// (setup/svd.cpp): one-sided Jacobi on the columns of A, singular values <= minCondition *
// max zeroed, A^+ = V diag(1/S) U^T. One warp per matrix, the matrix in shared memory.
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../include/wenoext_b200.h"
#include "plan_internal.h"

namespace {

struct PinvArgs {
    int32_t nMat, n, maxRows;
    const int32_t* rows;
    const int64_t* start;
    const double* A;
    double minCondition;
    double* Ainv;
    int32_t* nZeros;
    double* condition;
    int32_t* converged;
};

__device__ __forceinline__ double warpSum(double v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
    return v;
}

constexpr int PINV_WARPS = 4;

__global__ void __launch_bounds__(PINV_WARPS * 32) pinv_batched_kernel(const PinvArgs a) {
    extern __shared__ double sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = a.n;
    const size_t perWarp = (size_t)a.maxRows * n + (size_t)n * n + n;
    double* W = sm + perWarp * warp;  // m x n row-major, rotated in place
    double* V = W + (size_t)a.maxRows * n;
    double* S = V + (size_t)n * n;
    for (int mat = blockIdx.x * PINV_WARPS + warp; mat < a.nMat; mat += gridDim.x * PINV_WARPS) {
        const int m = a.rows[mat];
        const double* A = a.A + a.start[mat];
        for (int e = lane; e < m * n; e += 32) W[e] = A[e];
        for (int e = lane; e < n * n; e += 32) V[e] = (e / n == e % n) ? 1.0 : 0.0;
        __syncwarp();
        bool conv = false;
        for (int sweep = 0; sweep < 60 && !conv; ++sweep) {
            conv = true;
            for (int p = 0; p < n - 1; ++p)
                for (int q = p + 1; q < n; ++q) {
                    double alpha = 0, beta = 0, gamma = 0;
                    for (int i = lane; i < m; i += 32) {
                        const double wp = W[i * n + p], wq = W[i * n + q];
                        alpha = fma(wp, wp, alpha);
                        beta = fma(wq, wq, beta);
                        gamma = fma(wp, wq, gamma);
                    }
                    alpha = warpSum(alpha);
                    beta = warpSum(beta);
                    gamma = warpSum(gamma);
                    // (the butterfly gives every lane the same bits: the branch is uniform)
                    if (gamma == 0.0 || fabs(gamma) <= 1e-15 * sqrt(alpha * beta)) continue;
                    conv = false;
                    const double zeta = (beta - alpha) / (2.0 * gamma);
                    const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                    const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                    for (int i = lane; i < m; i += 32) {
                        const double wp = W[i * n + p], wq = W[i * n + q];
                        W[i * n + p] = c * wp - s * wq;
                        W[i * n + q] = s * wp + c * wq;
                    }
                    for (int i = lane; i < n; i += 32) {
                        const double vp = V[i * n + p], vq = V[i * n + q];
                        V[i * n + p] = c * vp - s * vq;
                        V[i * n + q] = s * vp + c * vq;
                    }
                    __syncwarp();
                }
        }
        // singular values = column norms; U = W / S
        double maxS = 0;
        for (int j = 0; j < n; ++j) {
            double nrm = 0;
            for (int i = lane; i < m; i += 32) nrm = fma(W[i * n + j], W[i * n + j], nrm);
            nrm = sqrt(warpSum(nrm));
            if (lane == 0) S[j] = nrm;
            maxS = fmax(maxS, nrm);
        }
        __syncwarp();
        const double minS = a.minCondition * maxS;
        int nZeros = 0;
        double mn = 1e15, mx = -1e15;
        for (int j = 0; j < n; ++j) {
            const double sj = S[j];
            if (sj <= minS) {
                ++nZeros;
            } else {
                mn = fmin(mn, sj);
                mx = fmax(mx, sj);
            }
        }
        // A^+[i][j] = sum_k V[i][k] * (1/S[k]) * U[j][k], zeroed singular values skipped
        double* P = a.Ainv + a.start[mat];
        for (int e = lane; e < n * m; e += 32) {
            const int i = e / m, j = e - i * m;
            double acc = 0;
            for (int k = 0; k < n; ++k) {
                const double sk = S[k];
                if (sk <= minS || sk == 0.0) continue;
                acc += V[i * n + k] * (1.0 / sk) * (W[j * n + k] / sk);
            }
            P[e] = acc;
        }
        if (lane == 0) {
            a.nZeros[mat] = nZeros;
            a.converged[mat] = conv ? 1 : 0;
            if (a.condition) a.condition[mat] = mx / mn;
        }
        __syncwarp();
    }
}

template <class T>
struct DevBuf {
    T* p = nullptr;
    ~DevBuf() {
        if (p) cudaFree(p);
    }
    cudaError_t alloc(size_t n) { return cudaMalloc((void**)&p, std::max<size_t>(n, 1) * sizeof(T)); }
};

}  // namespace

namespace wenoext {

// The batch on matrices that already sit on the device (ls_rows.cu assembles them there): launch geometry,
// shared-memory opt-in, one launch, synchronised on return.
int pinvBatchedOnDevice(int device, int32_t nMat, int32_t n, int maxRows, const int32_t* dRows,
                        const int64_t* dStart, const double* dA, double minCondition, double* dP,
                        int32_t* dZeros, double* dCond, int32_t* dConv, std::string& err) {
    if (nMat <= 0) return 0;
    auto fail = [&](const char* what, cudaError_t e) {
        err = std::string(what) + ": " + cudaGetErrorString(e);
        return WENOEXT_ERR_CUDA;
    };
    const size_t smem = PINV_WARPS * ((size_t)maxRows * n + (size_t)n * n + n) * sizeof(double);
    int smemMax = 0;
    cudaDeviceGetAttribute(&smemMax, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    if (smem > (size_t)smemMax) {
        err = "pinv: matrices do not fit the shared memory";
        return WENOEXT_ERR_INVALID;
    }
    PinvArgs a{nMat, n, maxRows, dRows, dStart, dA, minCondition, dP, dZeros, dCond, dConv};
    cudaError_t e = cudaFuncSetAttribute(pinv_batched_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail("pinv: shared memory", e);
    int nSM = 148;
    cudaDeviceGetAttribute(&nSM, cudaDevAttrMultiProcessorCount, device);
    const int blocksPerSM = std::max<int>(1, std::min<size_t>(8, (size_t)smemMax / std::max<size_t>(smem, 1)));
    const int grid = std::min((nMat + PINV_WARPS - 1) / PINV_WARPS, nSM * blocksPerSM);
    cudaEvent_t t0, t1;
    cudaEventCreate(&t0);
    cudaEventCreate(&t1);
    cudaEventRecord(t0);
    pinv_batched_kernel<<<grid, PINV_WARPS * 32, smem>>>(a);
    cudaEventRecord(t1);
    e = cudaDeviceSynchronize();
    if (e == cudaSuccess && getenv("WENOEXT_SETUP_TIMING")) {
        float ms = 0;
        cudaEventElapsedTime(&ms, t0, t1);
        fprintf(stderr, "[wenoext setup] pinv_batched_kernel: %d matrices (up to %d x %d) in %.3f ms\n", nMat,
                maxRows, n, ms);
    }
    cudaEventDestroy(t0);
    cudaEventDestroy(t1);
    if (e != cudaSuccess) return fail("pinv_batched_kernel", e);
    return 0;
}

int pinvBatchedImpl(int device, int32_t nMat, int32_t n, const int32_t* rows, const int64_t* start,
                      const double* A, double minCondition, double* Ainv, int32_t* nZeros, double* condition,
                      int32_t* converged, std::string& err) {
    if (nMat <= 0) return 0;
    auto fail = [&](const char* what, cudaError_t e) {
        err = std::string(what) + ": " + cudaGetErrorString(e);
        return WENOEXT_ERR_CUDA;
    };
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail("pinv: cudaSetDevice", e);
    int maxRows = 0;
    for (int i = 0; i < nMat; ++i) maxRows = std::max(maxRows, rows[i]);
    const size_t total = (size_t)start[nMat];
    DevBuf<int32_t> dRows, dZ, dConv;
    DevBuf<int64_t> dStart;
    DevBuf<double> dA, dP, dCond;
    if ((e = dRows.alloc(nMat)) != cudaSuccess || (e = dZ.alloc(nMat)) != cudaSuccess ||
        (e = dConv.alloc(nMat)) != cudaSuccess || (e = dStart.alloc((size_t)nMat + 1)) != cudaSuccess ||
        (e = dA.alloc(total)) != cudaSuccess || (e = dP.alloc(total)) != cudaSuccess ||
        (e = dCond.alloc(nMat)) != cudaSuccess)
        return fail("pinv: cudaMalloc", e);
    cudaMemcpy(dRows.p, rows, (size_t)nMat * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dStart.p, start, ((size_t)nMat + 1) * 8, cudaMemcpyHostToDevice);
    e = cudaMemcpy(dA.p, A, total * 8, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return fail("pinv: upload", e);
    if (int rc = pinvBatchedOnDevice(device, nMat, n, maxRows, dRows.p, dStart.p, dA.p, minCondition, dP.p, dZ.p,
                                     dCond.p, dConv.p, err))
        return rc;
    cudaMemcpy(Ainv, dP.p, total * 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(nZeros, dZ.p, (size_t)nMat * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(converged, dConv.p, (size_t)nMat * 4, cudaMemcpyDeviceToHost);
    if (condition) cudaMemcpy(condition, dCond.p, (size_t)nMat * 8, cudaMemcpyDeviceToHost);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail("pinv: download", e);
    return 0;
}

}  // namespace wenoext

extern "C" int wenoext_pinv_batched(int32_t nMat, int32_t n, const int32_t* rows, const int64_t* start,
                                    const double* A, double minCondition, double* Ainv, int32_t* nZeros,
                                    double* condition, int32_t* converged, int device) {
    if (nMat < 0 || n <= 0 || (nMat > 0 && (!rows || !start || !A || !Ainv || !nZeros || !converged))) {
        wenoext::setError("pinv_batched: bad argument");
        return WENOEXT_ERR_INVALID;
    }
    for (int i = 0; i < nMat; ++i)
        if (rows[i] < 1 || start[i + 1] - start[i] != (int64_t)rows[i] * n) {
            wenoext::setError("pinv_batched: start[] does not match rows[] x n");
            return WENOEXT_ERR_INVALID;
        }
    std::string err;
    const int rc = wenoext::pinvBatchedImpl(device, nMat, n, rows, start, A, minCondition, Ainv, nZeros,
                                              condition, converged, err);
    if (rc) wenoext::setError(err);
    return rc;
}
