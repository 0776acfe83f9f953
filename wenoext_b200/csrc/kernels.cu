// Hand-written sm_100a kernels of the WENOEXT reconstruction hot path.
//
// weno_cell_kernel fuses, per cell, what the reference does in three separate
// passes with a [nCells][K][C] intermediate in between:
//   calcCoeff   (WENOCoeff.C:104-165)  coeffs_s = A_s * (psi_j - psi_i)
//   calcWeight  (WENOCoeff.C:168-221)  IS = c^T B c, gamma, weighted sum
//   face loop + sumFlux (WENOUpwindFit.C:256-323) and the limiters
//   (WENOUpwindFit.C:81-108,172-179; WENOUpwindFit01.C:66-181)
// Every internal face is written by exactly one cell (its upwind cell), so the
// kernel is cell-centric and nothing but the per-face result goes back to HBM.
//
// Structure: persistent CTAs (one per SM). One producer warp streams whole cell
// records (device_plan.h) into a shared-memory ring with 1-D TMA bulk copies
// (cp.async.bulk ... mbarrier::complete_tx, SASS UBLKCP); consumer warps take
// one cell each. The kernel is HBM-bound: tensor cores are deliberately unused
// (independent 9x21 GEMVs, no operand reuse).
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_plan.h"

namespace wenoext {

namespace {

constexpr int TILE = 32;  // consecutive cells handed to one CTA at a time

__device__ __forceinline__ uint32_t smemAddr(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbarInit(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count));
}
__device__ __forceinline__ void mbarArriveExpectTx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbarArrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smemAddr(bar)) : "memory");
}
__device__ __forceinline__ void mbarWait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smemAddr(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ uint64_t evictFirstPolicy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// 1-D TMA bulk copy global -> shared, completion on an mbarrier
__device__ __forceinline__ void tmaLoad1D(void* dst, const void* src, uint32_t bytes,
                                          uint64_t* bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
        "[%0], [%1], %2, [%3], %4;" ::"r"(smemAddr(dst)),
        "l"(src), "r"(bytes), "r"(smemAddr(bar)), "l"(policy)
        : "memory");
}

__device__ __forceinline__ double intPowD(double base, uint32_t e) {
    double t = 1.0;
    for (uint32_t i = 0; i < e; ++i) t *= base;  // WENOCoeff.H:149-157
    return t;
}

struct WarpScratch {
    double* bsm;    // [max(maxTotN, maxSt*K)][CB]  b values, later IS products
    double* coef;   // [maxSt*K][CB]
    double* gam;    // [maxSt][CB]
    double* W;      // [K][CB] weighted coefficients
    double* poly;   // [maxFc][CB]
    double* theta;  // [4]
    int32_t* stN;   // [maxSt]
    int32_t* stA;   // [maxSt] offset (doubles) of A_s in the record
    int32_t* stB;   // [maxSt] offset (entries) of b_s
};

__host__ __device__ inline size_t alignUp(size_t x, size_t a) { return (x + a - 1) / a * a; }

__host__ __device__ inline size_t warpScratchBytes(int K, int CB, int maxTotN, int maxSt,
                                                   int maxFc) {
    size_t nb = (size_t)(maxTotN > maxSt * K ? maxTotN : maxSt * K);
    size_t d = nb * CB + (size_t)maxSt * K * CB + (size_t)maxSt * CB + (size_t)K * CB +
               (size_t)maxFc * CB + 4;
    size_t i = 3 * (size_t)maxSt;
    return alignUp(d * 8 + i * 4, 16);
}

template <int CB>
__device__ __forceinline__ void loadVals(const double* __restrict__ p, double (&v)[CB]) {
    if constexpr (CB == 1) {
        v[0] = __ldg(p);
    } else if constexpr (CB == 2) {
        double2 t = __ldg(reinterpret_cast<const double2*>(p));
        v[0] = t.x;
        v[1] = t.y;
    } else {
        double2 t = __ldg(reinterpret_cast<const double2*>(p));
        v[0] = t.x;
        v[1] = t.y;
        v[2] = __ldg(p + 2);
    }
}

template <int CB>
__global__ void __launch_bounds__(1024, 1) weno_cell_kernel(const __grid_constant__ CellKernelArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int K = a.K;
    const int NW = a.nConsumerWarps;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    unsigned char* ring = smem;
    uint64_t* fullBar = reinterpret_cast<uint64_t*>(smem + (size_t)a.nSlots * a.slotBytes);
    uint64_t* emptyBar = fullBar + a.nSlots;
    unsigned char* scratchBase = reinterpret_cast<unsigned char*>(emptyBar + a.nSlots);
    scratchBase = reinterpret_cast<unsigned char*>(alignUp((size_t)scratchBase, 16));

    if (threadIdx.x == 0) {
        for (int s = 0; s < a.nSlots; ++s) {
            mbarInit(&fullBar[s], 1);
            mbarInit(&emptyBar[s], 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // static work split: tiles of TILE consecutive cells, round-robin over CTAs
    const int nTiles = (a.nWork + TILE - 1) / TILE;
    int nItems = 0;
    if (nTiles > (int)blockIdx.x) {
        const int nMy = (nTiles - (int)blockIdx.x - 1) / (int)gridDim.x + 1;
        const int lastTile = blockIdx.x + (nMy - 1) * gridDim.x;
        nItems = nMy * TILE - (lastTile == nTiles - 1 ? nTiles * TILE - a.nWork : 0);
    }
    auto itemToWork = [&](int n) {
        return ((int)blockIdx.x + (n / TILE) * (int)gridDim.x) * TILE + (n % TILE);
    };

    if (warp == NW) {
        // ===== producer: one lane issues the bulk copies =====
        if (lane == 0) {
            const uint64_t pol = evictFirstPolicy();
            for (int n = 0; n < nItems; ++n) {
                const int slot = n % a.nSlots;
                const int round = n / a.nSlots;
                if (round > 0) mbarWait(&emptyBar[slot], (round - 1) & 1);
                const int w = itemToWork(n);
                const int cell = a.cellList ? a.cellList[w] : w;
                const int64_t off = a.recOff[cell];
                const uint32_t bytes = (uint32_t)(a.recOff[cell + 1] - off);
                mbarArriveExpectTx(&fullBar[slot], bytes);
                tmaLoad1D(ring + (size_t)slot * a.slotBytes, a.records + off, bytes,
                          &fullBar[slot], pol);
            }
        }
        return;
    }
    if (warp > NW) return;

    // ===== consumers: one warp per cell =====
    WarpScratch ws;
    {
        const size_t per = warpScratchBytes(K, CB, a.maxTotN, a.maxSt, a.maxFc);
        unsigned char* p = scratchBase + (size_t)warp * per;
        const size_t nb = (size_t)(a.maxTotN > a.maxSt * K ? a.maxTotN : a.maxSt * K);
        ws.bsm = reinterpret_cast<double*>(p);
        ws.coef = ws.bsm + nb * CB;
        ws.gam = ws.coef + (size_t)a.maxSt * K * CB;
        ws.W = ws.gam + (size_t)a.maxSt * CB;
        ws.poly = ws.W + (size_t)K * CB;
        ws.theta = ws.poly + (size_t)a.maxFc * CB;
        ws.stN = reinterpret_cast<int32_t*>(ws.theta + 4);
        ws.stA = ws.stN + a.maxSt;
        ws.stB = ws.stA + a.maxSt;
    }
    const int nBndBase = a.nInternalFaces;

    for (int n = warp; n < nItems; n += NW) {
        const int slot = n % a.nSlots;
        const int round = n / a.nSlots;
        const int w = itemToWork(n);
        const int cell = a.cellList ? a.cellList[w] : w;
        // centre value (broadcast load) while the record is in flight
        double cen[CB];
        loadVals<CB>(a.vals + (size_t)cell * a.VS, cen);

        mbarWait(&fullBar[slot], round & 1);
        const unsigned char* rec = ring + (size_t)slot * a.slotBytes;
        const RecHeader hd = *reinterpret_cast<const RecHeader*>(rec);
        const int nSt = hd.nSt, nFc = hd.nFc, totN = hd.totN;
        const double* Amat = reinterpret_cast<const double*>(rec + sizeof(RecHeader));
        // A size is only known after the prefix sum below; the int block is found via totN
        const double* Bm = Amat + (size_t)K * totN;
        const double* faceD = Bm + K * K;
        const double* quirk = faceD + (size_t)nFc * (K + 1);
        const int32_t* Ns = reinterpret_cast<const int32_t*>(quirk + K);
        const int32_t* idx = Ns + nSt;
        const int32_t* faceI = idx + totN;

        // per-stencil offsets (nSt <= 32 guaranteed by the plan)
        {
            int N = lane < nSt ? Ns[lane] : 0;
            int incl = N;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            if (lane < nSt) {
                ws.stN[lane] = N;
                ws.stB[lane] = incl - N;
                ws.stA[lane] = (incl - N) * K;
            }
        }
        // gather: b[e][c] = psi(id_e)[c] - psi(cell)[c]      (WENOCoeff.C:138-156)
        for (int e = lane; e < totN; e += 32) {
            double v[CB];
            loadVals<CB>(a.vals + (size_t)idx[e] * a.VS, v);
#pragma unroll
            for (int c = 0; c < CB; ++c) ws.bsm[e * CB + c] = v[c] - cen[c];
        }
        __syncwarp();
        // phase 1: coeffs[s][k][c] = sum_j A_s[k][j] * b_s[j][c]   (WENOCoeff.C:159)
        const int nPairs = nSt * K;
        for (int p = lane; p < nPairs; p += 32) {
            const int s = p / K, k = p - s * K;
            const int N = ws.stN[s];
            const double* Ar = Amat + ws.stA[s] + (size_t)k * N;
            const double* bp = ws.bsm + (size_t)ws.stB[s] * CB;
            double acc[CB];
#pragma unroll
            for (int c = 0; c < CB; ++c) acc[c] = 0.0;
            for (int j = 0; j < N; ++j) {
                const double av = Ar[j];
#pragma unroll
                for (int c = 0; c < CB; ++c) acc[c] = fma(av, bp[j * CB + c], acc[c]);
            }
#pragma unroll
            for (int c = 0; c < CB; ++c) ws.coef[p * CB + c] = acc[c];
        }
        __syncwarp();
        // phase 2a: t[s][k][c] = c_k * (B c)_k                  (WENOCoeff.C:190-196)
        for (int p = lane; p < nPairs; p += 32) {
            const int s = p / K, k = p - s * K;
            const double* cs = ws.coef + (size_t)s * K * CB;
            double u[CB];
#pragma unroll
            for (int c = 0; c < CB; ++c) u[c] = 0.0;
            for (int q = 0; q < K; ++q) {
                const double bv = Bm[k * K + q];
#pragma unroll
                for (int c = 0; c < CB; ++c) u[c] = fma(bv, cs[q * CB + c], u[c]);
            }
#pragma unroll
            for (int c = 0; c < CB; ++c) ws.bsm[p * CB + c] = ws.coef[p * CB + c] * u[c];
        }
        __syncwarp();
        // phase 2b: IS and gamma per (stencil, component)      (WENOCoeff.C:196-205)
        for (int i = lane; i < nSt * CB; i += 32) {
            const int s = i / CB, c = i - s * CB;
            double IS = 0.0;
            for (int k = 0; k < K; ++k) IS += ws.bsm[(s * K + k) * CB + c];
            ws.gam[i] = (s == 0 ? a.dm : 1.0) / intPowD(a.eps + IS, a.pExp);
        }
        __syncwarp();
        // phase 2c: weighted combination                        (WENOCoeff.C:207-220)
        for (int i = lane; i < K * CB; i += 32) {
            const int k = i / CB, c = i - k * CB;
            double acc = 0.0, gs = 0.0;
            for (int s = 0; s < nSt; ++s) {
                const double g = ws.gam[s * CB + c];
                gs += g;
                acc = fma(g, ws.coef[(s * K + k) * CB + c], acc);
            }
            ws.W[i] = nSt > 0 ? acc / gs : 0.0;  // Cell::empty -> zero (WENOCoeff.C:333)
        }
        __syncwarp();
        // phase 3: faces
        const bool fit01 = a.scheme == 1;
        if (fit01) {
            // own polynomial at every internal face (WENOUpwindFit01.C:80-136)
            for (int i = lane; i < nFc * CB; i += 32) {
                const int f = i / CB, c = i - f * CB;
                const int kind = faceI[f * 3 + 2];
                double val = 0.0;
                if ((kind >> FACE_TYPE_SHIFT) == FACE_INTERNAL) {
                    const double* fd = faceD + (size_t)f * (K + 1);
                    const double fl = __ldg(a.flux + faceI[f * 3]);
                    const bool nei = kind & FACE_SIDE_NEI;
                    const double* basis = (nei && !(fl < 0) && a.bugCompat) ? quirk : fd;
                    double s = 0.0;
                    for (int k = 0; k < K; ++k) s = fma(ws.W[k * CB + c], basis[k], s);
                    val = cen[c] + s / fd[K];
                } else {
                    val = cen[c];
                }
                ws.poly[i] = val;
            }
            __syncwarp();
            if (lane < CB) {
                const int c = lane;
                double psi = 0.0;
#pragma unroll
                for (int cc = 0; cc < CB; ++cc)
                    if (cc == c) psi = cen[cc];
                double mx = psi, mn = psi;
                for (int f = 0; f < nFc; ++f) {
                    if ((faceI[f * 3 + 2] >> FACE_TYPE_SHIFT) != FACE_INTERNAL) continue;
                    const double t = ws.poly[f * CB + c];
                    if (t > mx)
                        mx = t;
                    else if (t < mn)
                        mn = t;
                }
                const double argMax = fabs(mx - psi) < 1e-9 ? 1.0 : fabs((1.0 - psi) / (mx - psi));
                const double argMin = fabs(mn - psi) < 1e-9 ? 1.0 : fabs((0.0 - psi) / (mn - psi));
                ws.theta[c] = fmin(fmin(argMax, argMin), 1.0);
            }
            __syncwarp();
        }
        for (int i = lane; i < nFc * CB; i += 32) {
            const int f = i / CB, c = i - f * CB;
            const int faceId = faceI[f * 3];
            const int other = faceI[f * 3 + 1];
            const int kind = faceI[f * 3 + 2];
            const int type = kind >> FACE_TYPE_SHIFT;
            const bool nei = kind & FACE_SIDE_NEI;
            const double fl = __ldg(a.flux + faceId);
            double psi = 0.0;
#pragma unroll
            for (int cc = 0; cc < CB; ++cc)
                if (cc == c) psi = cen[cc];
            // who writes: owner unless flux<0 (flux==0 -> owner writes zero), WENOUpwindFit.C:258-281
            const bool up = nei ? (fl < 0) : (fl > 0);
            const bool writer = type == FACE_INTERNAL ? (nei ? (fl < 0) : !(fl < 0)) : (fl > 0);
            if (fit01 && type == FACE_PROCESSOR && a.thetaPatch)
                a.thetaPatch[(size_t)(faceId - nBndBase) * a.nComp + a.c0 + c] = ws.theta[c];
            if (!writer) continue;
            double out = 0.0;
            if (up) {
                const double* fd = faceD + (size_t)f * (K + 1);
                double s = 0.0;
                for (int k = 0; k < K; ++k) s = fma(ws.W[k * CB + c], fd[k], s);
                out = s / fd[K];
                if (type == FACE_INTERNAL) {
                    if (fit01) {
                        out = ws.theta[c] * out;  // WENOUpwindFit01.C:167-181
                    } else if (a.bounded) {
                        // WENOUpwindFit.C:81-108, :172-179
                        const double psiN = __ldg(a.vals + (size_t)other * a.VS + c);
                        const double dpsi = fabs(psi - psiN);
                        if (fabs(out) > dpsi) out = fmin(dpsi / fabs(out), 1.0) * out;
                    }
                }
            }
            a.corr[(size_t)faceId * a.nComp + a.c0 + c] = out;
            if (a.faceValue && type == FACE_INTERNAL)
                a.faceValue[(size_t)faceId * a.nComp + a.c0 + c] = psi + out;
        }
        __syncwarp();
        if (lane == 0) mbarArrive(&emptyBar[slot]);
    }
}

}  // namespace

namespace {

// vf AoS [nCells][nComp] -> packed value buffer [nCells][VS] for one component
// batch (VS = 1, 2 or 4 so that a 3-vector gather is one 32-byte sector).
__global__ void pack_values_kernel(const double* __restrict__ vf, int nCells, int nComp, int c0,
                                   int CB, int VS, double* __restrict__ vals) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)nCells * VS) return;
    const int cell = (int)(i / VS), c = (int)(i - (int64_t)cell * VS);
    vals[i] = c < CB ? vf[(size_t)cell * nComp + c0 + c] : 0.0;
}

// collectData pack step (WENOCoeff.C:254-258): sendHaloData_[proc][k] = vf[sendHaloCellIDs[k]]
__global__ void pack_halo_kernel(const double* __restrict__ vals, int VS,
                                 const int32_t* __restrict__ sendIds, int64_t nSend,
                                 double* __restrict__ sendBuf) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nSend * VS) return;
    const int64_t e = i / VS;
    const int c = (int)(i - e * VS);
    sendBuf[i] = vals[(size_t)sendIds[e] * VS + c];
}

// Second half of coupledRiemannSolver (WENOUpwindFit.C:474-518) and the
// processor-patch part of the limiters (WENOUpwindFit.C:118-165,:195-207;
// WENOUpwindFit01.C:194-210), one thread per (coupled boundary face, component).
__global__ void patch_kernel(const PatchKernelArgs a) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)a.nBnd * a.CB) return;
    const int bf = (int)(i / a.CB), c = (int)(i - (int64_t)bf * a.CB);
    const int type = a.bfPatchType[bf];
    if (type < 0) return;
    const int cc = a.c0 + c;
    const int f = a.nInternalFaces + bf;
    const double fl = a.flux[f];
    double v = a.corr[(size_t)f * a.nComp + cc];
    bool changed = false;
    if (type == FACE_PROCESSOR) {
        if (fl < 0) {
            v = a.recv[(size_t)bf * a.nComp + cc];
            changed = true;
        }
        if (a.scheme == 1) {
            v = a.thetaPatch[(size_t)bf * a.nComp + cc] * v;
            changed = true;
        } else if (a.bounded) {
            const double own = a.vals[(size_t)a.owner[f] * a.VS + c];
            const double nb = a.patchNbrField ? a.patchNbrField[(size_t)bf * a.nComp + cc]
                                              : a.vals[(size_t)a.bfNbrCell[bf] * a.VS + c];
            const double psi = fl > 0 ? own : nb, psiN = fl > 0 ? nb : own;
            const double dpsi = fabs(psi - psiN);
            if (fabs(v) > dpsi) v = fmin(dpsi / fabs(v), 1.0) * v;
            changed = true;
        }
    } else if (type == FACE_CYCLIC) {
        if (fl < 0) {
            v = a.corr[(size_t)a.bfPartner[bf] * a.nComp + cc];
            changed = true;
        }
    }
    if (changed) a.corr[(size_t)f * a.nComp + cc] = v;
}

}  // namespace

size_t cellKernelSmemBytes(int K, int CB, int nSlots, int slotBytes, int maxTotN, int maxSt,
                           int maxFc, int nConsumerWarps) {
    size_t s = (size_t)nSlots * slotBytes + (size_t)nSlots * 16 + 16;
    s += (size_t)nConsumerWarps * warpScratchBytes(K, CB, maxTotN, maxSt, maxFc);
    return s;
}

void launchCellKernel(const CellKernelArgs& a, int CB, int gridBlocks, size_t smemBytes,
                      cudaStream_t st) {
    const int threads = (a.nConsumerWarps + 1) * 32;
    switch (CB) {
        case 1:
            cudaFuncSetAttribute(weno_cell_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smemBytes);
            weno_cell_kernel<1><<<gridBlocks, threads, smemBytes, st>>>(a);
            break;
        case 2:
            cudaFuncSetAttribute(weno_cell_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smemBytes);
            weno_cell_kernel<2><<<gridBlocks, threads, smemBytes, st>>>(a);
            break;
        default:
            cudaFuncSetAttribute(weno_cell_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smemBytes);
            weno_cell_kernel<3><<<gridBlocks, threads, smemBytes, st>>>(a);
            break;
    }
}

void launchPatchKernel(const PatchKernelArgs& a, cudaStream_t st) {
    const int64_t n = (int64_t)a.nBnd * a.CB;
    if (n <= 0) return;
    patch_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a);
}

void launchPackValues(const double* vf, int nCells, int nComp, int c0, int CB, int VS,
                      double* vals, cudaStream_t st) {
    const int64_t n = (int64_t)nCells * VS;
    if (n <= 0) return;
    pack_values_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(vf, nCells, nComp, c0, CB, VS,
                                                                    vals);
}

void launchPackHalo(const double* vals, int VS, const int32_t* sendIds, int64_t nSend,
                    double* sendBuf, cudaStream_t st) {
    const int64_t n = nSend * VS;
    if (n <= 0) return;
    pack_halo_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(vals, VS, sendIds, nSend,
                                                                  sendBuf);
}

}  // namespace wenoext
