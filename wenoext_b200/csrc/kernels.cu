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
// (cp.async.bulk ... mbarrier::complete_tx, SASS UBLKCP): the small meta part
// and the matrix part complete on separate mbarriers, so the gather of stencil
// values overlaps the arrival of the matrices. Consumer warps take CELLS cells per
// pass and give one lane to every (cell, stencil): the lane gathers its own
// stencil's value differences straight into registers (software-pipelined global
// loads), streams its matrix column by column from shared memory (128-bit loads,
// stencil index fastest, so the lanes of a cell read consecutive words) and keeps
// all K x CB coefficients in registers. Smoothness indicators and weights stay in
// registers; the weighted sum over the stencils of a cell is a warp-shuffle
// butterfly over the cell's lanes; afterwards every lane evaluates one face.
// No scratch shared memory and no block-level barriers. The kernel is HBM-bound:
// tensor cores are deliberately unused (independent 9x21 GEMVs, no operand reuse).
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_plan.h"

namespace wenoext {

namespace {

constexpr int TILE = 32;          // consecutive cells handed to one CTA at a time
constexpr int MAX_THREADS = 256;  // 7 consumer warps + producer (up to 255 registers)

__device__ __forceinline__ uint32_t smemAddr(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbarArriveExpectTx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbarArrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
#ifdef WENOEXT_KDEBUG
__device__ int32_t* g_dbg = nullptr;
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity) {
    for (long long it = 0; it < 4000000; ++it) {
        uint32_t ok;
        asm volatile(
            "{\n.reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
        if (ok) return;
    }
    if (g_dbg && atomicCAS(g_dbg, 0, 4) == 0) {
        g_dbg[1] = (int)bar; g_dbg[2] = (int)parity; g_dbg[10] = blockIdx.x; g_dbg[11] = threadIdx.x;
    }
}
#else
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
#endif
__device__ __forceinline__ uint64_t evictFirstPolicy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// 1-D TMA bulk copy global -> shared, completion on an mbarrier
__device__ __forceinline__ void tmaLoad1D(uint32_t dst, const void* src, uint32_t bytes,
                                          uint32_t bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
        "[%0], [%1], %2, [%3], %4;" ::"r"(dst),
        "l"(src), "r"(bytes), "r"(bar), "l"(policy)
        : "memory");
}
// explicit shared-space accesses (32-bit addresses; guarantees LDS/STS)
__device__ __forceinline__ double lds64(uint32_t a) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ double2 lds128(uint32_t a) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ int32_t lds32(uint32_t a) {
    int32_t v;
    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ int4 lds128i(uint32_t a) {
    int4 v;
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"(a));
    return v;
}

// WENOCoeff.H:149-157: repeated multiplication, exponent truncated to unsigned.
// The first 8 factors are predicated (no loop) - p = 4 is the default.
__device__ __forceinline__ double intPowD(double base, uint32_t e) {
    double t = 1.0;
#pragma unroll
    for (uint32_t i = 0; i < 8; ++i) t = i < e ? t * base : t;
    for (uint32_t i = 8; i < e; ++i) t *= base;
    return t;
}

template <int CB>
__device__ __forceinline__ void loadVals(const double* __restrict__ p, double (&v)[CB]) {
    if constexpr (CB == 1) {
        v[0] = __ldg(p);
    } else {
        double2 t = __ldg(reinterpret_cast<const double2*>(p));
        v[0] = t.x;
        v[1] = t.y;
        if constexpr (CB == 3) v[2] = __ldg(p + 2);
    }
}

__device__ __forceinline__ uint32_t ldsVolatileU32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void stsVolatileU32(uint32_t a, uint32_t v) {
    asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
// Ring slot protocol. Slots are not bound to a consumer, so before waiting for
// round r of a slot a consumer first waits until round r-1 has been RELEASED
// (release counter): an mbarrier parity wait alone cannot tell "round r done"
// from "round r-1 not done yet".
__device__ __forceinline__ void slotWait(uint32_t fullBar, uint32_t relAddr, uint32_t round) {
    while (ldsVolatileU32(relAddr) < round) {
    }
    mbarWait(fullBar, round & 1u);
}
__device__ __forceinline__ void slotRelease(uint32_t emptyBar, uint32_t relAddr, uint32_t round) {
    stsVolatileU32(relAddr, round + 1u);
    mbarArrive(emptyBar);
}

template <int K, int CB, int CELLS, int D>
__global__ void __launch_bounds__(MAX_THREADS, 1)
    weno_cell_kernel(const __grid_constant__ CellKernelArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int SS = 32 / CELLS;  // stencil lanes per cell
    constexpr int KP = K / 2;       // coefficient pairs per column
    constexpr bool ODD = (K & 1) != 0;
    const int NW = a.nWarps;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nMeta = a.nMetaSlots, nChunk = a.nChunkSlots, CH = a.nChunks;

    // shared memory: meta ring | chunk ring | mbarriers | release counters
    const uint32_t metaRing = smemAddr(smem);
    const uint32_t chunkRing = metaRing + (uint32_t)nMeta * a.metaSlotBytes;
    const uint32_t barMetaFull = chunkRing + (uint32_t)nChunk * a.chunkSlotBytes;
    const uint32_t barMetaEmpty = barMetaFull + (uint32_t)nMeta * 8;
    const uint32_t barChunkFull = barMetaEmpty + (uint32_t)nMeta * 8;
    const uint32_t barChunkEmpty = barChunkFull + (uint32_t)nChunk * 8;
    const uint32_t relMeta = barChunkEmpty + (uint32_t)nChunk * 8;
    const uint32_t relChunk = relMeta + (uint32_t)nMeta * 4;

    for (int i = threadIdx.x; i < 2 * (nMeta + nChunk); i += blockDim.x) mbarInit(barMetaFull + i * 8, 1);
    for (int i = threadIdx.x; i < nMeta + nChunk; i += blockDim.x) stsVolatileU32(relMeta + i * 4, 0u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    // static work split: tiles of TILE consecutive cells, round-robin over CTAs
    const int nTiles = (a.nWork + TILE - 1) / TILE;
    int nItems = 0;
    if (nTiles > (int)blockIdx.x) {
        const int nMy = (nTiles - (int)blockIdx.x - 1) / (int)gridDim.x + 1;
        const int lastTile = blockIdx.x + (nMy - 1) * gridDim.x;
        nItems = nMy * TILE - (lastTile == nTiles - 1 ? nTiles * TILE - a.nWork : 0);
    }
    auto itemToCell = [&](int n) {
        const int w = ((int)blockIdx.x + (n / TILE) * (int)gridDim.x) * TILE + (n % TILE);
        return a.cellList ? __ldg(a.cellList + w) : w;
    };

    if (warp == NW) {
        // ===== producer warp. All lanes prefetch the record descriptors of the next
        // 32 items. Per pass (CELLS consecutive items) lane l issues one bulk copy:
        // l < CELLS the meta part of cell l, then chunk ch of cell c in the order the
        // consumers need them (chunk-major). =====
        const uint64_t pol = evictFirstPolicy();
        auto fetch = [&](int n) {
            int4 r = make_int4(0, 0, 0, 0);
            if (n < nItems) r = __ldg(reinterpret_cast<const int4*>(a.recInfo + itemToCell(n)));
            return r;
        };
        const int perPass = CELLS * (1 + CH);  // bulk copies per pass
        int4 cur = fetch(lane);
        for (int n0 = 0; n0 < nItems; n0 += 32) {
            const int4 nxt = fetch(n0 + 32 + lane);
            for (int pp = 0; pp < 32 / CELLS && n0 + pp * CELLS < nItems; ++pp) {
                for (int r0 = 0; r0 < perPass; r0 += 32) {
                    const int r = r0 + lane;
                    const int role = r / CELLS;  // 0: meta, 1 + ch: chunk ch
                    const int c = r - role * CELLS;
                    const int srcLane = pp * CELLS + c;
                    const uint32_t offLo = (uint32_t)__shfl_sync(0xffffffffu, cur.x, srcLane);
                    const uint32_t offHi = (uint32_t)__shfl_sync(0xffffffffu, cur.y, srcLane);
                    const int meta = __shfl_sync(0xffffffffu, cur.z, srcLane);
                    const int packed = __shfl_sync(0xffffffffu, cur.w, srcLane);
                    const int item = n0 + srcLane;
                    if (r < perPass && item < nItems) {
                        const unsigned char* src = a.records + (((uint64_t)offHi << 32) | offLo);
                        const int nSt = packed & 0xffff, Nmax = (packed >> 16) & 0xffff;
                        if (role == 0) {
                            const int slot = item % nMeta;
                            const uint32_t round = (uint32_t)(item / nMeta);
                            if (round > 0) mbarWait(barMetaEmpty + slot * 8, (round - 1) & 1u);
                            const uint32_t fb = barMetaFull + slot * 8;
                            mbarArriveExpectTx(fb, (uint32_t)meta);
                            tmaLoad1D(metaRing + (uint32_t)slot * a.metaSlotBytes, src, (uint32_t)meta,
                                      fb, pol);
                        } else {
                            const int ch = role - 1;
                            const int pass = item / CELLS;
                            const int q = (pass * CH + ch) * CELLS + c;
                            const int slot = q % nChunk;
                            const uint32_t round = (uint32_t)(q / nChunk);
                            if (round > 0) mbarWait(barChunkEmpty + slot * 8, (round - 1) & 1u);
                            const uint32_t fb = barChunkFull + slot * 8;
                            const int cols = min(D, Nmax - ch * D);
                            if (cols > 0) {
                                const uint32_t bytes = (uint32_t)chunkBytes(K, nSt, cols);
                                mbarArriveExpectTx(fb, bytes);
                                tmaLoad1D(chunkRing + (uint32_t)slot * a.chunkSlotBytes,
                                          src + meta + (size_t)ch * chunkBytes(K, nSt, D), bytes, fb,
                                          pol);
                            } else {
                                mbarArrive(fb);  // nothing to load (empty cell / short stencils)
                            }
                        }
                    }
                    __syncwarp();
                }
            }
            cur = nxt;
        }
        return;
    }
    if (warp > NW) return;

    // ===== consumers: CELLS cells per warp pass, one lane per (cell, stencil) =====
    const int cslot = lane / SS;
    const int s = lane - cslot * SS;
    const bool fit01 = a.scheme == 1;
    const bool fitBounded = a.scheme == 0 && a.bounded;

    for (int pass = warp; pass * CELLS < nItems; pass += NW) {
        const int item = pass * CELLS + cslot;
        const bool valid = item < nItems;
        const int mslot = item % nMeta;
        const uint32_t mround = (uint32_t)(item / nMeta);
        const uint32_t rec = metaRing + (uint32_t)mslot * a.metaSlotBytes;
        const int cell = valid ? itemToCell(item) : 0;
        double cen[CB];
        loadVals<CB>(a.vals + (size_t)cell * a.VS, cen);

        if (valid) slotWait(barMetaFull + mslot * 8, relMeta + mslot * 4, mround);
        __syncwarp();
        int nSt = 0, nFc = 0, Nmax = 0;
        uint32_t oIdx = 0, oFace = 0, oB = 0;
        if (valid) {
            const int4 h0 = lds128i(rec), h1 = lds128i(rec + 16);
            nSt = h0.x;
            nFc = h0.y;
            Nmax = h0.z;
            oIdx = h1.x;
            oFace = h1.y;
            oB = h1.z;
        }
#ifdef WENOEXT_KDEBUG
        if (valid && (nSt < 0 || nSt > SS || nFc < 0 || nFc > 64 || Nmax < 0 || Nmax > a.maxN ||
                      oIdx > (uint32_t)a.metaSlotBytes || oFace > (uint32_t)a.metaSlotBytes ||
                      oB > (uint32_t)a.metaSlotBytes)) {
            if (atomicCAS(a.dbg, 0, 1) == 0) {
                a.dbg[1] = item; a.dbg[2] = mslot; a.dbg[3] = (int)mround; a.dbg[4] = cell;
                a.dbg[5] = nSt; a.dbg[6] = nFc; a.dbg[7] = Nmax; a.dbg[8] = (int)oIdx;
                a.dbg[10] = blockIdx.x; a.dbg[11] = threadIdx.x; a.dbg[12] = nItems;
            }
            nSt = nFc = Nmax = 0;
        }
#endif
        const bool act = s < nSt;
        const int N = act ? lds32(rec + (uint32_t)sizeof(RecHeader) + s * 4) : 0;
        const int Nw = __reduce_max_sync(0xffffffffu, Nmax);

        // first face of this lane: issue its global loads now, consumed at the end
        int4 pfFace = make_int4(0, -1, 0, 0);
        double pfFlux = 0.0;
        double pfPsiN[CB];
#pragma unroll
        for (int c = 0; c < CB; ++c) pfPsiN[c] = 0.0;
        if (s < nFc) {
            pfFace = lds128i(rec + oFace + s * 16);
#ifdef WENOEXT_KDEBUG
            if (pfFace.x < 0 || pfFace.x >= a.nFaces || pfFace.y < -1 || pfFace.y >= a.nVals) {
                if (atomicCAS(a.dbg, 0, 3) == 0) {
                    a.dbg[1] = item; a.dbg[2] = mslot; a.dbg[3] = s; a.dbg[4] = cell;
                    a.dbg[5] = pfFace.x; a.dbg[6] = pfFace.y; a.dbg[7] = pfFace.z; a.dbg[8] = nFc;
                }
                pfFace = make_int4(0, 0, 0, 0);
            }
#endif
            pfFlux = __ldg(a.flux + pfFace.x);
            if (fitBounded && (pfFace.z >> FACE_TYPE_SHIFT) == FACE_INTERNAL)
                loadVals<CB>(a.vals + (size_t)pfFace.y * a.VS, pfPsiN);
        }

        // gather pipeline: column j of this lane's stencil (WENOCoeff.C:138-156)
        const uint32_t idxBase = rec + oIdx + s * 4;
        const uint32_t idxStride = (uint32_t)nSt * 4;
        auto issue = [&](int j, double (&g)[CB]) {
            if (j < N) {
                int id = lds32(idxBase + (uint32_t)j * idxStride);
#ifdef WENOEXT_KDEBUG
                if (id < 0 || id >= a.nVals) {
                    if (atomicCAS(a.dbg, 0, 2) == 0) {
                        a.dbg[1] = item; a.dbg[2] = mslot; a.dbg[3] = j; a.dbg[4] = cell;
                        a.dbg[5] = nSt; a.dbg[6] = id; a.dbg[7] = N; a.dbg[8] = s;
                    }
                    id = 0;
                }
#endif
                loadVals<CB>(a.vals + (size_t)id * a.VS, g);
            } else {
#pragma unroll
                for (int c = 0; c < CB; ++c) g[c] = cen[c];
            }
        };
        double nx[D][CB];
#pragma unroll
        for (int d = 0; d < D; ++d) issue(d, nx[d]);

        // phase 1: coeffs[k][c] = sum_j A_s[k][j] * b_s[j][c]        (WENOCoeff.C:159)
        double acc[K][CB];
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int c = 0; c < CB; ++c) acc[k][c] = 0.0;
        const uint32_t pairStride = (uint32_t)nSt * 16;
        // Fast path (uniform meshes): every active lane of the warp has the same number
        // of columns, a multiple of D -> no per-column predicate. Lanes without a
        // stencil read the chunk of the first lane that has one (finite data) against
        // b == 0.
        const bool uniform = __all_sync(0xffffffffu, N == Nw || N == 0) && (Nw % D) == 0;
        const uint32_t have = __ballot_sync(0xffffffffu, N > 0);
        const int srcLane = have ? __ffs(have) - 1 : 0;
        const uint32_t pS = __shfl_sync(0xffffffffu, pairStride, srcLane);
        const uint32_t pStr = (uniform && N == 0) ? pS : pairStride;
        for (int ch = 0; ch < CH; ++ch) {
            const int j0 = ch * D;
            double cu[D][CB];
#pragma unroll
            for (int d = 0; d < D; ++d)
#pragma unroll
                for (int c = 0; c < CB; ++c) cu[d][c] = nx[d][c] - cen[c];
#pragma unroll
            for (int d = 0; d < D; ++d) issue(j0 + D + d, nx[d]);
            // this cell's chunk ch
            const int q = (pass * CH + ch) * CELLS + cslot;
            const int cs = q % nChunk;
            const uint32_t cround = (uint32_t)(q / nChunk);
            if (valid) slotWait(barChunkFull + cs * 8, relChunk + cs * 4, cround);
            __syncwarp();
            uint32_t Ap = chunkRing + (uint32_t)cs * a.chunkSlotBytes + (uint32_t)s * 16;
            if (j0 < Nw) {
                if (uniform) {
                    const uint32_t ApS = __shfl_sync(0xffffffffu, Ap, srcLane);
                    if (N == 0) Ap = ApS;
                    uint32_t Sp = Ap - (N == 0 ? 0u : (uint32_t)s * 8) + (uint32_t)D * KP * pStr;
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        uint32_t ap = Ap;
#pragma unroll
                        for (int q2 = 0; q2 < KP; ++q2) {
                            const double2 av = lds128(ap);
                            ap += pStr;
#pragma unroll
                            for (int c = 0; c < CB; ++c) {
                                acc[2 * q2][c] = fma(av.x, cu[d][c], acc[2 * q2][c]);
                                acc[2 * q2 + 1][c] = fma(av.y, cu[d][c], acc[2 * q2 + 1][c]);
                            }
                        }
                        if constexpr (ODD) {
                            const double a1 = lds64(Sp);
#pragma unroll
                            for (int c = 0; c < CB; ++c)
                                acc[K - 1][c] = fma(a1, cu[d][c], acc[K - 1][c]);
                        }
                        Ap += pStr * KP;
                        Sp += pStr / 2;
                    }
                } else {
                    const int cols = min(D, Nmax - j0);  // columns stored in this chunk
                    uint32_t Sp = Ap - (uint32_t)s * 8 + (uint32_t)(cols > 0 ? cols : 0) * KP * pairStride;
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        if (j0 + d < N) {
                            uint32_t ap = Ap;
#pragma unroll
                            for (int q2 = 0; q2 < KP; ++q2) {
                                const double2 av = lds128(ap);
                                ap += pairStride;
#pragma unroll
                                for (int c = 0; c < CB; ++c) {
                                    acc[2 * q2][c] = fma(av.x, cu[d][c], acc[2 * q2][c]);
                                    acc[2 * q2 + 1][c] = fma(av.y, cu[d][c], acc[2 * q2 + 1][c]);
                                }
                            }
                            if constexpr (ODD) {
                                const double a1 = lds64(Sp);
#pragma unroll
                                for (int c = 0; c < CB; ++c)
                                    acc[K - 1][c] = fma(a1, cu[d][c], acc[K - 1][c]);
                            }
                        }
                        Ap += pairStride * KP;
                        Sp += pairStride / 2;
                    }
                }
            }
            __syncwarp();
            if (valid && s == 0) slotRelease(barChunkEmpty + cs * 8, relChunk + cs * 4, cround);
        }
        // phase 2: IS = c^T (B c) per component                  (WENOCoeff.C:190-196)
        double g[CB];
        {
            double IS[CB];
#pragma unroll
            for (int c = 0; c < CB; ++c) IS[c] = 0.0;
            const uint32_t Bb = rec + oB;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                double u[CB];
#pragma unroll
                for (int c = 0; c < CB; ++c) u[c] = 0.0;
#pragma unroll
                for (int q = 0; q < K; ++q) {
                    const double bq = lds64(Bb + (uint32_t)(k * K + q) * 8);
#pragma unroll
                    for (int c = 0; c < CB; ++c) u[c] = fma(bq, acc[q][c], u[c]);
                }
#pragma unroll
                for (int c = 0; c < CB; ++c) IS[c] = fma(acc[k][c], u[c], IS[c]);
            }
            // gamma: dm for the first surviving stencil          (WENOCoeff.C:198-205)
#pragma unroll
            for (int c = 0; c < CB; ++c)
                g[c] = act ? (s == 0 ? a.dm : 1.0) / intPowD(a.eps + IS[c], a.pExp) : 0.0;
        }
        // weighted sum over the stencils of the cell: butterfly over its SS lanes
        // (WENOCoeff.C:207-220; the division by gammaSum is applied per face below)
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int c = 0; c < CB; ++c) acc[k][c] *= g[c];
#pragma unroll
        for (int m = 1; m < SS; m <<= 1) {
#pragma unroll
            for (int k = 0; k < K; ++k)
#pragma unroll
                for (int c = 0; c < CB; ++c)
                    acc[k][c] += __shfl_xor_sync(0xffffffffu, acc[k][c], m);
#pragma unroll
            for (int c = 0; c < CB; ++c) g[c] += __shfl_xor_sync(0xffffffffu, g[c], m);
        }
        // now acc = sum_s gamma_s c_s and g = gammaSum on every lane of the cell

        // phase 3: faces, one (or more) per lane
        const uint32_t fdBase = rec + oB + (uint32_t)(K * K) * 8;
        const uint32_t quirkA = fdBase + (uint32_t)(nFc * (K + 1)) * 8;
        // sumFlux with this cell's weighted polynomial (WENOUpwindFit.C:293-323)
        // (sum_s gamma_s c_s) / gammaSum / refFacAr is evaluated with reciprocals
        // (one division per component and one per face instead of two per value)
        double invG[CB];
#pragma unroll
        for (int c = 0; c < CB; ++c) invG[c] = nSt > 0 ? 1.0 / g[c] : 0.0;  // Cell::empty -> zero
        auto polyAt = [&](uint32_t basis, double refFacAr, double (&out)[CB]) {
            double S[CB];
#pragma unroll
            for (int c = 0; c < CB; ++c) S[c] = 0.0;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const double bk = lds64(basis + k * 8);
#pragma unroll
                for (int c = 0; c < CB; ++c) S[c] = fma(acc[k][c], bk, S[c]);
            }
            const double rinv = 1.0 / refFacAr;
#pragma unroll
            for (int c = 0; c < CB; ++c) out[c] = (S[c] * invG[c]) * rinv;
        };
        double theta[CB];
#pragma unroll
        for (int c = 0; c < CB; ++c) theta[c] = 1.0;
        if (fit01) {
            // WENOUpwindFit01.C:66-161: extrema of the own polynomial over the
            // internal faces of the cell
            double mx[CB], mn[CB];
#pragma unroll
            for (int c = 0; c < CB; ++c) mx[c] = mn[c] = cen[c];
            for (int f = s; f < nFc; f += SS) {
                const int4 fi = f == s ? pfFace : lds128i(rec + oFace + f * 16);
                if ((fi.z >> FACE_TYPE_SHIFT) != FACE_INTERNAL) continue;
                const double fl = f == s ? pfFlux : __ldg(a.flux + fi.x);
                const uint32_t fd = fdBase + (uint32_t)(f * (K + 1)) * 8;
                const bool nei = fi.z & FACE_SIDE_NEI;
                const uint32_t basis = (nei && !(fl < 0) && a.bugCompat) ? quirkA : fd;
                double pv[CB];
                polyAt(basis, lds64(fd + K * 8), pv);
#pragma unroll
                for (int c = 0; c < CB; ++c) {
                    const double t = cen[c] + pv[c];
                    mx[c] = t > mx[c] ? t : mx[c];
                    mn[c] = t < mn[c] ? t : mn[c];
                }
            }
#pragma unroll
            for (int m = 1; m < SS; m <<= 1)
#pragma unroll
                for (int c = 0; c < CB; ++c) {
                    const double omx = __shfl_xor_sync(0xffffffffu, mx[c], m);
                    const double omn = __shfl_xor_sync(0xffffffffu, mn[c], m);
                    mx[c] = omx > mx[c] ? omx : mx[c];
                    mn[c] = omn < mn[c] ? omn : mn[c];
                }
#pragma unroll
            for (int c = 0; c < CB; ++c) {
                const double psi = cen[c];
                const double argMax =
                    fabs(mx[c] - psi) < 1e-9 ? 1.0 : fabs((1.0 - psi) / (mx[c] - psi));
                const double argMin =
                    fabs(mn[c] - psi) < 1e-9 ? 1.0 : fabs((0.0 - psi) / (mn[c] - psi));
                theta[c] = fmin(fmin(argMax, argMin), 1.0);
            }
        }
        for (int f = s; f < nFc; f += SS) {
            const int4 fi = f == s ? pfFace : lds128i(rec + oFace + f * 16);
            const double fl = f == s ? pfFlux : __ldg(a.flux + fi.x);
            const int faceId = fi.x, type = fi.z >> FACE_TYPE_SHIFT;
            const bool nei = fi.z & FACE_SIDE_NEI;
            // who writes: owner unless flux<0 (flux==0 -> owner writes zero), WENOUpwindFit.C:258-281
            const bool up = nei ? (fl < 0) : (fl > 0);
            const bool writer = type == FACE_INTERNAL ? (nei ? (fl < 0) : !(fl < 0)) : (fl > 0);
            if (fit01 && type == FACE_PROCESSOR && a.thetaPatch) {
#pragma unroll
                for (int c = 0; c < CB; ++c)
                    a.thetaPatch[(size_t)(faceId - a.nInternalFaces) * a.nComp + a.c0 + c] = theta[c];
            }
            if (!writer) continue;
            double out[CB];
#pragma unroll
            for (int c = 0; c < CB; ++c) out[c] = 0.0;
            if (up) {
                const uint32_t fd = fdBase + (uint32_t)(f * (K + 1)) * 8;
                polyAt(fd, lds64(fd + K * 8), out);
                if (type == FACE_INTERNAL) {
                    if (fit01) {
#pragma unroll
                        for (int c = 0; c < CB; ++c) out[c] = theta[c] * out[c];  // WENOUpwindFit01.C:167-181
                    } else if (a.bounded) {
                        // WENOUpwindFit.C:81-108, :172-179
                        double psiN[CB];
                        if (f == s) {
#pragma unroll
                            for (int c = 0; c < CB; ++c) psiN[c] = pfPsiN[c];
                        } else {
                            loadVals<CB>(a.vals + (size_t)fi.y * a.VS, psiN);
                        }
#pragma unroll
                        for (int c = 0; c < CB; ++c) {
                            const double dpsi = fabs(cen[c] - psiN[c]);
                            if (fabs(out[c]) > dpsi) out[c] = fmin(dpsi / fabs(out[c]), 1.0) * out[c];
                        }
                    }
                }
            }
            double* dst = a.corr + (size_t)faceId * a.nComp + a.c0;
#pragma unroll
            for (int c = 0; c < CB; ++c) dst[c] = out[c];
            if (a.faceValue && type == FACE_INTERNAL) {
                double* fv = a.faceValue + (size_t)faceId * a.nComp + a.c0;
#pragma unroll
                for (int c = 0; c < CB; ++c) fv[c] = cen[c] + out[c];
            }
        }
        __syncwarp();
        if (valid && s == 0) slotRelease(barMetaEmpty + mslot * 8, relMeta + mslot * 4, mround);
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

size_t cellKernelSmemBytes(const CellKernelArgs& a) {
    return (size_t)a.nMetaSlots * (a.metaSlotBytes + 20) + (size_t)a.nChunkSlots * (a.chunkSlotBytes + 20) + 64;
}

bool cellKernelSupportsK(int K) { return K == 5 || K == 9 || K == 14 || K == 19; }

namespace {
template <int K, int CB, int CELLS, int D>
int launchT(const CellKernelArgs& a, int gridBlocks, size_t smemBytes, cudaStream_t st) {
    static size_t configured = 0;  // per instantiation
    if (smemBytes > configured) {
        if (cudaFuncSetAttribute(weno_cell_kernel<K, CB, CELLS, D>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smemBytes) != cudaSuccess)
            return 1;
        configured = smemBytes;
    }
    const int threads = (a.nWarps + 1) * 32;
    weno_cell_kernel<K, CB, CELLS, D><<<gridBlocks, threads, smemBytes, st>>>(a);
    return 0;
}
template <int K, int D>
int launchK(const CellKernelArgs& a, int CB, int gridBlocks, size_t smemBytes, cudaStream_t st) {
    if (a.cellsPerWarp == 4) {
        switch (CB) {
            case 1: return launchT<K, 1, 4, D>(a, gridBlocks, smemBytes, st);
            case 2: return launchT<K, 2, 4, D>(a, gridBlocks, smemBytes, st);
            default: return launchT<K, 3, 4, D>(a, gridBlocks, smemBytes, st);
        }
    }
    switch (CB) {
        case 1: return launchT<K, 1, 1, D>(a, gridBlocks, smemBytes, st);
        case 2: return launchT<K, 2, 1, D>(a, gridBlocks, smemBytes, st);
        default: return launchT<K, 3, 1, D>(a, gridBlocks, smemBytes, st);
    }
}
}  // namespace

// K (the polynomial space) is a template parameter because every lane keeps the
// K x CB coefficients of its stencil in registers: 3-D order 2 and 2-D order 3: 9,
// 3-D order 3: 19, 2-D order 2: 5, 2-D order 4: 14.
int launchCellKernel(const CellKernelArgs& a, int CB, int gridBlocks, size_t smemBytes,
                     cudaStream_t st) {
    if ((a.nWarps + 1) * 32 > MAX_THREADS || (a.cellsPerWarp != 4 && a.cellsPerWarp != 1)) return 1;
    switch (a.K) {
        case 5: return launchK<5, 7>(a, CB, gridBlocks, smemBytes, st);
        case 9: return launchK<9, 7>(a, CB, gridBlocks, smemBytes, st);
        case 14: return launchK<14, 4>(a, CB, gridBlocks, smemBytes, st);
        case 19: return launchK<19, 4>(a, CB, gridBlocks, smemBytes, st);
        default: return 1;
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
