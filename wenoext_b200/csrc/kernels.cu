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
// values overlaps the arrival of the matrices. A group of G warps takes one cell:
// every lane owns one row (stencil, coefficient) of the stacked matrices, reads
// it conflict-free from the interleaved record and keeps CB component
// accumulators in registers. The kernel is HBM-bound: tensor cores are
// deliberately unused (independent 9x21 GEMVs, no operand reuse).
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_plan.h"

namespace wenoext {

namespace {

constexpr int TILE = 32;          // consecutive cells handed to one CTA at a time
constexpr int MAX_THREADS = 544;  // 16 consumer warps + producer

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
__device__ __forceinline__ void sts64(uint32_t a, double v) {
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory");
}
__device__ __forceinline__ void sts128(uint32_t a, double x, double y) {
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(a), "d"(x), "d"(y) : "memory");
}
// barrier of one cell group (G warps); G == 1 is a warp barrier
__device__ __forceinline__ void groupSync(int G, int grp) {
    if (G == 1)
        __syncwarp();
    else
        asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "r"(G * 32) : "memory");
}

__device__ __forceinline__ double intPowD(double base, uint32_t e) {
    double t = 1.0;
    for (uint32_t i = 0; i < e; ++i) t *= base;  // WENOCoeff.H:149-157
    return t;
}

__host__ __device__ inline size_t alignUp(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Per-group scratch: component vectors are stored as a (c0,c1) double2 plane plus
// a c2 plane so that one LDS.128 + one LDS.64 fetch three components.
template <int CB>
struct Vec {
    uint32_t p01, p2;  // shared addresses of the planes
    __device__ __forceinline__ void load(int i, double (&v)[CB]) const {
        if constexpr (CB == 1) {
            v[0] = lds64(p01 + i * 8);
        } else {
            double2 t = lds128(p01 + i * 16);
            v[0] = t.x;
            v[1] = t.y;
            if constexpr (CB == 3) v[2] = lds64(p2 + i * 8);
        }
    }
    __device__ __forceinline__ void store(int i, const double (&v)[CB]) const {
        if constexpr (CB == 1) {
            sts64(p01 + i * 8, v[0]);
        } else {
            sts128(p01 + i * 16, v[0], v[1]);
            if constexpr (CB == 3) sts64(p2 + i * 8, v[2]);
        }
    }
    __device__ __forceinline__ uint32_t addr(int i, int c) const {
        if constexpr (CB == 1) return p01 + i * 8;
        return c < 2 ? p01 + i * 16 + c * 8 : p2 + i * 8;
    }
};

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
template <int CB>
__device__ __forceinline__ double pick(const double (&v)[CB], int c) {
    double r = v[0];
    if constexpr (CB > 1) r = c == 1 ? v[1] : r;
    if constexpr (CB > 2) r = c == 2 ? v[2] : r;
    return r;
}

struct GroupScratch {
    uint32_t b01, b2, c01, c2, gam, W, poly, theta;
};
__host__ __device__ inline size_t scratchLayout(int K, int CB, int maxTotN, int maxRows, int maxSt,
                                                int maxFc, GroupScratch* g, uint32_t base) {
    const int nb = maxTotN > maxRows ? maxTotN : maxRows;
    size_t o = 0;
    const size_t w01 = CB == 1 ? 8 : 16;
    if (g) g->b01 = base + (uint32_t)o;
    o += alignUp((size_t)nb * w01, 16);
    if (g) g->b2 = base + (uint32_t)o;
    if (CB == 3) o += alignUp((size_t)nb * 8, 16);
    if (g) g->c01 = base + (uint32_t)o;
    o += alignUp((size_t)maxRows * w01, 16);
    if (g) g->c2 = base + (uint32_t)o;
    if (CB == 3) o += alignUp((size_t)maxRows * 8, 16);
    if (g) g->gam = base + (uint32_t)o;
    o += alignUp((size_t)maxSt * CB * 8, 16);
    if (g) g->W = base + (uint32_t)o;
    o += alignUp((size_t)K * CB * 8, 16);
    if (g) g->poly = base + (uint32_t)o;
    o += alignUp((size_t)maxFc * CB * 8, 16);
    if (g) g->theta = base + (uint32_t)o;
    o += 32;
    return o;
}

template <int KT, int CB>
__global__ void __launch_bounds__(MAX_THREADS, 1)
    weno_cell_kernel(const __grid_constant__ CellKernelArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int K = KT ? KT : a.K;
    const int G = a.G, L = G * 32, NG = a.nGroups;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nSlots = a.nSlots;

    const uint32_t ring = smemAddr(smem);
    const uint32_t bars = ring + (uint32_t)nSlots * a.slotBytes;  // per slot: fullMeta, fullA, empty
    const uint32_t scratch0 = (bars + (uint32_t)nSlots * 24 + 15u) & ~15u;

    if (threadIdx.x == 0) {
        for (int s = 0; s < 3 * nSlots; ++s) mbarInit(bars + s * 8, 1);
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
    auto itemToCell = [&](int n) {
        const int w = ((int)blockIdx.x + (n / TILE) * (int)gridDim.x) * TILE + (n % TILE);
        return a.cellList ? __ldg(a.cellList + w) : w;
    };

    if (warp == NG * G) {
        // ===== producer warp: all lanes prefetch the record descriptors of the next
        // 32 items, lane 0 waits for free slots and issues the bulk copies =====
        const uint64_t pol = evictFirstPolicy();
        auto fetch = [&](int n) {
            int4 r = make_int4(0, 0, 0, 0);
            if (n < nItems) r = __ldg(reinterpret_cast<const int4*>(a.recInfo + itemToCell(n)));
            return r;
        };
        int4 cur = fetch(lane);
        int slot = 0, round = 0;
        for (int n0 = 0; n0 < nItems; n0 += 32) {
            const int4 nxt = fetch(n0 + 32 + lane);
            const int cnt = min(32, nItems - n0);
            for (int i = 0; i < cnt; ++i) {
                const uint32_t offLo = (uint32_t)__shfl_sync(0xffffffffu, cur.x, i);
                const uint32_t offHi = (uint32_t)__shfl_sync(0xffffffffu, cur.y, i);
                const int tot = __shfl_sync(0xffffffffu, cur.z, i);
                const int meta = __shfl_sync(0xffffffffu, cur.w, i);
                if (lane == 0) {
                    const uint64_t off = ((uint64_t)offHi << 32) | offLo;
                    if (round > 0) mbarWait(bars + (slot * 3 + 2) * 8, (round - 1) & 1);
                    const uint32_t dst = ring + (uint32_t)slot * a.slotBytes;
                    const uint32_t bMeta = bars + (slot * 3) * 8, bA = bMeta + 8;
                    mbarArriveExpectTx(bMeta, (uint32_t)meta);
                    tmaLoad1D(dst, a.records + off, (uint32_t)meta, bMeta, pol);
                    const int aB = tot - meta;
                    if (aB > 0) {
                        mbarArriveExpectTx(bA, (uint32_t)aB);
                        tmaLoad1D(dst + meta, a.records + off + meta, (uint32_t)aB, bA, pol);
                    } else {
                        mbarArrive(bA);
                    }
                }
                if (++slot == nSlots) {
                    slot = 0;
                    ++round;
                }
            }
            cur = nxt;
        }
        return;
    }
    if (warp > NG * G) return;

    // ===== consumers: a group of G warps per cell, one lane per matrix row =====
    const int grp = warp / G;
    const int gl = (warp - grp * G) * 32 + lane;
    GroupScratch gs;
    {
        const size_t per = scratchLayout(K, CB, a.maxTotN, a.maxRows, a.maxSt, a.maxFc, nullptr, 0);
        scratchLayout(K, CB, a.maxTotN, a.maxRows, a.maxSt, a.maxFc, &gs,
                      scratch0 + (uint32_t)(grp * per));
    }
    const Vec<CB> bv{gs.b01, gs.b2};  // b values, later the IS products t
    const Vec<CB> cv{gs.c01, gs.c2};  // coefficients per (stencil, k)
    const int nBndBase = a.nInternalFaces;
    const bool fit01 = a.scheme == 1;
    const bool fitBounded = a.scheme == 0 && a.bounded;

    int slot = grp % nSlots, round = grp / nSlots;
    for (int n = grp; n < nItems; n += NG) {
        const int cell = itemToCell(n);
        double cen[CB];
        loadVals<CB>(a.vals + (size_t)cell * a.VS, cen);

        const uint32_t rec = ring + (uint32_t)slot * a.slotBytes;
        mbarWait(bars + (slot * 3) * 8, round & 1);
        const int nSt = lds32(rec), nFc = lds32(rec + 4), totN = lds32(rec + 8);
        const RecLayout lay = recLayout(K, L, nSt, totN, nFc);
        const int nRows = nSt * K;
        const int nFaceItems = nFc * CB;

        // face data of this lane's first (face, component) item: issue the global
        // loads now, they are consumed at the very end
        int pfFace = 0, pfOther = -1, pfKind = 0;
        double pfFlux = 0.0, pfPsiN = 0.0;
        if (gl < nFaceItems) {
            const int f = gl / CB, c = gl - f * CB;
            const uint32_t fa = rec + lay.oFaceI + f * 12;
            pfFace = lds32(fa);
            pfOther = lds32(fa + 4);
            pfKind = lds32(fa + 8);
            pfFlux = __ldg(a.flux + pfFace);
            if (fitBounded && (pfKind >> FACE_TYPE_SHIFT) == FACE_INTERNAL)
                pfPsiN = __ldg(a.vals + (size_t)pfOther * a.VS + c);
        }

        // gather: b[e][c] = psi(id_e)[c] - psi(cell)[c]      (WENOCoeff.C:138-156)
        for (int e = gl; e < totN; e += L) {
            const int id = lds32(rec + lay.oIdx + e * 4);
            double v[CB];
            loadVals<CB>(a.vals + (size_t)id * a.VS, v);
#pragma unroll
            for (int c = 0; c < CB; ++c) v[c] -= cen[c];
            bv.store(e, v);
        }
        groupSync(G, grp);
        mbarWait(bars + (slot * 3 + 1) * 8, round & 1);

        // phase 1: coeffs[s][k][c] = sum_j A_s[k][j] * b_s[j][c]   (WENOCoeff.C:159)
        for (int r0 = 0; r0 < nRows; r0 += L) {
            const int p = r0 + gl;
            if (p < nRows) {
                const int s = p / K;
                const int N = lds32(rec + lay.oN + s * 4);
                const int bo = lds32(rec + lay.oStOff + s * 4);
                const int rbase = lds32(rec + lay.oRnd + (r0 / L) * 8 + 4);
                const int R = min(L, nRows - r0);
                uint32_t Ap = rec + lay.metaBytes + (uint32_t)(rbase + gl) * 8;
                const uint32_t stride = (uint32_t)R * 8;
                double acc[CB];
#pragma unroll
                for (int c = 0; c < CB; ++c) acc[c] = 0.0;
#pragma unroll 7
                for (int j = 0; j < N; ++j) {
                    const double av = lds64(Ap);
                    Ap += stride;
                    double b[CB];
                    bv.load(bo + j, b);
#pragma unroll
                    for (int c = 0; c < CB; ++c) acc[c] = fma(av, b[c], acc[c]);
                }
                cv.store(p, acc);
            }
        }
        groupSync(G, grp);
        // phase 2a: t[s][k][c] = c_k * (B c)_k                  (WENOCoeff.C:190-196)
        for (int r0 = 0; r0 < nRows; r0 += L) {
            const int p = r0 + gl;
            if (p < nRows) {
                const int s = p / K, k = p - s * K;
                const uint32_t Bp = rec + lay.oB + (uint32_t)(k * K) * 8;
                double u[CB];
#pragma unroll
                for (int c = 0; c < CB; ++c) u[c] = 0.0;
#pragma unroll(KT ? KT : 4)
                for (int q = 0; q < K; ++q) {
                    const double bq = lds64(Bp + q * 8);
                    double cq[CB];
                    cv.load(s * K + q, cq);
#pragma unroll
                    for (int c = 0; c < CB; ++c) u[c] = fma(bq, cq[c], u[c]);
                }
                double own[CB];
                cv.load(p, own);
#pragma unroll
                for (int c = 0; c < CB; ++c) u[c] *= own[c];
                bv.store(p, u);
            }
        }
        groupSync(G, grp);
        // phase 2b: IS and gamma per (stencil, component)      (WENOCoeff.C:196-205)
        for (int i = gl; i < nSt * CB; i += L) {
            const int s = i / CB, c = i - s * CB;
            double IS = 0.0;
            for (int k = 0; k < K; ++k) IS += lds64(bv.addr(s * K + k, c));
            sts64(gs.gam + i * 8, (s == 0 ? a.dm : 1.0) / intPowD(a.eps + IS, a.pExp));
        }
        groupSync(G, grp);
        // phase 2c: weighted combination                        (WENOCoeff.C:207-220)
        for (int i = gl; i < K * CB; i += L) {
            const int k = i / CB, c = i - k * CB;
            double acc = 0.0, gsum = 0.0;
            for (int s = 0; s < nSt; ++s) {
                const double g = lds64(gs.gam + (s * CB + c) * 8);
                gsum += g;
                acc = fma(g, lds64(cv.addr(s * K + k, c)), acc);
            }
            sts64(gs.W + i * 8, nSt > 0 ? acc / gsum : 0.0);  // Cell::empty -> zero (WENOCoeff.C:333)
        }
        groupSync(G, grp);
        // phase 3: faces. Items (face, component) of the cell, one per lane.
        if (fit01) {
            // own polynomial at every internal face (WENOUpwindFit01.C:80-136)
            for (int i = gl; i < nFaceItems; i += L) {
                const int f = i / CB, c = i - f * CB;
                const uint32_t fa = rec + lay.oFaceI + f * 12;
                const int kind = lds32(fa + 8);
                double val = pick<CB>(cen, c);
                if ((kind >> FACE_TYPE_SHIFT) == FACE_INTERNAL) {
                    const uint32_t fd = rec + lay.oFaceD + (uint32_t)(f * (K + 1)) * 8;
                    const double fl = i == gl ? pfFlux : __ldg(a.flux + lds32(fa));
                    const bool nei = kind & FACE_SIDE_NEI;
                    const uint32_t basis =
                        (nei && !(fl < 0) && a.bugCompat) ? rec + lay.oQuirk : fd;
                    double s = 0.0;
                    for (int k = 0; k < K; ++k)
                        s = fma(lds64(gs.W + (k * CB + c) * 8), lds64(basis + k * 8), s);
                    val += s / lds64(fd + K * 8);
                }
                sts64(gs.poly + i * 8, val);
            }
            groupSync(G, grp);
            if (gl < CB) {
                const int c = gl;
                const double psi = pick<CB>(cen, c);
                double mx = psi, mn = psi;
                for (int f = 0; f < nFc; ++f) {
                    if ((lds32(rec + lay.oFaceI + f * 12 + 8) >> FACE_TYPE_SHIFT) != FACE_INTERNAL)
                        continue;
                    const double t = lds64(gs.poly + (f * CB + c) * 8);
                    if (t > mx)
                        mx = t;
                    else if (t < mn)
                        mn = t;
                }
                const double argMax = fabs(mx - psi) < 1e-9 ? 1.0 : fabs((1.0 - psi) / (mx - psi));
                const double argMin = fabs(mn - psi) < 1e-9 ? 1.0 : fabs((0.0 - psi) / (mn - psi));
                sts64(gs.theta + c * 8, fmin(fmin(argMax, argMin), 1.0));
            }
            groupSync(G, grp);
        }
        for (int i = gl; i < nFaceItems; i += L) {
            const int f = i / CB, c = i - f * CB;
            int faceId = pfFace, other = pfOther, kind = pfKind;
            double fl = pfFlux;
            if (i != gl) {
                const uint32_t fa = rec + lay.oFaceI + f * 12;
                faceId = lds32(fa);
                other = lds32(fa + 4);
                kind = lds32(fa + 8);
                fl = __ldg(a.flux + faceId);
            }
            const int type = kind >> FACE_TYPE_SHIFT;
            const bool nei = kind & FACE_SIDE_NEI;
            const double psi = pick<CB>(cen, c);
            // who writes: owner unless flux<0 (flux==0 -> owner writes zero), WENOUpwindFit.C:258-281
            const bool up = nei ? (fl < 0) : (fl > 0);
            const bool writer = type == FACE_INTERNAL ? (nei ? (fl < 0) : !(fl < 0)) : (fl > 0);
            const double th = fit01 ? lds64(gs.theta + c * 8) : 1.0;
            if (fit01 && type == FACE_PROCESSOR && a.thetaPatch)
                a.thetaPatch[(size_t)(faceId - nBndBase) * a.nComp + a.c0 + c] = th;
            if (!writer) continue;
            double out = 0.0;
            if (up) {
                const uint32_t fd = rec + lay.oFaceD + (uint32_t)(f * (K + 1)) * 8;
                double s = 0.0;
                for (int k = 0; k < K; ++k)
                    s = fma(lds64(gs.W + (k * CB + c) * 8), lds64(fd + k * 8), s);
                out = s / lds64(fd + K * 8);
                if (type == FACE_INTERNAL) {
                    if (fit01) {
                        out = th * out;  // WENOUpwindFit01.C:167-181
                    } else if (a.bounded) {
                        // WENOUpwindFit.C:81-108, :172-179
                        const double psiN =
                            i == gl ? pfPsiN : __ldg(a.vals + (size_t)other * a.VS + c);
                        const double dpsi = fabs(psi - psiN);
                        if (fabs(out) > dpsi) out = fmin(dpsi / fabs(out), 1.0) * out;
                    }
                }
            }
            a.corr[(size_t)faceId * a.nComp + a.c0 + c] = out;
            if (a.faceValue && type == FACE_INTERNAL)
                a.faceValue[(size_t)faceId * a.nComp + a.c0 + c] = psi + out;
        }
        groupSync(G, grp);
        if (gl == 0) mbarArrive(bars + (slot * 3 + 2) * 8);
        slot += NG;
        while (slot >= nSlots) {
            slot -= nSlots;
            ++round;
        }
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

size_t groupScratchBytes(int K, int CB, int maxTotN, int maxRows, int maxSt, int maxFc) {
    return scratchLayout(K, CB, maxTotN, maxRows, maxSt, maxFc, nullptr, 0);
}

size_t cellKernelSmemBytes(const CellKernelArgs& a, int CB) {
    size_t s = (size_t)a.nSlots * a.slotBytes + (size_t)a.nSlots * 24 + 16;
    s += (size_t)a.nGroups * groupScratchBytes(a.K, CB, a.maxTotN, a.maxRows, a.maxSt, a.maxFc);
    return s;
}

namespace {
template <int KT, int CB>
int launchT(const CellKernelArgs& a, int gridBlocks, size_t smemBytes, cudaStream_t st) {
    static size_t configured = 0;  // per instantiation
    if (smemBytes > configured) {
        if (cudaFuncSetAttribute(weno_cell_kernel<KT, CB>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smemBytes) != cudaSuccess)
            return 1;
        configured = smemBytes;
    }
    const int threads = (a.nGroups * a.G + 1) * 32;
    weno_cell_kernel<KT, CB><<<gridBlocks, threads, smemBytes, st>>>(a);
    return 0;
}
template <int KT>
int launchK(const CellKernelArgs& a, int CB, int gridBlocks, size_t smemBytes, cudaStream_t st) {
    switch (CB) {
        case 1: return launchT<KT, 1>(a, gridBlocks, smemBytes, st);
        case 2: return launchT<KT, 2>(a, gridBlocks, smemBytes, st);
        default: return launchT<KT, 3>(a, gridBlocks, smemBytes, st);
    }
}
}  // namespace

// K is a template parameter for the common polynomial spaces (3-D order 2: 9,
// 3-D order 3: 19, 2-D order 2: 5); anything else runs the runtime-K instance.
int launchCellKernel(const CellKernelArgs& a, int CB, int gridBlocks, size_t smemBytes,
                     cudaStream_t st) {
    if ((a.nGroups * a.G + 1) * 32 > MAX_THREADS) return 1;
    switch (a.K) {
        case 9: return launchK<9>(a, CB, gridBlocks, smemBytes, st);
        case 19: return launchK<19>(a, CB, gridBlocks, smemBytes, st);
        case 5: return launchK<5>(a, CB, gridBlocks, smemBytes, st);
        default: return launchK<0>(a, CB, gridBlocks, smemBytes, st);
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
