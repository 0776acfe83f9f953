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
// Structure: persistent CTAs (one per SM), 8 warps, each running its own pipeline.
// A warp takes tiles of 16 consecutive cells from a global counter and processes 4
// cells per pass, one lane per (cell, stencil). Lanes 0..3 issue 1-D TMA bulk copies
// (cp.async.bulk ... mbarrier::complete_tx, SASS UBLKCP) of the parts of "their"
// cell's record (device_plan.h) into the warp's private shared memory: head (this
// pass and the next), tail, and a ring of matrix chunk steps that is refilled by the
// warp itself as it consumes it. A lane gathers its own stencil's value differences
// straight into registers (global loads issued two chunk steps ahead), streams its
// matrix column by column from shared memory (128-bit loads, stencil index fastest, so
// the lanes of a cell read consecutive words) and keeps all K x CB coefficients in
// registers. Smoothness indicators and weights stay in registers; the weighted sum
// over the stencils of a cell is a warp-shuffle butterfly over the cell's lanes;
// afterwards every lane evaluates one face. No scratch shared memory, no block-level
// barriers after start-up, no cross-warp synchronisation. The kernel is HBM-bound:
// tensor cores are deliberately unused (independent 9x21 GEMVs, no operand reuse).
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <stdint.h>

#include "device_plan.h"

namespace wenoext {

namespace {

#ifndef WENOEXT_MAX_THREADS
#define WENOEXT_MAX_THREADS 256
#endif
constexpr int MAX_THREADS = WENOEXT_MAX_THREADS;  // 8 warps per CTA (up to 255 registers per thread)

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
// developer build: bounded wait, so that a protocol error shows up as wrong results and a
// report instead of a hang
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
    printf("[kdebug] mbarrier wait timed out: bar %u parity %u block %d thread %d\n", bar, parity,
           (int)blockIdx.x, (int)threadIdx.x);
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
__device__ __forceinline__ int32_t ldAcquireSys(const int32_t* p) {
    int32_t v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// 1/x for normal, finite x of moderate magnitude: hardware seed (about 20 bits) and two
// Newton steps, accurate to about 1 ulp; no special-case path. Used where the operand is a
// sum of weights or a face area; x outside [1e-290, 1e290] goes through the IEEE division.
__device__ __forceinline__ double fastRcp(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    return fma(r, e, r);
}
__device__ __forceinline__ double safeRcp(double x) {
    const double ax = fabs(x);
    return (ax > 1e-290 && ax < 1e290) ? fastRcp(x) : 1.0 / x;
}
__device__ __forceinline__ uint64_t evictFirstPolicy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// one lane of the (converged) warp
__device__ __forceinline__ bool electOne() {
    uint32_t p;
    asm volatile(
        "{\n.reg .pred q;\n"
        "elect.sync _|q, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, q;\n}\n"
        : "=r"(p));
    return p != 0;
}
// 2-D TMA tile copy global -> shared: box of the tensor map at element (c0, row c1), rows land
// densely one after the other; completion on an mbarrier
__device__ __forceinline__ void tmaLoad2D(uint32_t dst, const void* map, int c0, int c1, uint32_t bar,
                                          uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint "
        "[%0], [%1, {%2, %3}], [%4], %5;" ::"r"(dst),
        "l"(map), "r"(c0), "r"(c1), "r"(bar), "l"(policy)
        : "memory");
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

__device__ __forceinline__ uint64_t evictLastPolicy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// Cell values are re-read by every neighbouring stencil while 26 GB of records
// stream through the L2 once: keep them resident (evict_last) against the stream.
// The gathers use the coherent path (ld.global, no .nc): on parallel plans the halo region of `vals` is
// written by peer GPUs while the gated launch runs, and .nc is only defined for data that is read-only
// for the kernel's lifetime; the ld.acquire.sys on the halo flag orders these loads against the peers'
// writes. Same speed as .nc here (A/B in profiles/r02_kernel_experiments.md), and no per-load predicate.
template <int CB>
__device__ __forceinline__ void loadVals(const double* __restrict__ p, double (&v)[CB], uint64_t pol) {
    if constexpr (CB == 1) {
        asm volatile("ld.global.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v[0]) : "l"(p), "l"(pol));
    } else if constexpr (CB == 3) {
        // a 3-vector sits in a 32-byte slot (VS = 4): one 256-bit load, the pad is dropped
        double pad;
        asm volatile("ld.global.L2::cache_hint.v4.f64 {%0, %1, %2, %3}, [%4], %5;"
                     : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(pad)
                     : "l"(p), "l"(pol));
    } else {
        asm volatile("ld.global.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;" : "=d"(v[0]), "=d"(v[1]) : "l"(p), "l"(pol));
    }
}

// One elected-lane-free pipeline per warp: lanes 0..CELLS-1 each issue the bulk
// copies of "their" cell of the pass (head, tail, chunk steps); all lanes wait on
// the warp's own mbarriers. A buffer is refilled by the warp itself right after
// it has consumed it, so program order (plus __syncwarp) is the only "empty"
// signal needed: no producer warp, no cross-warp synchronisation.
template <int K, int CB, int CELLS, int D>
#ifdef WENOEXT_MAXNREG  // developer A/B switch: explicit register cap instead of ptxas' own choice
__global__ void __maxnreg__(WENOEXT_MAXNREG)
#else
__global__ void __launch_bounds__(MAX_THREADS, 1)
#endif
    weno_cell_kernel(const __grid_constant__ CellKernelArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int SS = 32 / CELLS;  // stencil lanes per cell (8, 6 with two idle lanes, or 32)
    constexpr int TILE = CELLS == 1 ? 16 : 4 * CELLS;  // consecutive work items a warp takes at a time
    constexpr int KP = K / 2;       // coefficient pairs per column
    constexpr bool ODD = (K & 1) != 0;
    constexpr bool LOOKAHEAD = K * CB <= 27;  // RecInfo of the pass three ahead loaded a pass early
    const int NW = a.nWarps;
    // broadcast through a shuffle: tells the compiler the warp index is warp-uniform
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int CH = a.nChunks, R = a.ringSteps;

    // shared memory of this warp: head[2][CELLS] | tail[CELLS] | ring[R][CELLS]; the
    // mbarriers of all warps follow the data: head x2, tail, ring x R per warp
    const uint32_t wbase = smemAddr(smem) + (uint32_t)warp * (uint32_t)a.warpBytes;
    const uint32_t headBuf = wbase;
    const uint32_t tailBuf = headBuf + 2u * CELLS * (uint32_t)a.headSlotBytes;
    const uint32_t ringBuf = tailBuf + (uint32_t)CELLS * (uint32_t)a.tailSlotBytes;
    const uint32_t bars = smemAddr(smem) + (uint32_t)NW * (uint32_t)a.warpBytes +
                          (uint32_t)warp * (uint32_t)(3 + R) * 8u;
    const uint32_t barTail = bars + 16, barRing = bars + 24;
    if (lane < 3 + R) mbarInit(bars + lane * 8, CELLS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    // Dynamic work split: a warp takes tiles of TILE consecutive work items from a global
    // counter (one tile ahead, so the atomic's latency is hidden). A CTA that starts late
    // (e.g. while a NCCL kernel holds its SM) simply takes fewer tiles.
    constexpr int PPT = TILE / CELLS;  // passes per tile
    // Two ranges of work items: A = [0, gateItem) (interior cells) and B = [gateItem, nWork)
    // (boundary cells, whose stencils gather halo values); without a gate everything is A.
    // Tiles of A come from tileCounter[0]. Once A is exhausted a warp looks at the halo flags
    // (one per sending rank, raised by that rank's halo_push_kernel over NVLink once its
    // values sit in this GPU's `vals`): all raised -> it goes on with tiles of B from
    // tileCounter[1]; not yet -> it stops, and the host's follow-up launch (behind
    // halo_wait_kernel, same counters, nGate == 0) takes what is left of B. The persistent
    // grid never spins on a flag.
    const bool split = a.gateItem >= 0;
    const int nA = split ? a.gateItem : a.nWork;
    const int nTilesA = (nA + TILE - 1) / TILE;
    const int nTiles = nTilesA + (split ? (a.nWork - nA + TILE - 1) / TILE : 0);
    // The atomic's result stays untouched in lane 0 (pendingTile) for half a tile before
    // it is broadcast, so nobody ever waits for its latency.
    auto grabTile = [&]() {
        int t = 0;
        if (lane == 0) t = atomicAdd(a.tileCounter, 1);
        return t;
    };
    // lane 0's raw counter value -> tile index on every lane (>= nTiles: no more work)
    auto resolveTile = [&](int raw) {
        int t = __shfl_sync(0xffffffffu, raw, 0);
        if (t < nTilesA) return t;
        // range B: lane q looks at the flag of sender q (acquire: the halo gathers that follow
        // must see what that sender pushed before it raised the flag)
        bool ok = split;
        if (lane < a.nGate) ok = ldAcquireSys(a.gateFlags + lane) >= a.gateEpoch;
        const bool ready = __all_sync(0xffffffffu, ok);
        if (lane == 0) t = ready ? nTilesA + atomicAdd(a.tileCounter + 1, 1) : nTiles;
        return __shfl_sync(0xffffffffu, t, 0);
    };
    // a halo that never arrived (halo_wait_kernel timed out): leave the boundary cells alone
#ifndef WENOEXT_NO_ABORT  // developer A/B switch
    if (a.abortFlag && *reinterpret_cast<const volatile int32_t*>(a.abortFlag) != 0) return;
#endif
    int curTile = resolveTile(grabTile());
    if (curTile >= nTiles) return;
    int pendingTile = grabTile();  // becomes nextTile at pass 1 of the current tile
    int nextTile = nTiles;
    static_assert(PPT >= 4, "the next tile is broadcast at pass 1 and first needed by the look-ahead of pass 1 (3 ahead)");
    int tp = 0;  // pass of the current tile being consumed
    // work item of cell c of the pass k passes after the one being consumed (k <= 2);
    // -1 beyond the end of the work
    auto workItem = [&](int k, int c) {
        const int t = tp + k;
        const int tile = t < PPT ? curTile : nextTile;
        const int off = (t & (PPT - 1)) * CELLS + c;
        const bool inA = tile < nTilesA;
        const int w = inA ? tile * TILE + off : nA + (tile - nTilesA) * TILE + off;
        return (tile < nTiles && w < (inA ? nA : a.nWork)) ? a.workBase + w : -1;
    };
    auto workToCell = [&](int w) { return a.cellList ? __ldg(a.cellList + w) : w; };

    // ---- issue side (lanes 0..CELLS-1; lane c serves cell c of a pass) ----
    const uint64_t pol = evictFirstPolicy();
    const uint64_t polV = evictLastPolicy();
    // row `id` of the packed values: one 32 x 32 -> 64 bit multiply-add on the base (ids are >= 0)
    const uint32_t valRowBytes = (uint32_t)a.VS * 8u;
    // (TRIM: the changes of round 2's last kernel pass apply to the instances with K * CB <= 27 only; the
    // K = 19 and larger instances sit at 255 registers with a stack frame, and every one of these changes
    // moved their spills to worse places: c4 7.36 -> 7.64 ms, profiles/r02_kernel_experiments.md)
    auto valRow = [&](int id) {
        if constexpr (K * CB <= 27)
            return reinterpret_cast<const double*>(reinterpret_cast<const unsigned char*>(a.vals) +
                                                   (uint64_t)(uint32_t)id * valRowBytes);
        else
            return a.vals + (size_t)id * a.VS;
    };
    const bool issuer = lane < CELLS;
    // RecInfo of cell `lane` of the pass k passes ahead; w < 0: no such cell
    // Fast passes: the CELLS cells are consecutive cells of the layout the tensor maps were
    // built for (plan.cu: buildTmaMaps). One elected lane then fetches each part of all of them
    // with ONE 2-D copy (box = part x CELLS rows of the record matrix); the rows land densely,
    // so the slots of such a pass are tmaBytes[part] apart instead of the padded slot strides.
    // fastOf: cell 0 of the pass, or -1 (per-lane 1-D copies, the general path).
    const bool tma2d = CELLS == 4 && a.tmaMaps != nullptr;
    // two halves, so that the load can be issued a whole pass before its result is looked at
    // (looking at it right away cost one full memory latency per pass: 4 % of the stall samples)
    auto loadInfo = [&](int k, int& cellIdx) {
        int4 v = make_int4(0, 0, 0, -1);
        const int w = issuer ? workItem(k, lane) : -1;
        cellIdx = -1;
        if (w >= 0) {
            cellIdx = workToCell(w);
            v = __ldg(reinterpret_cast<const int4*>(a.recInfo + cellIdx));
        }
        return v;
    };
    auto finishInfo = [&](const int4& v, int cellIdx, int& fastOf) {
        fastOf = -1;
        if (tma2d) {
            const int c0 = __shfl_sync(0xffffffffu, cellIdx, 0);
            const bool ok = !issuer || (cellIdx == c0 + lane && v.w == a.tmaW && (uint32_t)v.z == a.tmaZ);
            if (__all_sync(0xffffffffu, ok) && c0 >= 0) fastOf = c0;
        }
        return v;
    };
    auto fetchInfo = [&](int k, int& fastOf) {
        int cellIdx;
        const int4 v = loadInfo(k, cellIdx);
        return finishInfo(v, cellIdx, fastOf);
    };
    auto issue2D = [&](uint32_t dst, int part, int cell0, uint32_t bar) {
        if (issuer) mbarArriveExpectTx(bar, a.tmaBytes[part]);
        if (electOne())
            tmaLoad2D(dst, reinterpret_cast<const unsigned char*>(a.tmaMaps) + 128 * part, (int)a.tmaElemOff[part],
                      cell0, bar, pol);
    };
    auto recPtr = [&](const int4& ri) {
        return a.records + (((uint64_t)(uint32_t)ri.y << 32) | (uint32_t)ri.x);
    };
    auto issueHead = [&](const int4& ri, int fastOf, int hb) {
        if (fastOf >= 0) {
            issue2D(headBuf + (uint32_t)(hb * CELLS) * a.headSlotBytes, 0, fastOf, bars + hb * 8);
            return;
        }
        if (!issuer) return;
        const uint32_t bar = bars + hb * 8;
        if (ri.w >= 0) {
            const uint32_t bytes = (uint32_t)ri.z & 0xffffu;
            mbarArriveExpectTx(bar, bytes);
            tmaLoad1D(headBuf + (uint32_t)(hb * CELLS + lane) * a.headSlotBytes, recPtr(ri), bytes, bar, pol);
        } else {
            mbarArrive(bar);
        }
    };
    auto issueTail = [&](const int4& ri, int fastOf) {
        if (fastOf >= 0) {
            issue2D(tailBuf, 1, fastOf, barTail);
            return;
        }
        if (!issuer) return;
        const uint32_t bytes = ri.w >= 0 ? ((uint32_t)ri.z >> 16) : 0u;
        if (bytes) {
            mbarArriveExpectTx(barTail, bytes);
            tmaLoad1D(tailBuf + (uint32_t)lane * a.tailSlotBytes, recPtr(ri) + ((uint32_t)ri.z & 0xffffu),
                      bytes, barTail, pol);
        } else {
            mbarArrive(barTail);
        }
    };
    // chunk cursor: the next step to issue is chunk cCh of this warp's pass cPass
    int4 cInfo;
    const unsigned char* cSrc = nullptr;
    int cCh = 0, cPass = 0;
    uint32_t cFullBytes = 0;  // size of a chunk with D columns and D index rows
    int cNFull = 0;           // chunks [0, cNFull) are full
    int cFast = -1;  // the cursor's pass is a fast pass: its cell 0
    auto cursorEnter = [&](const int4& ri, int fastOf) {
        cFast = fastOf;
        cInfo = ri;
        cSrc = recPtr(ri) + ((uint32_t)ri.z & 0xffffu) + ((uint32_t)ri.z >> 16);
        cCh = 0;
        const int nSt = ri.w & 0xffff, Nmax = ri.w >> 16;
        cFullBytes = (uint32_t)((D * nSt * K * 8 + D * nSt * 4 + 15) & ~15);
        cNFull = ri.w >= 0 ? max(0, Nmax / D - GATHER_AHEAD) : 0;
    };
    auto issueChunk = [&](int slot) {
        if (cFast >= 0) {
            const uint32_t bar = barRing + slot * 8;
            if (a.tmaBytes[2 + cCh])
                issue2D(ringBuf + (uint32_t)(slot * CELLS) * a.chunkSlotBytes, 2 + cCh, cFast, bar);
            else if (issuer)
                mbarArrive(bar);
            ++cCh;
            return;
        }
        if (issuer) {
            const uint32_t bar = barRing + slot * 8;
            uint32_t bytes = cFullBytes;
            if (cCh >= cNFull)
                bytes = cInfo.w >= 0 ? (uint32_t)chunkBytes(K, cInfo.w & 0xffff, cInfo.w >> 16, cCh) : 0u;
            if (bytes) {
                mbarArriveExpectTx(bar, bytes);
                tmaLoad1D(ringBuf + (uint32_t)(slot * CELLS + lane) * a.chunkSlotBytes, cSrc, bytes, bar, pol);
                cSrc += bytes;
            } else {
                mbarArrive(bar);
            }
        }
        ++cCh;
    };

    int fCur, fNext, fNN;
    int4 infoCur = fetchInfo(0, fCur), infoNext = fetchInfo(1, fNext), infoNN = fetchInfo(2, fNN);
    issueHead(infoCur, fCur, 0);
    cursorEnter(infoCur, fCur);
    for (int r = 0; r < R; ++r) issueChunk(r);  // R <= CH
    if (cCh == CH) {
        cPass = 1;
        cursorEnter(infoNext, fNext);
    }
    // (cells beyond the end of the work have info.w < 0: their copies degenerate to plain
    // arrivals, so the pipeline can run past the last pass without anything in flight)

    // ---- consumer side: CELLS cells per pass, one lane per (cell, stencil) ----
    const int cslot = lane / SS;  // (== CELLS for the idle lanes of the 6-lane mapping)
    const int s = lane - cslot * SS;
    // lanes (s+1, s+2, s+3) mod SS of the own cell: all-reduce over a group of 6
    const int rot1 = (cslot * SS + (s + 1) % SS) & 31, rot2 = (cslot * SS + (s + 2) % SS) & 31,
              rot3 = (cslot * SS + (s + 3) % SS) & 31;
    // sum / max / min over the SS lanes of a cell, result on every lane of the cell
    auto groupSum = [&](double v) {
        if constexpr ((SS & (SS - 1)) == 0) {
#pragma unroll
            for (int m = 1; m < SS; m <<= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
            return v;
        } else {
            static_assert(SS == 6 || (SS & (SS - 1)) == 0, "lane groups of 6, 8 or 32");
            const double t = v + __shfl_sync(0xffffffffu, v, rot3);  // pairs (s, s+3)
            return t + __shfl_sync(0xffffffffu, t, rot1) + __shfl_sync(0xffffffffu, t, rot2);
        }
    };
    auto groupMax = [&](double v) {
        if constexpr ((SS & (SS - 1)) == 0) {
#pragma unroll
            for (int m = 1; m < SS; m <<= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, m));
            return v;
        } else {
            const double t = fmax(v, __shfl_sync(0xffffffffu, v, rot3));
            return fmax(t, fmax(__shfl_sync(0xffffffffu, t, rot1), __shfl_sync(0xffffffffu, t, rot2)));
        }
    };
    const bool fit01 = a.scheme == 1;
    const bool fitBounded = a.scheme == 0 && a.bounded;
    int rs = 0;           // ring slot of the next step to consume
    uint32_t rpar = 0;    // its mbarrier phase parity

    for (int i = 0;; ++i) {
        const int item = cslot < CELLS ? workItem(0, cslot) : -1;
        const bool valid = item >= 0;
        const int hb = i & 1;
        // slots of a fast pass are packed (the rows of a 2-D copy land densely)
        const bool fastCur = fCur >= 0;
        const uint32_t rec = headBuf + (uint32_t)(hb * CELLS) * a.headSlotBytes +
                             (uint32_t)cslot * (fastCur ? a.tmaBytes[0] : (uint32_t)a.headSlotBytes);
        const uint32_t tail = tailBuf + (uint32_t)cslot * (fastCur ? a.tmaBytes[1] : (uint32_t)a.tailSlotBytes);
        const int cell = valid ? workToCell(item) : 0;
        double cen[CB];
        loadVals<CB>(valRow(cell), cen, polV);
        // RecInfo of the pass three ahead: with the state of this pass it is workItem(3, .), which is
        // what fetchInfo(2, .) gives once tp has moved on at the bottom of the pass (the next tile is
        // resolved at tp == 1, before any pass needs it here)
        // (only where registers allow: the K = 19 and larger instances sit at the 255-register limit and
        // would pay for the five live registers with spills in the matrix loop)
        int cellN3 = -1;
        int4 rawN3 = make_int4(0, 0, 0, -1);
        if constexpr (LOOKAHEAD) rawN3 = loadInfo(3, cellN3);

        // the other head slot and the tail slot were last read by the previous pass
        issueHead(infoNext, fNext, hb ^ 1);
        issueTail(infoCur, fCur);

        mbarWait(bars + hb * 8, (uint32_t)(i >> 1) & 1u);
        int nSt = 0, nFc = 0, Nmax = 0;
        uint32_t oFace = 0, oIdx = 0;
        if (valid) {
            const int4 h0 = lds128i(rec), h1 = lds128i(rec + 16);
            nSt = h0.x;
            nFc = h0.y;
            Nmax = h0.z;
            oFace = h1.x;
            oIdx = h1.y;
        }
#ifdef WENOEXT_KDEBUG
        if (valid && (nSt < 0 || nSt > SS || nFc < 0 || nFc > 64 || Nmax < 0 || Nmax > a.maxN ||
                      oIdx > (uint32_t)a.headSlotBytes || oFace > (uint32_t)a.headSlotBytes)) {
            if (atomicCAS(a.dbg, 0, 1) == 0) {
                a.dbg[1] = item; a.dbg[2] = hb; a.dbg[3] = i; a.dbg[4] = cell;
                a.dbg[5] = nSt; a.dbg[6] = nFc; a.dbg[7] = Nmax; a.dbg[8] = (int)oIdx;
                a.dbg[10] = blockIdx.x; a.dbg[11] = threadIdx.x; a.dbg[12] = curTile;
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
                    a.dbg[1] = item; a.dbg[2] = hb; a.dbg[3] = s; a.dbg[4] = cell;
                    a.dbg[5] = pfFace.x; a.dbg[6] = pfFace.y; a.dbg[7] = pfFace.z; a.dbg[8] = nFc;
                }
                pfFace = make_int4(0, 0, 0, 0);
            }
#endif
            pfFlux = __ldg(a.flux + pfFace.x);
            if (fitBounded && (pfFace.z >> FACE_TYPE_SHIFT) == FACE_INTERNAL)
                loadVals<CB>(valRow(pfFace.y), pfPsiN, polV);
        }

        // gather of column j of this lane's stencil (WENOCoeff.C:138-156), branch-free:
        // lanes without a stencil and columns beyond the cell's stencils load the cell's
        // own value (difference exactly 0)
        const uint32_t rowBytes = (uint32_t)nSt * 4;
        auto gather = [&](uint32_t idxBlock, int j0, double (&g)[D][CB]) {
#pragma unroll
            for (int d = 0; d < D; ++d) {
                int id = cell;
                if (act && j0 + d < Nmax) id = lds32(idxBlock + (uint32_t)d * rowBytes + s * 4);
#ifdef WENOEXT_KDEBUG
                if (id < 0 || id >= a.nVals) {
                    if (atomicCAS(a.dbg, 0, 2) == 0) {
                        a.dbg[1] = item; a.dbg[2] = j0 + d; a.dbg[3] = i; a.dbg[4] = cell;
                        a.dbg[5] = nSt; a.dbg[6] = id; a.dbg[7] = N; a.dbg[8] = s;
                    }
                    id = 0;
                }
#endif
                loadVals<CB>(valRow(id), g[d], polV);
            }
        };
        // two buffers: the columns of step ch are gathered GATHER_AHEAD = 2 steps earlier
        static_assert(GATHER_AHEAD == 2, "the chunk loop below alternates two gather buffers");
        double nxA[D][CB], nxB[D][CB];
        gather(rec + oIdx, 0, nxA);
        gather(rec + oIdx + (uint32_t)D * rowBytes, D, nxB);

        // phase 1: coeffs[k][c] = sum_j A_s[k][j] * b_s[j][c]        (WENOCoeff.C:159)
        double acc[K][CB];
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int c = 0; c < CB; ++c) acc[k][c] = 0.0;
        const uint32_t pairStride = (uint32_t)nSt * 16;
        // Uniform meshes: every active lane of the warp has the same number of columns.
        // Lanes without a stencil then read the matrix of the first stencil of their own
        // cell (same shared-memory words as their neighbours, finite data) against b == 0.
        const bool uniform = __all_sync(0xffffffffu, N == Nw || N == 0);
        const uint32_t have = __ballot_sync(0xffffffffu, N > 0);
        const int own0 = cslot * SS;
        const int srcLane = ((have >> own0) & 1u) ? own0 : (have ? __ffs(have) - 1 : 0);
        const bool borrow = uniform && N == 0;
        const uint32_t pS = __shfl_sync(0xffffffffu, pairStride, srcLane);
        const uint32_t pStr = borrow ? pS : pairStride;
        // whose matrix this lane reads: its own (cell slot, stencil), or the lender's; on a
        // uniform mesh the lender's cell has Nmax == Nw
        const int cslotE = borrow ? srcLane / SS : cslot;
        const uint32_t sE = (uint32_t)(borrow ? srcLane % SS : s);
        const int NmaxE = borrow ? Nw : Nmax;
        auto step = [&](const int ch, double (&nx)[D][CB]) {
            const int j0 = ch * D;
            double cu[D][CB];
#pragma unroll
            for (int d = 0; d < D; ++d)
#pragma unroll
                for (int c = 0; c < CB; ++c) cu[d][c] = nx[d][c] - cen[c];
            // this pass's chunk step ch
            mbarWait(barRing + rs * 8, rpar);
            const uint32_t cStride = fastCur ? a.tmaBytes[2 + ch] : (uint32_t)a.chunkSlotBytes;
            const uint32_t ringSlot = ringBuf + (uint32_t)(rs * CELLS) * a.chunkSlotBytes;
            const uint32_t chunk = ringSlot + (uint32_t)cslot * cStride;
            const int cols = max(0, min(D, Nmax - j0));  // columns stored in this cell's chunk
            // gathers of the columns two steps ahead (their indices travel with this chunk)
            if (j0 + GATHER_AHEAD * D < Nw)
                gather(chunk + (uint32_t)chunkIdxOffset(K, nSt, cols), j0 + GATHER_AHEAD * D, nx);
            const uint32_t chunkE = ringSlot + (uint32_t)cslotE * cStride;
            uint32_t Ap = chunkE + sE * 16;
            uint32_t Sp = chunkE + (uint32_t)max(0, min(D, NmaxE - j0)) * KP * pStr + sE * 8;
            const int colsW = min(D, Nw - j0);  // warp-wide column count of this step
            if (uniform && colsW == D) {
                // full step, no predicates
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
                        for (int c = 0; c < CB; ++c) acc[K - 1][c] = fma(a1, cu[d][c], acc[K - 1][c]);
                    }
                    Ap += pStr * KP;
                    Sp += pStr / 2;
                }
            } else if (colsW > 0) {
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    if (uniform ? (d < colsW) : (j0 + d < N)) {
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
                    }
                    Ap += pStr * KP;
                    Sp += pStr / 2;
                }
            }
            __syncwarp();
            // refill the slot just consumed with the step R ahead in this warp's stream
            issueChunk(rs);
            if (cCh == CH) {
                ++cPass;
                if (cPass - i == 1)
                    cursorEnter(infoNext, fNext);
                else
                    cursorEnter(infoNN, fNN);
            }
            if (++rs == R) {
                rs = 0;
                rpar ^= 1u;
            }
        };
        for (int ch = 0; ch < CH; ch += 2) {
            step(ch, nxA);
            if (ch + 1 < CH) step(ch + 1, nxB);
        }
#ifdef WENOEXT_PREFETCH_NEXT
        // The first gathers of a pass wait a whole L2 round trip right behind the head (the top
        // long_scoreboard line of every capture). The next pass's head has been in flight since the
        // start of this pass: pull the cell values its first WENOEXT_PREFETCH_NEXT columns will gather
        // into L1 now, under the tail phase (no registers live across it).
        {
            const int itemN = cslot < CELLS ? workItem(1, cslot) : -1;
            mbarWait(bars + (hb ^ 1) * 8, (uint32_t)((i + 1) >> 1) & 1u);
            if (itemN >= 0) {
                const uint32_t recN = headBuf + (uint32_t)((hb ^ 1) * CELLS) * a.headSlotBytes +
                                      (uint32_t)cslot * (fNext >= 0 ? a.tmaBytes[0] : (uint32_t)a.headSlotBytes);
                const int4 h0 = lds128i(recN), h1 = lds128i(recN + 16);
                if (s < h0.x) {
#pragma unroll
                    for (int d = 0; d < WENOEXT_PREFETCH_NEXT; ++d)
                        if (d < h0.z) {
                            const int id = lds32(recN + (uint32_t)h1.y + (uint32_t)(d * h0.x + s) * 4);
                            asm volatile("prefetch.global.L1 [%0];" ::"l"(a.vals + (size_t)id * a.VS));
                        }
                }
                if (s == 0) asm volatile("prefetch.global.L1 [%0];" ::"l"(a.vals + (size_t)workToCell(itemN) * a.VS));
            }
        }
#endif
        // phase 2: IS = c^T (B c) per component                  (WENOCoeff.C:190-196)
        mbarWait(barTail, (uint32_t)i & 1u);
        double g[CB];
        {
            double IS[CB];
#pragma unroll
            for (int c = 0; c < CB; ++c) IS[c] = 0.0;
            // B arrives folded (diagonal, then B[k][q] + B[q][k] for q > k, device_plan.h):
            // IS = sum_k c_k (B_kk c_k + sum_{q>k} (B_kq + B_qk) c_q), K(K+1)/2 products per
            // component instead of K^2
            // (the triangle is read pairwise, 128 bits at a time: t counts its entries in storage order)
            double2 bpair = make_double2(0.0, 0.0);
            int t = 0;
            uint32_t Bb = tail;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                double u[CB];
#pragma unroll
                for (int c = 0; c < CB; ++c) u[c] = 0.0;
#pragma unroll
                for (int q = k; q < K; ++q) {
                    double bq;
                    if constexpr (LOOKAHEAD) {
                        if ((t & 1) == 0)
                            bpair = t + 1 < symTriSize(K) ? lds128(tail + (uint32_t)t * 8)
                                                           : make_double2(lds64(tail + (uint32_t)t * 8), 0.0);
                        bq = (t & 1) ? bpair.y : bpair.x;
                    } else {
                        bq = lds64(Bb);  // one word at a time off a running pointer
                        Bb += 8;
                    }
                    ++t;
#pragma unroll
                    for (int c = 0; c < CB; ++c) u[c] = fma(bq, acc[q][c], u[c]);
                }
#pragma unroll
                for (int c = 0; c < CB; ++c) IS[c] = fma(acc[k][c], u[c], IS[c]);
            }
            // gamma: dm for the first surviving stencil          (WENOCoeff.C:198-205)
            // intPow (WENOCoeff.H:149-157): repeated multiplication in the reference's order
            double pw[CB], bs[CB];
#pragma unroll
            for (int c = 0; c < CB; ++c) {
                bs[c] = a.eps + IS[c];
                pw[c] = 1.0;
            }
            if (LOOKAHEAD && a.pExp == 4) {
                // the default p: the same products in the same order (1.0 * bs is exact), no loop
#pragma unroll
                for (int c = 0; c < CB; ++c) pw[c] = ((bs[c] * bs[c]) * bs[c]) * bs[c];
            } else {
                for (uint32_t e = 0; e < a.pExp; ++e)
#pragma unroll
                    for (int c = 0; c < CB; ++c) pw[c] *= bs[c];
            }
#pragma unroll
            for (int c = 0; c < CB; ++c) g[c] = act ? (s == 0 ? a.dm : 1.0) * safeRcp(pw[c]) : 0.0;
        }
        // weighted sum over the stencils of the cell: butterfly over its SS lanes
        // (WENOCoeff.C:207-220; the division by gammaSum is applied per face below)
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int c = 0; c < CB; ++c) acc[k][c] *= g[c];
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int c = 0; c < CB; ++c) acc[k][c] = groupSum(acc[k][c]);
#pragma unroll
        for (int c = 0; c < CB; ++c) g[c] = groupSum(g[c]);
        // now acc = sum_s gamma_s c_s and g = gammaSum on every lane of the cell

        // phase 3: faces, one (or more) per lane
        const uint32_t fdBase = tail + (uint32_t)symTriSize(K) * 8;
        const uint32_t quirkA = fdBase + (uint32_t)(nFc * (K + 1)) * 8;
        // sumFlux with this cell's weighted polynomial (WENOUpwindFit.C:293-323)
        // (sum_s gamma_s c_s) / gammaSum / refFacAr is evaluated with reciprocals
        // (one division per component and one per face instead of two per value)
        double invG[CB];
#pragma unroll
        for (int c = 0; c < CB; ++c) invG[c] = nSt > 0 ? safeRcp(g[c]) : 0.0;  // Cell::empty -> zero
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
            const double rinv = safeRcp(refFacAr);
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
            for (int c = 0; c < CB; ++c) {
                mx[c] = groupMax(mx[c]);
                mn[c] = -groupMax(-mn[c]);
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
                            loadVals<CB>(valRow(fi.y), psiN, polV);
                        }
#pragma unroll
                        for (int c = 0; c < CB; ++c) {
                            const double dpsi = fabs(cen[c] - psiN[c]);
                            // theta = min(dpsi / |corr|, 1) with |corr| > dpsi: corr * theta is corr's sign on dpsi
                            // (the product differs from it by rounding only; no division, no reciprocal)
                            if constexpr (LOOKAHEAD) {
                                if (fabs(out[c]) > dpsi) out[c] = copysign(dpsi, out[c]);
                            } else {
                                if (fabs(out[c]) > dpsi) out[c] = fmin(dpsi * safeRcp(fabs(out[c])), 1.0) * out[c];
                            }
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
        if (++tp == PPT) {
            tp = 0;
            curTile = nextTile;
            if (curTile >= nTiles) break;
            pendingTile = grabTile();
        } else if (tp == 1) {
            nextTile = resolveTile(pendingTile);
        }
        infoCur = infoNext;
        infoNext = infoNN;
        fCur = fNext;
        fNext = fNN;
        if constexpr (LOOKAHEAD)
            infoNN = finishInfo(rawN3, cellN3, fNN);
        else
            infoNN = fetchInfo(2, fNN);
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

// dst[i][:] = src[rows[i]][:] (faces written late by boundary cells, sent to the host compactly)
__global__ void gather_rows_kernel(const double* __restrict__ src, const int32_t* __restrict__ rows,
                                   int64_t n, int nComp, double* __restrict__ dst) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * nComp) return;
    const int64_t r = i / nComp;
    dst[i] = src[(size_t)rows[r] * nComp + (i - r * nComp)];
}

// fvc::surfaceIntegrate(faceFlux * interpolate(vf)) (wenoext_div): one thread per
// (cell, component), faces in ascending index so the sum has OpenFOAM's order.
__global__ void surface_integrate_kernel(const DivKernelArgs a) {
    // (cells [cell0, cell0 + nCells): the pipelined host-buffer path integrates the cells of a slab as soon as
    // their faces are final)
    const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i0 >= (int64_t)a.nCells * a.nComp) return;
    const int64_t i = i0 + (int64_t)a.cell0 * a.nComp;
    const int cell = (int)(i / a.nComp), c = (int)(i - (int64_t)cell * a.nComp);
    const double own = a.vf[i];
    double sum = 0.0;
    for (int k = a.cellFaceStart[cell]; k < a.cellFaceStart[cell + 1]; ++k) {
        const int f = a.cellFaces[k];
        const double fl = a.flux[f];
        if (f < a.nInternalFaces) {
            // interpolate(): weights()*vf_P + (1 - weights())*vf_N + correction(), weights() = pos(flux)
            const int P_ = a.owner[f], N_ = a.neighbour[f];
            const double up = a.vf[(size_t)(fl >= 0 ? P_ : N_) * a.nComp + c];
            const double v = fl * (up + a.corr[(size_t)f * a.nComp + c]);
            sum += P_ == cell ? v : -v;
        } else {
            const int bf = f - a.nInternalFaces;
            double val;
            if (a.bfPatchType[bf] >= 0) {
                // coupled patch: weights() = pos(flux), value across from the caller or the halo
                double nb = own;
                if (a.patchNbrField)
                    nb = a.patchNbrField[(size_t)bf * a.nComp + c];
                else if (a.bfNbrCell && a.bfNbrCell[bf] >= 0 && a.bfNbrCell[bf] < a.nCells)
                    nb = a.vf[(size_t)a.bfNbrCell[bf] * a.nComp + c];  // cyclic: a local cell
                else if (a.bfNbrCell && a.bfNbrCell[bf] >= 0 && c >= a.c0 && c < a.c0 + a.CB)
                    nb = a.vals[(size_t)a.bfNbrCell[bf] * a.VS + (c - a.c0)];  // halo value
                val = (fl > 0 ? own : nb) + a.corr[(size_t)f * a.nComp + c];
            } else {
                val = a.bndValue[(size_t)bf * a.nComp + c];
            }
            sum += fl * val;
        }
    }
    a.div[i] = sum / a.cellVolumes[cell];
}

}  // namespace

size_t cellKernelSmemBytes(const CellKernelArgs& a) {
    return (size_t)a.nWarps * ((size_t)a.warpBytes + (size_t)(3 + a.ringSteps) * 8);
}

// nDvt of the supported polynomial spaces (WENOBase.C:1133-1140): 3-D orders 1-4 are 3, 9,
// 19, 34; 2-D orders 1-5 are 2, 5, 9, 14, 20
bool cellKernelSupportsK(int K) {
    return K == 2 || K == 3 || K == 5 || K == 9 || K == 14 || K == 19 || K == 20 || K == 34;
}
// components batched per matrix read: the K x CB coefficients live in registers
int cellKernelMaxCB(int K) { return K <= 20 ? 3 : 1; }
int cellKernelMaxWarps() { return MAX_THREADS / 32; }

namespace {
template <int K, int CB, int CELLS, int D>
int launchT(const CellKernelArgs& a, int gridBlocks, size_t smemBytes, cudaStream_t st) {
    // the opt-in to large dynamic shared memory is per function AND per device
    constexpr int MAX_DEVICES = 64;
    static size_t configured[MAX_DEVICES] = {0};  // per instantiation
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 1;
    if (dev < 0 || dev >= MAX_DEVICES || smemBytes > configured[dev]) {
        if (cudaFuncSetAttribute(weno_cell_kernel<K, CB, CELLS, D>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smemBytes) != cudaSuccess)
            return 1;
        if (dev >= 0 && dev < MAX_DEVICES) configured[dev] = smemBytes;
    }
    const int threads = a.nWarps * 32;
    weno_cell_kernel<K, CB, CELLS, D><<<gridBlocks, threads, smemBytes, st>>>(a);
    return 0;
}
template <int K, int D>
int launchK(const CellKernelArgs& a, int CB, int gridBlocks, size_t smemBytes, cudaStream_t st) {
    if constexpr (K > 20) {  // cellKernelMaxCB(K) == 1
        if (CB != 1) return 1;
        if (a.cellsPerWarp == 5) return launchT<K, 1, 5, D>(a, gridBlocks, smemBytes, st);
        if (a.cellsPerWarp == 4) return launchT<K, 1, 4, D>(a, gridBlocks, smemBytes, st);
        return launchT<K, 1, 1, D>(a, gridBlocks, smemBytes, st);
    } else {
        if (a.cellsPerWarp == 5) {
            switch (CB) {
                case 1: return launchT<K, 1, 5, D>(a, gridBlocks, smemBytes, st);
                case 2: return launchT<K, 2, 5, D>(a, gridBlocks, smemBytes, st);
                default: return launchT<K, 3, 5, D>(a, gridBlocks, smemBytes, st);
            }
        }
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
}
}  // namespace

// K (the polynomial space) is a template parameter because every lane keeps the
// K x CB coefficients of its stencil in registers (cellKernelSupportsK lists the spaces).
int launchCellKernel(const CellKernelArgs& a, int CB, int gridBlocks, size_t smemBytes,
                     cudaStream_t st) {
    if (a.nWarps * 32 > MAX_THREADS || a.nWarps < 1 || a.ringSteps < 1 || a.ringSteps > a.nChunks ||
        a.ringSteps > 27 || (a.cellsPerWarp != 5 && a.cellsPerWarp != 4 && a.cellsPerWarp != 1))
        return 1;
    switch (a.K) {
#ifdef WENOEXT_DEV_ONLY_K  // developer builds: one polynomial space only (compile time)
        case WENOEXT_DEV_ONLY_K: return launchK<WENOEXT_DEV_ONLY_K, chunkCols(WENOEXT_DEV_ONLY_K)>(a, CB, gridBlocks, smemBytes, st);
        default: return 1;
    }
#else
        case 2: return launchK<2, chunkCols(2)>(a, CB, gridBlocks, smemBytes, st);
        case 3: return launchK<3, chunkCols(3)>(a, CB, gridBlocks, smemBytes, st);
        case 5: return launchK<5, chunkCols(5)>(a, CB, gridBlocks, smemBytes, st);
        case 9: return launchK<9, chunkCols(9)>(a, CB, gridBlocks, smemBytes, st);
        case 14: return launchK<14, chunkCols(14)>(a, CB, gridBlocks, smemBytes, st);
        case 19: return launchK<19, chunkCols(19)>(a, CB, gridBlocks, smemBytes, st);
        case 20: return launchK<20, chunkCols(20)>(a, CB, gridBlocks, smemBytes, st);
        case 34: return launchK<34, chunkCols(34)>(a, CB, gridBlocks, smemBytes, st);
        default: return 1;
    }
#endif
}

void launchGatherRows(const double* src, const int32_t* rows, int64_t n, int nComp, double* dst,
                      cudaStream_t st) {
    if (n <= 0) return;
    gather_rows_kernel<<<(unsigned)((n * nComp + 255) / 256), 256, 0, st>>>(src, rows, n, nComp, dst);
}

// dst[rows[i]][:] = src[rows[i]][:] — dst may be page-locked HOST memory (zero-copy stores over the
// host link): puts the few faces that boundary cells finalise late straight into the caller's result
// buffer, behind the bulk copies of the ranges that contain them
__global__ void scatter_rows_kernel(const double* __restrict__ src, const int32_t* __restrict__ rows, int64_t n,
                                    int nComp, double* __restrict__ dst) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * nComp) return;
    const int64_t r = i / nComp;
    const size_t at = (size_t)rows[r] * nComp + (size_t)(i - r * nComp);
    dst[at] = src[at];
}
void launchScatterRows(const double* src, const int32_t* rows, int64_t n, int nComp, double* dst,
                       cudaStream_t st) {
    if (n <= 0) return;
    scatter_rows_kernel<<<(unsigned)((n * nComp + 255) / 256), 256, 0, st>>>(src, rows, n, nComp, dst);
}

void launchDivKernel(const DivKernelArgs& a, cudaStream_t st) {
    const int64_t n = (int64_t)a.nCells * a.nComp;
    if (n <= 0) return;
    surface_integrate_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a);
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

// Exchange #1 as a one-sided push (one process per GPU, peers' buffers mapped with CUDA
// IPC): sendHaloData of WENOCoeff.C:254-258 is written straight into each receiver's value
// buffer over NVLink, behind its local cells, where the receiver's gathers expect it. The
// last block to finish raises the receivers' flags (threadfence + block counter), so a
// raised flag means every value of this step has arrived.
__global__ void halo_push_kernel(const __grid_constant__ HaloPushArgs a) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t total = a.off[a.nPeers] * a.VS;
    if (i < total) {
        const int64_t k = i / a.VS;
        const int c = (int)(i - k * a.VS);
        int p = 0;
        while (k >= a.off[p + 1]) ++p;
        const int64_t j = k - a.off[p];
        // sendIds == nullptr: contiguous rows starting at src0[p] (exchange #2: a processor patch of corr)
        const int64_t row = a.sendIds ? (int64_t)a.sendIds[a.src0[p] + j] : a.src0[p] + j;
        a.dst[p][j * a.VS + c] = a.vals[(size_t)row * a.VS + c];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(a.done, 1) == (int)gridDim.x - 1) {
        *a.done = 0;
        __threadfence_system();
        for (int p = 0; p < a.nPeers; ++p)
            asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(a.flag[p]), "r"(a.epoch) : "memory");
    }
}
void launchHaloPush(const HaloPushArgs& a, cudaStream_t st) {
    const int64_t total = a.off[a.nPeers] * a.VS;
    halo_push_kernel<<<(unsigned)std::max<int64_t>(1, (total + 255) / 256), 256, 0, st>>>(a);
}

// Host-side stream order behind the halo: one lane per sending rank waits for its flag.
// Nothing this GPU still has to do is needed for the flags to rise (the senders push on
// their own), so the wait cannot deadlock. A sender may legitimately be late (a rank writing
// results, say), so the patience is long: timeoutNs (WENOEXT_HALO_TIMEOUT_S, default 600 s;
// 0 = wait for ever, like a blocking receive) before the launch fails instead of hanging.
// On a timeout the kernel records 1 + (index of the late sender) in *statusDev (device memory) and
// *status (host-mapped memory) and returns: the follow-up cell launch sees it and does nothing, the host turns it into
// WENOEXT_ERR_TIMEOUT. The CUDA context stays usable (no trap).
__global__ void halo_wait_kernel(const int32_t* flags, int n, int32_t epoch, long long timeoutNs,
                                 int32_t* status, int32_t* statusDev) {
    if ((int)threadIdx.x >= n) return;
    long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (ldAcquireSys(flags + threadIdx.x) < epoch) {
        __nanosleep(200);
        long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (timeoutNs > 0 && t1 - t0 > timeoutNs) {
            // device copy: read by the cell kernels that follow (a host-mapped word is far too slow to
            // be read by every thread of a launch); host-mapped copy: read by the host after the call
            if (statusDev) atomicCAS(statusDev, 0, 1 + (int)threadIdx.x);
            if (status) atomicCAS(status, 0, 1 + (int)threadIdx.x);
            return;
        }
    }
}
void launchHaloWait(const int32_t* flags, int n, int32_t epoch, cudaStream_t st, int32_t* status,
                    int32_t* statusDev) {
    static const long long timeoutNs = [] {
        const char* e = getenv("WENOEXT_HALO_TIMEOUT_S");
        return (long long)((e ? atof(e) : 600.0) * 1e9);
    }();
    if (n > 0) halo_wait_kernel<<<1, 32, 0, st>>>(flags, n, epoch, timeoutNs, status, statusDev);
}

void launchPackHalo(const double* vals, int VS, const int32_t* sendIds, int64_t nSend,
                    double* sendBuf, cudaStream_t st) {
    const int64_t n = nSend * VS;
    if (n <= 0) return;
    pack_halo_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(vals, VS, sendIds, nSend,
                                                                  sendBuf);
}

}  // namespace wenoext
