/* CPU ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's per-step reconstruction path, used by
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/--impl reference
 * legs as the checker. It is never linked into or called from the product
 * library (wenoext_b200/), which has no CPU path at all.
 *
 * Parity status: the reference library cannot be built here (every hot-path
 * translation unit needs OpenFOAM, which is absent) and its own tests pin no
 * per-face value for correction() (tests/src/WENOUpwindFit-Test.C:309-310 are
 * inequalities only). The restatement is therefore pinned against
 *   (a) oracle/_ref: the same loops with the three matrix products done by the
 *       reference's vendored Blaze 3.8 expression templates, compiled from
 *       /root/reference/blaze-3.8 (see oracle/Makefile), and
 *   (b) property tests the algorithm implies (exact polynomial reproduction,
 *       zero/constant fields, 0..1 bounds for WENOUpwindFit01).
 * "parity unpinned" by reference golden vectors — see DESIGN.md.
 *
 * Each function cites the reference lines it follows (paths relative to the
 * reference root). ORACLE_REAL may be set to `long double` to build the
 * extended-precision variant that bounds both implementations' round-off.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/wenoext_b200.h"

#ifndef ORACLE_REAL
#define ORACLE_REAL double
#endif
typedef ORACLE_REAL real;

#ifndef ORACLE_PREFIX
#define ORACLE_PREFIX weno_oracle
#endif
#define CAT_(a, b) a##_##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(ORACLE_PREFIX, name)

static real absr(real x) { return x < 0 ? -x : x; }

/* WENOCoeff.H:149-157 */
static real intPow(real base, unsigned int exponent) {
    real temp = 1.0;
    for (unsigned int i = 0; i < exponent; ++i) temp *= base;
    return temp;
}

/* offset of processor `proc` in the flat halo array (receive order by proc id) */
static int64_t haloOffset(const wenoext_plan_desc* d, int proc) {
    int64_t off = 0;
    for (int p = 0; p < proc; ++p) off += d->receiveHaloSize[p];
    return off;
}

/* WENOCoeff::getWENOPol + calcCoeff + calcWeight (WENOCoeff.C:299-354, :104-165,
 * :168-221). haloData: receiveHaloData_[proc][id] flattened by proc id.
 * coeffsWeighted: [nCells][nDvt][nComp], zero for Cell::empty cells. */
int FN(getWENOPol_range)(const wenoext_plan_desc* d, int nComp, double p_, double dm_,
                         double epsilon_, const double* vf, const double* haloData,
                         double* coeffsWeighted, int cell0, int cell1) {
    const int K = d->nDvt;
    const unsigned int pExp = (unsigned int)p_;
    int64_t* hoff = (int64_t*)malloc(sizeof(int64_t) * (d->nProcs + 1));
    for (int p = 0; p <= d->nProcs; ++p) hoff[p] = haloOffset(d, p);
    /* scratch sized for the largest stencil */
    int64_t maxN = 1, maxS = 1;
    for (int64_t s = 0; s < d->nStencils; ++s) {
        int64_t n = d->stencilEntryStart[s + 1] - d->stencilEntryStart[s];
        if (n > maxN) maxN = n;
    }
    for (int c = 0; c < d->nCells; ++c) {
        int64_t n = d->cellStencilStart[c + 1] - d->cellStencilStart[c];
        if (n > maxS) maxS = n;
    }
    real* bJ = (real*)malloc(sizeof(real) * maxN * nComp);
    real* coeffs = (real*)malloc(sizeof(real) * maxS * K * nComp);
    real* sumB = (real*)malloc(sizeof(real) * K * nComp);
    real* acc = (real*)malloc(sizeof(real) * K * nComp);
    real* gammaSum = (real*)malloc(sizeof(real) * nComp);
    memset(coeffsWeighted + (size_t)cell0 * K * nComp, 0,
           sizeof(double) * (size_t)(cell1 - cell0) * K * nComp);

    for (int cellI = cell0; cellI < cell1; ++cellI) {
        const int64_t s0 = d->cellStencilStart[cellI], s1 = d->cellStencilStart[cellI + 1];
        /* WENOCoeff.C:333 */
        if (d->stencilsID[d->stencilEntryStart[s0]] == WENOEXT_CELL_EMPTY) continue;
        /* calcCoeff */
        int coeffIndex = 0;
        for (int64_t s = s0; s < s1; ++s) {
            const int64_t e0 = d->stencilEntryStart[s];
            const int64_t size = d->stencilEntryStart[s + 1] - e0;
            if (d->stencilsID[e0] == WENOEXT_CELL_DELETED) continue; /* :121 */
            const int m = d->stencilMatrix[s];
            const double* A = d->matrixData + d->matrixStart[m];
            const int N = d->matrixCols[m];
            for (int64_t j = 1; j < size; ++j) {
                const int32_t id = d->stencilsID[e0 + j];
                const int32_t pm = d->cellToProcMap ? d->cellToProcMap[e0 + j] : WENOEXT_CELL_LOCAL;
                for (int c = 0; c < nComp; ++c) {
                    if (pm == WENOEXT_CELL_LOCAL) /* :141-147 */
                        bJ[(j - 1) * nComp + c] =
                            (real)vf[(size_t)id * nComp + c] - (real)vf[(size_t)cellI * nComp + c];
                    else if (pm != WENOEXT_CELL_DELETED) /* :148-155 */
                        bJ[(j - 1) * nComp + c] = (real)haloData[(size_t)(hoff[pm] + id) * nComp + c] -
                                                  (real)vf[(size_t)cellI * nComp + c];
                }
            }
            /* coeffsList[coeffIndex] = A*bJ_  (:159) */
            real* cf = coeffs + (size_t)coeffIndex * K * nComp;
            for (int k = 0; k < K; ++k)
                for (int c = 0; c < nComp; ++c) {
                    real sum = 0;
                    for (int j = 0; j < N; ++j) sum += (real)A[(size_t)k * N + j] * bJ[j * nComp + c];
                    cf[k * nComp + c] = sum;
                }
            coeffIndex++;
        }
        /* calcWeight */
        const double* B = d->B + (size_t)cellI * K * K;
        for (int i = 0; i < K * nComp; ++i) acc[i] = 0;
        for (int c = 0; c < nComp; ++c) gammaSum[c] = 0;
        for (int stencilI = 0; stencilI < coeffIndex; ++stencilI) {
            const real* cf = coeffs + (size_t)stencilI * K * nComp;
            for (int k = 0; k < K; ++k) /* sumB = B*coeffs (:190) */
                for (int c = 0; c < nComp; ++c) {
                    real sum = 0;
                    for (int q = 0; q < K; ++q) sum += (real)B[k * K + q] * cf[q * nComp + c];
                    sumB[k * nComp + c] = sum;
                }
            for (int c = 0; c < nComp; ++c) {
                real smoothInd = 0; /* trans(colCoeff)*colSumB (:196) */
                for (int k = 0; k < K; ++k) smoothInd += cf[k * nComp + c] * sumB[k * nComp + c];
                real gamma;
                if (stencilI == 0) /* :198-205 */
                    gamma = (real)dm_ / intPow((real)epsilon_ + smoothInd, pExp);
                else
                    gamma = (real)1.0 / intPow((real)epsilon_ + smoothInd, pExp);
                gammaSum[c] += gamma;
                for (int k = 0; k < K; ++k) acc[k * nComp + c] += gamma * cf[k * nComp + c];
            }
        }
        double* out = coeffsWeighted + (size_t)cellI * K * nComp;
        for (int k = 0; k < K; ++k) /* :216-220 */
            for (int c = 0; c < nComp; ++c) out[k * nComp + c] = (double)(acc[k * nComp + c] / gammaSum[c]);
    }
    free(bJ);
    free(coeffs);
    free(sumB);
    free(acc);
    free(gammaSum);
    free(hoff);
    return 0;
}

int FN(getWENOPol)(const wenoext_plan_desc* d, int nComp, double p_, double dm_, double epsilon_,
                   const double* vf, const double* haloData, double* coeffsWeighted) {
    return FN(getWENOPol_range)(d, nComp, p_, dm_, epsilon_, vf, haloData, coeffsWeighted, 0,
                                d->nCells);
}

/* WENOUpwindFit::sumFlux (WENOUpwindFit.C:293-323); cube = intBasTrans()[f][side] */
static void sumFlux(const wenoext_plan_desc* d, int nComp, const double* coeffcI,
                    const double* cube, real* flux) {
    const int s = d->polOrder + 1;
    for (int c = 0; c < nComp; ++c) flux[c] = 0;
    int nCoeff = 0;
    for (int n = 0; n <= d->dim[0]; ++n)
        for (int m = 0; m <= d->dim[1]; ++m)
            for (int l = 0; l <= d->dim[2]; ++l)
                if ((n + m + l) <= d->polOrder && (n + m + l) > 0) {
                    const real I = cube[(size_t)n * s * s + m * s + l];
                    for (int c = 0; c < nComp; ++c) flux[c] += (real)coeffcI[nCoeff * nComp + c] * I;
                    nCoeff++;
                }
}

static const double* cubeOf(const wenoext_plan_desc* d, int64_t face, int side) {
    const int s = d->polOrder + 1;
    return d->intBasTrans + ((size_t)face * 2 + side) * s * s * s;
}

static int isCoupled(const wenoext_patch* p) {
    return p->type == WENOEXT_PATCH_PROCESSOR || p->type == WENOEXT_PATCH_CYCLIC;
}

/* correction() up to and including the first half of coupledRiemannSolver
 * (WENOUpwindFit.C:256-282 and :425-470). corr [nFaces][nComp] is fully
 * written (zero where the reference leaves pTraits::zero). btsfUD
 * [nFaces-nInternalFaces][nComp] is what swapData would send. */
/* internal-face loop of correction() (WENOUpwindFit.C:256-282) over a face range
 * (the threaded CPU baseline splits it; corr must be pre-zeroed) */
int FN(internalFaces_range)(const wenoext_plan_desc* d, int nComp, const double* faceFlux,
                            const double* coeffsWeighted, double* corr, int face0, int face1) {
    const int K = d->nDvt;
    real flux[16];
    if (nComp > 16) return 1;
    for (int faceI = face0; faceI < face1; ++faceI) {
        if (faceFlux[faceI] > 0) {
            const int P = d->owner[faceI];
            sumFlux(d, nComp, coeffsWeighted + (size_t)P * K * nComp, cubeOf(d, faceI, 0), flux);
            for (int c = 0; c < nComp; ++c)
                corr[(size_t)faceI * nComp + c] = (double)(flux[c] / (real)d->refFacAr[faceI]);
        } else if (faceFlux[faceI] < 0) {
            const int N = d->neighbour[faceI];
            sumFlux(d, nComp, coeffsWeighted + (size_t)N * K * nComp, cubeOf(d, faceI, 1), flux);
            for (int c = 0; c < nComp; ++c)
                corr[(size_t)faceI * nComp + c] = (double)(flux[c] / (real)d->refFacAr[faceI]);
        }
    }
    return 0;
}

int FN(phaseA)(const wenoext_plan_desc* d, int nComp, const double* faceFlux,
               const double* coeffsWeighted, double* corr, double* btsfUD) {
    const int K = d->nDvt;
    const int nInt = d->nInternalFaces;
    real flux[16];
    if (nComp > 16) return 1;
    memset(corr, 0, sizeof(double) * (size_t)d->nFaces * nComp);
    memset(btsfUD, 0, sizeof(double) * (size_t)(d->nFaces - nInt) * nComp);
    FN(internalFaces_range)(d, nComp, faceFlux, coeffsWeighted, corr, 0, nInt);
    for (int patchI = 0; patchI < d->nPatches; ++patchI) {
        const wenoext_patch* pt = &d->patches[patchI];
        if (!isCoupled(pt)) continue; /* :433 */
        for (int i = 0; i < pt->size; ++i) {
            const int f = pt->start + i;
            if (faceFlux[f] > 0) {
                const int own = d->owner[f];
                sumFlux(d, nComp, coeffsWeighted + (size_t)own * K * nComp, cubeOf(d, f, 0), flux);
                for (int c = 0; c < nComp; ++c) {
                    double v = (double)(flux[c] / (real)d->refFacAr[f]);
                    btsfUD[(size_t)(f - nInt) * nComp + c] = v;
                    corr[(size_t)f * nComp + c] = v;
                }
            }
        }
    }
    return 0;
}

/* WENOUpwindFit::calcLimiter (WENOUpwindFit.C:38-209) */
/* internal-face part of WENOUpwindFit::calcLimiter (:81-108, :172-179) over a range */
int FN(limiterFitInternal_range)(const wenoext_plan_desc* d, int nComp, const double* vf,
                                 const double* faceFlux, double* corr, int face0, int face1) {
    for (int faceI = face0; faceI < face1; ++faceI) {
        const int P = d->owner[faceI], N = d->neighbour[faceI];
        const int up = faceFlux[faceI] > 0 ? P : N, dn = faceFlux[faceI] > 0 ? N : P;
        for (int c = 0; c < nComp; ++c) {
            real psi = vf[(size_t)up * nComp + c], psiN = vf[(size_t)dn * nComp + c];
            real t = corr[(size_t)faceI * nComp + c];
            real theta = 1.0;
            if (absr(t) > absr(psi - psiN)) {
                theta = absr(psi - psiN) / absr(t);
                if (theta > 1.0) theta = 1.0;
            }
            corr[(size_t)faceI * nComp + c] = (double)(theta * t);
        }
    }
    return 0;
}

static void limiterFit(const wenoext_plan_desc* d, int nComp, const double* vf,
                       const double* faceFlux, const double* patchNbrField, double* corr) {
    const int nInt = d->nInternalFaces;
    FN(limiterFitInternal_range)(d, nComp, vf, faceFlux, corr, 0, nInt);
    for (int patchI = 0; patchI < d->nPatches; ++patchI) {
        const wenoext_patch* pt = &d->patches[patchI];
        /* theta is computed on every coupled patch (:118-165) but only
         * processor patches are scaled (:195-207) */
        if (pt->type != WENOEXT_PATCH_PROCESSOR) continue;
        for (int i = 0; i < pt->size; ++i) {
            const int f = pt->start + i;
            const int own = d->owner[f];
            for (int c = 0; c < nComp; ++c) {
                real a = vf[(size_t)own * nComp + c];
                real b = patchNbrField[(size_t)(f - nInt) * nComp + c];
                real psi = faceFlux[f] > 0 ? a : b, psiN = faceFlux[f] > 0 ? b : a;
                real t = corr[(size_t)f * nComp + c];
                real theta = 1.0;
                if (absr(t) > absr(psi - psiN)) {
                    theta = absr(psi - psiN) / absr(t);
                    if (theta > 1.0) theta = 1.0;
                }
                corr[(size_t)f * nComp + c] = (double)(theta * t);
            }
        }
    }
}

/* WENOUpwindFit01::calcLimiter (WENOUpwindFit01.C:37-212) */
static void limiterFit01(const wenoext_plan_desc* d, int nComp, int bugCompat, const double* vf,
                         const double* faceFlux, const double* coeffsWeighted, double* corr) {
    const int K = d->nDvt;
    const int nInt = d->nInternalFaces;
    const int s = d->polOrder + 1;
    double* theta = (double*)calloc((size_t)d->nCells * nComp, sizeof(double));
    double* zeroCube = (double*)calloc((size_t)s * s * s, sizeof(double));
    real flux[16];
    for (int cellI = 0; cellI < d->nCells; ++cellI) {
        const double* cw = coeffsWeighted + (size_t)cellI * K * nComp;
        for (int cI = 0; cI < nComp; ++cI) {
            const real vfc = vf[(size_t)cellI * nComp + cI];
            real maxP = vfc, minP = vfc;
            for (int fI = d->cellFaceStart[cellI]; fI < d->cellFaceStart[cellI + 1]; ++fI) {
                const int f = d->cellFaces[fI];
                if (f >= nInt) continue;
                real tsfPci = maxP; /* "dummy value so it has the correct type" :85 */
                if (cellI == d->owner[f]) {
                    if (faceFlux[f] > 0) {
                        tsfPci = (real)corr[(size_t)f * nComp + cI] + vfc;
                    } else {
                        sumFlux(d, nComp, cw, cubeOf(d, f, 0), flux);
                        tsfPci = vfc + flux[cI] / (real)d->refFacAr[f];
                    }
                } else if (cellI == d->neighbour[f]) {
                    if (faceFlux[f] < 0) {
                        tsfPci = (real)corr[(size_t)f * nComp + cI] + vfc;
                    } else {
                        /* :125 indexes intBasTrans() with the CELL index */
                        const double* cube;
                        if (bugCompat)
                            cube = cellI < d->nFaces ? cubeOf(d, cellI, 1) : zeroCube;
                        else
                            cube = cubeOf(d, f, 1);
                        sumFlux(d, nComp, cw, cube, flux);
                        tsfPci = vfc + flux[cI] / (real)d->refFacAr[f];
                    }
                }
                if (tsfPci > maxP)
                    maxP = tsfPci;
                else if (tsfPci < minP)
                    minP = tsfPci;
            }
            real argMax, argMin;
            if (absr(maxP - vfc) < 1E-9)
                argMax = 1.0;
            else
                argMax = absr(((real)1.0 - vfc) / (maxP - vfc));
            if (absr(minP - vfc) < 1E-9)
                argMin = 1.0;
            else
                argMin = absr(((real)0.0 - vfc) / (minP - vfc));
            real th = argMax < argMin ? argMax : argMin;
            if (th > 1.0) th = 1.0;
            theta[(size_t)cellI * nComp + cI] = (double)th;
        }
    }
    for (int faceI = 0; faceI < nInt; ++faceI) {
        const int up = faceFlux[faceI] > 0 ? d->owner[faceI] : d->neighbour[faceI];
        for (int c = 0; c < nComp; ++c)
            corr[(size_t)faceI * nComp + c] =
                (double)((real)theta[(size_t)up * nComp + c] * (real)corr[(size_t)faceI * nComp + c]);
    }
    for (int patchI = 0; patchI < d->nPatches; ++patchI) {
        const wenoext_patch* pt = &d->patches[patchI];
        if (pt->type != WENOEXT_PATCH_PROCESSOR) continue;
        for (int i = 0; i < pt->size; ++i) {
            const int f = pt->start + i;
            const int own = d->owner[f];
            for (int c = 0; c < nComp; ++c)
                corr[(size_t)f * nComp + c] =
                    (double)((real)theta[(size_t)own * nComp + c] * (real)corr[(size_t)f * nComp + c]);
        }
    }
    free(theta);
    free(zeroCube);
}

/* Second half of coupledRiemannSolver (WENOUpwindFit.C:474-546) after swapData
 * delivered btsfUDrecv (the neighbour ranks' btsfUD on processor patches), then
 * the limiter of the selected scheme. */
int FN(phaseB)(const wenoext_plan_desc* d, int scheme, int bounded, int nComp, int bugCompat,
               const double* vf, const double* faceFlux, const double* coeffsWeighted,
               const double* patchNbrField, const double* btsfUDrecv, double* corr) {
    const int nInt = d->nInternalFaces;
    if (nComp > 16) return 1;
    for (int patchI = 0; patchI < d->nPatches; ++patchI) {
        const wenoext_patch* pt = &d->patches[patchI];
        if (pt->type == WENOEXT_PATCH_PROCESSOR) {
            for (int i = 0; i < pt->size; ++i) {
                const int f = pt->start + i;
                if (faceFlux[f] < 0)
                    for (int c = 0; c < nComp; ++c)
                        corr[(size_t)f * nComp + c] = btsfUDrecv[(size_t)(f - nInt) * nComp + c];
            }
        } else if (pt->type == WENOEXT_PATCH_CYCLIC) {
            const wenoext_patch* nb = &d->patches[pt->nbr];
            for (int i = 0; i < pt->size; ++i) {
                const int f = pt->start + i;
                if (faceFlux[f] < 0)
                    for (int c = 0; c < nComp; ++c)
                        corr[(size_t)f * nComp + c] = corr[(size_t)(nb->start + i) * nComp + c];
            }
        }
    }
    if (scheme == WENOEXT_SCHEME_UPWINDFIT01)
        limiterFit01(d, nComp, bugCompat, vf, faceFlux, coeffsWeighted, corr);
    else if (bounded)
        limiterFit(d, nComp, vf, faceFlux, patchNbrField, corr);
    return 0;
}

/* Whole correction() for one rank given its halo values and the neighbours'
 * patch values (both NULL in a serial run). */
int FN(correction)(const wenoext_plan_desc* d, int scheme, int bounded, int nComp, double p_,
                   double dm_, double epsilon_, int bugCompat, const double* vf,
                   const double* haloData, const double* faceFlux, const double* patchNbrField,
                   const double* btsfUDrecv, double* corr) {
    double* cw = (double*)malloc(sizeof(double) * (size_t)d->nCells * d->nDvt * nComp);
    double* ud = (double*)malloc(sizeof(double) * (size_t)(d->nFaces - d->nInternalFaces + 1) * nComp);
    int rc = FN(getWENOPol)(d, nComp, p_, dm_, epsilon_, vf, haloData, cw);
    if (!rc) rc = FN(phaseA)(d, nComp, faceFlux, cw, corr, ud);
    if (!rc)
        rc = FN(phaseB)(d, scheme, bounded, nComp, bugCompat, vf, faceFlux, cw, patchNbrField,
                        btsfUDrecv ? btsfUDrecv : ud, corr);
    free(cw);
    free(ud);
    return rc;
}
