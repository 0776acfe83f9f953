/* CPU ORACLE (Blaze-backed variant) — TEST INFRASTRUCTURE ONLY.
 *
 * Same cell loop as weno_oracle.c, but the three matrix products of the hot path
 * are evaluated by the reference's own arithmetic library — the Blaze 3.8 headers
 * vendored in the reference tree (compiled from /root/reference/blaze-3.8 by
 * oracle/Makefile; nothing is copied) — with the container types the reference
 * uses, so that padding, SIMD partial sums and FMA contraction match what
 * libWENOEXT does on a host of the same SIMD width:
 *   A, B    blaze::DynamicMatrix<double>              (matrixDB.H:69, geometryWENO.H)
 *   bJ_, coeffs  blaze::DynamicMatrix<double, columnMajor>   (WENOCoeff.H:73, :109)
 *   coeffs = A*bJ_ (WENOCoeff.C:159); sumB = B*coeffs (:190);
 *   smoothInd = trans(column(coeffs,c))*column(sumB,c) (:194-196)
 * The matrices live in a de-duplicated pool exactly like the reference's matrixDB
 * (what the reference really runs: cache-resident on uniform meshes).
 * The rest of libWENOEXT cannot be compiled here (needs OpenFOAM), see DESIGN.md.
 */
#include <blaze/Math.h>

#include <cstdint>
#include <cstring>
#include <vector>

#include "../include/wenoext_b200.h"

namespace {
using RowMat = blaze::DynamicMatrix<double>;
using ColMat = blaze::DynamicMatrix<double, blaze::columnMajor>;

struct BlazeOracle {
    const wenoext_plan_desc* d;
    std::vector<RowMat> pool;  // LSmatrix DB
    std::vector<RowMat> B;     // B()[cell]
    std::vector<int64_t> hoff;
};

inline double intPow(double base, unsigned int exponent) {
    double temp = 1.0;
    for (unsigned int i = 0; i < exponent; ++i) temp *= base;
    return temp;
}
}  // namespace

extern "C" void* weno_oracle_blaze_create(const wenoext_plan_desc* d) {
    BlazeOracle* o = new BlazeOracle;
    o->d = d;
    const int K = d->nDvt;
    o->pool.resize(d->nUniqueMatrices);
    for (int m = 0; m < d->nUniqueMatrices; ++m) {
        const int N = d->matrixCols[m];
        RowMat M(K, N);
        const double* src = d->matrixData + d->matrixStart[m];
        for (int i = 0; i < K; ++i)
            for (int j = 0; j < N; ++j) M(i, j) = src[(size_t)i * N + j];
        o->pool[m] = std::move(M);
    }
    o->B.resize(d->nCells);
    for (int c = 0; c < d->nCells; ++c) {
        RowMat M(K, K);
        const double* src = d->B + (size_t)c * K * K;
        for (int i = 0; i < K; ++i)
            for (int j = 0; j < K; ++j) M(i, j) = src[i * K + j];
        o->B[c] = std::move(M);
    }
    o->hoff.assign(d->nProcs + 1, 0);
    if (d->nProcs > 1)
        for (int p = 0; p < d->nProcs; ++p) o->hoff[p + 1] = o->hoff[p] + d->receiveHaloSize[p];
    return o;
}

extern "C" void weno_oracle_blaze_destroy(void* h) { delete static_cast<BlazeOracle*>(h); }

/* WENOCoeff::getWENOPol for cells [cell0, cell1) (the reference loops over all
 * cells; the range lets the harness emulate one MPI rank per core). */
extern "C" int weno_oracle_blaze_getWENOPol_range(void* h, int nComp, double p_, double dm_,
                                                  double epsilon_, const double* vf,
                                                  const double* haloData, double* coeffsWeighted,
                                                  int cell0, int cell1) {
    BlazeOracle* o = static_cast<BlazeOracle*>(h);
    const wenoext_plan_desc* d = o->d;
    const int K = d->nDvt;
    const unsigned int pExp = (unsigned int)p_;
    ColMat bJ;
    std::vector<ColMat> coeffsList(10);
    std::vector<double> gammaSum(nComp), gamma(nComp);
    for (int cellI = cell0; cellI < cell1; ++cellI) {
        double* out = coeffsWeighted + (size_t)cellI * K * nComp;
        std::memset(out, 0, sizeof(double) * K * nComp);
        const int64_t s0 = d->cellStencilStart[cellI], s1 = d->cellStencilStart[cellI + 1];
        if (d->stencilsID[d->stencilEntryStart[s0]] == WENOEXT_CELL_EMPTY) continue;
        if ((int64_t)coeffsList.size() < s1 - s0) coeffsList.resize(s1 - s0);
        int coeffIndex = 0;
        for (int64_t s = s0; s < s1; ++s) {
            const int64_t e0 = d->stencilEntryStart[s];
            const int64_t size = d->stencilEntryStart[s + 1] - e0;
            if (d->stencilsID[e0] == WENOEXT_CELL_DELETED) continue;
            const RowMat& A = o->pool[d->stencilMatrix[s]];
            bJ.resize(A.columns(), nComp);
            for (int64_t j = 1; j < size; ++j) {
                const int32_t id = d->stencilsID[e0 + j];
                const int32_t pm = d->cellToProcMap ? d->cellToProcMap[e0 + j] : WENOEXT_CELL_LOCAL;
                if (pm == WENOEXT_CELL_LOCAL) {
                    for (int c = 0; c < nComp; ++c)
                        bJ(j - 1, c) = vf[(size_t)id * nComp + c] - vf[(size_t)cellI * nComp + c];
                } else if (pm != WENOEXT_CELL_DELETED) {
                    for (int c = 0; c < nComp; ++c)
                        bJ(j - 1, c) = haloData[(size_t)(o->hoff[pm] + id) * nComp + c] -
                                       vf[(size_t)cellI * nComp + c];
                }
            }
            coeffsList[coeffIndex] = A * bJ;
            coeffIndex++;
        }
        const RowMat& B = o->B[cellI];
        for (int c = 0; c < nComp; ++c) gammaSum[c] = gamma[c] = 0.0;
        for (int stencilI = 0; stencilI < coeffIndex; ++stencilI) {
            const ColMat& coeffs = coeffsList[stencilI];
            auto sumB = B * coeffs;
            for (int compI = 0; compI < nComp; ++compI) {
                const auto colCoeff = column(coeffs, compI);
                const auto colSumB = column(sumB, compI);
                const double smoothInd = trans(colCoeff) * colSumB;
                if (stencilI == 0)
                    gamma[compI] = dm_ / (intPow(epsilon_ + smoothInd, pExp));
                else
                    gamma[compI] = 1.0 / (intPow(epsilon_ + smoothInd, pExp));
                gammaSum[compI] += gamma[compI];
                for (int k = 0; k < K; ++k) out[k * nComp + compI] += gamma[compI] * colCoeff[k];
            }
        }
        for (int k = 0; k < K; ++k)
            for (int c = 0; c < nComp; ++c) out[k * nComp + c] /= gammaSum[c];
    }
    return 0;
}

extern "C" int weno_oracle_blaze_getWENOPol(void* h, int nComp, double p_, double dm_,
                                            double epsilon_, const double* vf,
                                            const double* haloData, double* coeffsWeighted) {
    BlazeOracle* o = static_cast<BlazeOracle*>(h);
    return weno_oracle_blaze_getWENOPol_range(h, nComp, p_, dm_, epsilon_, vf, haloData,
                                              coeffsWeighted, 0, o->d->nCells);
}

/* Memory image of a row-major blaze::DynamicMatrix<double> holding A (rows x cols): what
 * BlazeIO.C:126-129 writes to the cache (M.data(), M.spacing()*M.rows() doubles). Returns
 * the spacing; `out` receives the image if it has room for `cap` doubles. Pins the padded
 * blob format of tests/test_wenobase_cache.py against the real container. */
extern "C" int weno_oracle_blaze_image(int rows, int cols, const double* A, double* out, int cap) {
    blaze::DynamicMatrix<double> M(rows, cols);
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) M(i, j) = A[i * cols + j];
    const int spacing = (int)M.spacing();
    if (out && cap >= spacing * rows) memcpy(out, M.data(), sizeof(double) * spacing * rows);
    return spacing;
}
