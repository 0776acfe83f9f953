#include "svd.h"

#include <algorithm>
#include <cmath>

namespace wenoext {

SvdResult svd(const std::vector<double>& A, int m, int n, double minCondition) {
    SvdResult r;
    r.m = m;
    r.n = n;
    // work on columns: W = A (m x n), V = I
    std::vector<double> W(A);
    std::vector<double> V((size_t)n * n, 0.0);
    for (int i = 0; i < n; ++i) V[(size_t)i * n + i] = 1.0;
    const int maxSweeps = 60;
    bool conv = false;
    for (int sweep = 0; sweep < maxSweeps && !conv; ++sweep) {
        conv = true;
        for (int p = 0; p < n - 1; ++p)
            for (int q = p + 1; q < n; ++q) {
                double alpha = 0, beta = 0, gamma = 0;
                for (int i = 0; i < m; ++i) {
                    double wp = W[(size_t)i * n + p], wq = W[(size_t)i * n + q];
                    alpha += wp * wp;
                    beta += wq * wq;
                    gamma += wp * wq;
                }
                if (gamma == 0.0 || std::fabs(gamma) <= 1e-15 * std::sqrt(alpha * beta)) continue;
                conv = false;
                double zeta = (beta - alpha) / (2.0 * gamma);
                double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                double c = 1.0 / std::sqrt(1.0 + t * t), s = c * t;
                for (int i = 0; i < m; ++i) {
                    double wp = W[(size_t)i * n + p], wq = W[(size_t)i * n + q];
                    W[(size_t)i * n + p] = c * wp - s * wq;
                    W[(size_t)i * n + q] = s * wp + c * wq;
                }
                for (int i = 0; i < n; ++i) {
                    double vp = V[(size_t)i * n + p], vq = V[(size_t)i * n + q];
                    V[(size_t)i * n + p] = c * vp - s * vq;
                    V[(size_t)i * n + q] = s * vp + c * vq;
                }
            }
    }
    r.converged = conv;
    r.S.assign(n, 0.0);
    r.U.assign((size_t)m * n, 0.0);
    for (int j = 0; j < n; ++j) {
        double nrm = 0;
        for (int i = 0; i < m; ++i) nrm += W[(size_t)i * n + j] * W[(size_t)i * n + j];
        nrm = std::sqrt(nrm);
        r.S[j] = nrm;
        if (nrm > 0)
            for (int i = 0; i < m; ++i) r.U[(size_t)i * n + j] = W[(size_t)i * n + j] / nrm;
    }
    r.V = V;
    // OpenFOAM SVD: zero singular values <= minCondition * max(S)
    double maxS = 0;
    for (double s : r.S) maxS = std::max(maxS, s);
    const double minS = minCondition * maxS;
    for (double& s : r.S)
        if (s <= minS) {
            s = 0;
            r.nZeros++;
        }
    return r;
}

std::vector<double> SvdResult::pseudoInverse() const {
    std::vector<double> P((size_t)n * m, 0.0);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < m; ++j) {
            double s = 0;
            for (int k = 0; k < n; ++k)
                if (S[k] != 0) s += V[(size_t)i * n + k] * (1.0 / S[k]) * U[(size_t)j * n + k];
            P[(size_t)i * m + j] = s;
        }
    return P;
}

double SvdResult::condition() const {
    double mn = 1e15, mx = -1e15;
    for (double s : S) {
        if (s == 0) continue;
        mn = std::min(mn, s);
        mx = std::max(mx, s);
    }
    return mx / mn;
}

}  // namespace wenoext
