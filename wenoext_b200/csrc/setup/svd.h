// One-sided Jacobi SVD and the truncated pseudo-inverse the setup needs.
// Stands in for OpenFOAM's `SVD(A, minCondition)` / `VSinvUt()` used at
// libWENOEXT/WENOBase/WENOBase.C:527,:575 (OpenFOAM is not vendored).
#pragma once
#include <vector>

namespace wenoext {

struct SvdResult {
    int m = 0, n = 0;
    std::vector<double> U;  // m x n row-major
    std::vector<double> S;  // n (zeroed below minCondition * max)
    std::vector<double> V;  // n x n row-major
    int nZeros = 0;
    bool converged = true;
    // V * diag(1/S) * U^T, n x m row-major
    std::vector<double> pseudoInverse() const;
    double condition() const;  // max/min over non-zero S (WENOBase.C:418-434)
};

// A is m x n row-major, m >= n.
SvdResult svd(const std::vector<double>& A, int m, int n, double minCondition);

}  // namespace wenoext
