// Batched truncated pseudo-inverse on the GPU (csrc/pinv.cu): the device counterpart of
// svd() + SvdResult::pseudoInverse() (svd.h) for many matrices at once.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace wenoext {

// A_i: rows[i] x n row-major at A + start[i]; Ainv_i: n x rows[i] row-major at Ainv + start[i].
// Returns 0 or a WENOEXT_ERR_* code with `err` set.
int pinvBatchedDevice(int device, int32_t nMat, int32_t n, const int32_t* rows, const int64_t* start,
                      const double* A, double minCondition, double* Ainv, int32_t* nZeros, double* condition,
                      int32_t* converged, std::string& err);

// Host image of wenoext_ls_batch (include/wenoext_b200.h): LS rows, matrices and pseudo-inverses in one
// device batch (csrc/ls_rows.cu).
struct LsBatchHost {
    int32_t polOrder = 0, nDvt = 0, dim[3] = {0, 0, 0};
    std::vector<double> points, cellCentres, cellVolumes;
    std::vector<int64_t> triStart;
    std::vector<int32_t> triPoints;
    std::vector<double> JInv, refPoint, refDet, volI;
    std::vector<int32_t> frameCell, pairFrame, pairCell, rows;
    std::vector<int64_t> start, rowPair;
};
int lsPinvBatchedDevice(int device, const LsBatchHost& b, double minCondition, double* Ainv, int32_t* nZeros,
                        double* condition, int32_t* converged, std::string& err);

}  // namespace wenoext
