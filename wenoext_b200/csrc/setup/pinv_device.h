// Batched truncated pseudo-inverse on the GPU (csrc/pinv.cu): the device counterpart of
// svd() + SvdResult::pseudoInverse() (svd.h) for many matrices at once.
#pragma once
#include <cstdint>
#include <string>

namespace wenoext {

// A_i: rows[i] x n row-major at A + start[i]; Ainv_i: n x rows[i] row-major at Ainv + start[i].
// Returns 0 or a WENOEXT_ERR_* code with `err` set.
int pinvBatchedDevice(int device, int32_t nMat, int32_t n, const int32_t* rows, const int64_t* start,
                      const double* A, double minCondition, double* Ainv, int32_t* nZeros, double* condition,
                      int32_t* converged, std::string& err);

}  // namespace wenoext
