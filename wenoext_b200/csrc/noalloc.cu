// wenoext_correction_noalloc: correction() for callers that may not allocate device memory (code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in
// the reference.
//
// The fused cell kernel needs two pieces of scratch memory that the ordinary entry points keep in
// cudaMalloc'ed buffers of the plan: the packed value buffer `vals` ([nCells][VS] doubles, written by
// the pack kernel and gathered by the cell kernel) and the pair of tile counters of the persistent
// grid. Here both live in ONE file-scope __device__ array of fixed size, which every GPU that loads the
// library holds until the process exits:
//   [0, 256)                       tile counters (zeroed by every call)
//   [256, 256 + nCells*VS*8)       packed values (fully rewritten by every call)
// Every call on a device shares the array and it keeps the last call's contents: calls that overlap on
// the GPU (two streams, two plans on one device, two host threads) race on it. Serialise them.
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/wenoext_b200.h"
#include "device_plan.h"
#include "plan_internal.h"

namespace {

constexpr size_t kCounterBytes = 256;
// WENOEXT_NOALLOC_MAX_VALUE_BYTES (public header): 8 MiB of packed values = 64^3 cells of a vector
// field (VS = 4) or 2^20 cells of a scalar field
static __device__ __align__(256) unsigned char g_noallocScratch[kCounterBytes + WENOEXT_NOALLOC_MAX_VALUE_BYTES];

// In host code `g_noallocScratch` is only a stand-in for the symbol, not a device address: the
// address is looked up once per device (with that device current) and cached here.
constexpr int kMaxDevices = 64;
unsigned char* g_scratchAddr[kMaxDevices] = {nullptr};
cudaError_t g_scratchErr[kMaxDevices] = {cudaSuccess};
bool g_scratchTried[kMaxDevices] = {false};

// the pack step of the no-alloc path: device code of this file may name the array directly
// (vf AoS [nCells][nComp] -> [nCells][VS], pad lanes zero; same arithmetic as pack_values_kernel)
__global__ void pack_values_scratch_kernel(const double* __restrict__ vf, int nCells, int nComp, int CB,
                                           int VS) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)nCells * VS) return;
    double* vals = reinterpret_cast<double*>(g_noallocScratch + kCounterBytes);
    const int cell = (int)(i / VS), c = (int)(i - (int64_t)cell * VS);
    vals[i] = c < CB ? vf[(size_t)cell * nComp + c] : 0.0;
}

}  // namespace

namespace wenoext {

// Set-up, once per device, with that device current (called by wenoext_plan_create after its
// cudaSetDevice): resolves the device address of the scratch array. Allocates nothing.
void noAllocResolve(int device) {
    if (device < 0 || device >= kMaxDevices || g_scratchTried[device]) return;
    g_scratchTried[device] = true;
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noallocScratch);
    g_scratchErr[device] = e;
    if (e == cudaSuccess)
        g_scratchAddr[device] = static_cast<unsigned char*>(p);
    else
        cudaGetLastError();
}

}  // namespace wenoext

using namespace wenoext;

extern "C" int wenoext_correction_noalloc(wenoext_plan* P, int scheme, int bounded, int nComp,
                                          const double* vf, const double* faceFlux, double* corr,
                                          void* stream) {
    if (!P || !vf || !faceFlux || !corr) {
        setError("correction_noalloc: NULL argument");
        return WENOEXT_ERR_INVALID;
    }
    if (scheme != WENOEXT_SCHEME_UPWINDFIT && scheme != WENOEXT_SCHEME_UPWINDFIT01) {
        setError("correction_noalloc: bad scheme");
        return WENOEXT_ERR_INVALID;
    }
    if (P->nProcs > 1 || P->hasProcessor) {
        setError("correction_noalloc: serial plans only (the halo and patch exchanges use plan-owned buffers)");
        return WENOEXT_ERR_INVALID;
    }
    const int CB = nComp;
    if (nComp < 1 || nComp > cellKernelMaxCB(P->K)) {
        setError("correction_noalloc: nComp must be 1.." + std::to_string(cellKernelMaxCB(P->K)) +
                 " (one component batch)");
        return WENOEXT_ERR_INVALID;
    }
    const int VS = CB == 3 ? 4 : CB;
    const uint64_t need = (uint64_t)P->nCells * VS * 8;
    if (need > (uint64_t)WENOEXT_NOALLOC_MAX_VALUE_BYTES) {
        setError("correction_noalloc: " + std::to_string(P->nCells) + " cells x " + std::to_string(VS) +
                 " doubles = " + std::to_string(need) + " bytes of packed values exceed the static scratch of " +
                 std::to_string((uint64_t)WENOEXT_NOALLOC_MAX_VALUE_BYTES) +
                 " bytes (WENOEXT_NOALLOC_MAX_VALUE_BYTES); use wenoext_correction");
        return WENOEXT_ERR_INVALID;
    }
#ifdef WENOEXT_KDEBUG
    setError("correction_noalloc: not available in WENOEXT_KDEBUG builds (the debug path allocates)");
    return WENOEXT_ERR_INVALID;
#endif
    const int dev = P->device;
    if (dev < 0 || dev >= kMaxDevices || !g_scratchAddr[dev]) {
        setError(std::string("correction_noalloc: the static scratch array could not be resolved on this device: ") +
                 (dev >= 0 && dev < kMaxDevices ? cudaGetErrorString(g_scratchErr[dev]) : "device index out of range"));
        return WENOEXT_ERR_CUDA;
    }
    if (cudaSetDevice(dev) != cudaSuccess) {
        cudaGetLastError();
        setError("correction_noalloc: cudaSetDevice failed");
        return WENOEXT_ERR_CUDA;
    }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* scratch = g_scratchAddr[dev];
    int32_t* counters = reinterpret_cast<int32_t*>(scratch);
    double* vals = reinterpret_cast<double*>(scratch + kCounterBytes);

    // the call record the launch helpers read (device pointers throughout)
    wenoext_call& c = P->call;
    c = wenoext_call();
    c.scheme = scheme;
    c.bounded = bounded ? 1 : 0;
    c.nComp = nComp;
    c.c0 = 0;
    c.CB = CB;
    c.VS = VS;
    c.vf = vf;
    c.flux = faceFlux;
    c.corr = corr;
    c.st = st;
    P->lastLaunches = 0;

    // boundaryField starts as pTraits::zero (WENOUpwindFit.C:230-246)
    if (P->nBnd &&
        cudaMemsetAsync(corr + (size_t)P->nInt * nComp, 0, (size_t)P->nBnd * nComp * 8, st) != cudaSuccess) {
        cudaGetLastError();
        setError("correction_noalloc: cudaMemsetAsync failed");
        return WENOEXT_ERR_CUDA;
    }
    const int64_t n = (int64_t)P->nCells * VS;
    pack_values_scratch_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(vf, P->nCells, nComp, CB, VS);
    P->lastLaunches++;

    // the fused cell kernel with its scratch in the static array: the plan's own buffers are swapped
    // out for the duration of the enqueue (launchCells zeroes the counters it is given and allocates
    // only when the plan has none; per-launch timing would create events, so it is off here)
    double* const savedVals = P->vals;
    int32_t* const savedCounter = P->tileCounter;
    const bool savedTiming = P->timing;
    P->vals = vals;
    P->tileCounter = counters;
    P->timing = false;
    int rc = launchCells(P, nullptr, P->nCells);
    if (!rc && P->hasCoupled) rc = phase2(P);  // cyclic patches: reads and writes `corr` only
    P->vals = savedVals;
    P->tileCounter = savedCounter;
    P->timing = savedTiming;
    return rc;
}
