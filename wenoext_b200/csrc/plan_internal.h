// Internal definition of the opaque wenoext_plan (one plan <-> one GPU <-> one
// host thread at a time; SURVEY.md 8b "Threading").
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/wenoext_b200.h"
#include "device_plan.h"

namespace wenoext {
void setError(const std::string& s);
extern thread_local std::string g_lastError;

// NCCL entry points resolved with dlopen (no link-time dependency, so the
// library loads on machines without NCCL or without a GPU).
struct NcclId {
    char internal[128];
};
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, NcclId /*ncclUniqueId by value*/, int) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
NcclApi* ncclApi();  // nullptr (and last error set) if libnccl cannot be loaded
}  // namespace wenoext

struct wenoext_call {
    int scheme = 0, bounded = 0, nComp = 0, c0 = 0, CB = 0, VS = 0;
    const double *vf = nullptr, *flux = nullptr, *pnf = nullptr;  // device pointers
    double *corr = nullptr, *fv = nullptr;
    cudaStream_t st = nullptr;
};

struct wenoext_plan {
    int device = 0;
    int nSM = 0;
    int smemLimit = 0;
    int32_t polOrder = 0, K = 0, nCells = 0, nFaces = 0, nInt = 0, nBnd = 0;
    int32_t nProcs = 1, myProc = 0;
    // records
    unsigned char* records = nullptr;
    wenoext::RecInfo* recInfo = nullptr;
    int32_t maxN = 0;
    int32_t* dbg = nullptr;  // WENOEXT_KDEBUG builds only
    int32_t* tileCounter = nullptr;  // work counter of the cell kernel, zeroed before every launch
    int64_t recordBytesTotal = 0, maxRecordBytes = 0, maxHeadBytes = 0, maxTailBytes = 0;
    int32_t maxSt = 0, maxFc = 0;
    // 2-D bulk copies: uniform record stride (0: records packed, no fast path), the layout the
    // tensor maps were built for, the maps (device) and the part table of that layout
    int64_t recStride = 0;
    uint32_t tmaZ = 0;
    int32_t tmaW = 0;
    void* tmaMaps = nullptr;
    uint32_t tmaBytes[18] = {0}, tmaElemOff[18] = {0};
    int32_t* cellListInterior = nullptr;
    int32_t* cellListBoundary = nullptr;
    int32_t nInterior = 0, nBoundaryCells = 0;
    // One process per GPU: exchange #1 as a one-sided push into the receivers' value buffers
    // (CUDA IPC over NVLink, halo_push_kernel) instead of NCCL send/recv, so that no
    // communication kernel competes with the persistent cell kernel for SMs. Interior and
    // boundary cells then go in one launch, the boundary tiles behind the halo gate.
    bool p2p = false;
    double* valsIpc[2] = {nullptr, nullptr};  // double-buffered value buffers the peers write into
    int32_t* haloFlags = nullptr;             // [nRecvPeers]: epoch of the last halo each sender delivered
    int32_t nRecvPeers = 0;
    int32_t* pushDone = nullptr;              // block counters of halo_push_kernel: [0] exchange #1, [1] exchange #2
    bool patchPush = false;                   // exchange #2 over peer memory too (else grouped ncclSend/ncclRecv)
    int32_t nPatchSenders = 0;                // flags [32, 32 + nPatchSenders) of haloFlags
    static constexpr int kPatchMaxComp = 9;   // the IPC receive buffer holds [nBnd][9] doubles (tensor)
    struct PushPeer {
        int32_t proc = 0;
        double* vals[2] = {nullptr, nullptr};  // peer's valsIpc, mapped here
        int32_t* flags = nullptr;              // peer's haloFlags, mapped here
        int64_t haloBase = 0;                  // peer's nCells + its haloOff[this rank]
        int32_t slot = 0;                      // this rank's index among the peer's senders
        // exchange #2 (processor-patch face values) pushed the same way: the peer's receive buffer,
        // where this rank's patch starts in it (boundary-face index), this rank's patch, flag slot
        double* recvPatch = nullptr;
        int64_t patchOff = -1;
        int32_t patchStart = 0, patchSize = 0, patchSlot = -1;
    };
    std::vector<PushPeer> pushPeers;
    int32_t haloEpoch = 0;
    // halo_wait_kernel's verdict, host-mapped: 0, or 1 + index of the sender that never delivered
    int32_t* haloStatusH = nullptr;
    int32_t* haloStatusD = nullptr;
    int32_t* haloStatusDev = nullptr;  // the same verdict in device memory, read by the cell kernels
    bool mergedLaunch = false;  // this call's boundary cells already went with the interior launch
    int cfgMerge = 1;           // WENOEXT_MERGE=0: separate interior / boundary launches
    int cfgP2P = 1;             // WENOEXT_P2P=0: NCCL send/recv for exchange #1
    int cfgPatchPush = 1;       // WENOEXT_PATCH_PUSH=0: grouped ncclSend/ncclRecv for exchange #2
    // faces / patches
    int32_t* owner = nullptr;
    int32_t* bfPatchType = nullptr;
    int32_t* bfPartner = nullptr;
    int32_t* bfNbrCell = nullptr;
    bool hasCoupled = false, hasProcessor = false, hasNbrCell = false;
    std::vector<wenoext_patch> patches;
    // halo exchange #1
    std::vector<int64_t> haloOff, sendOff;  // [nProcs+1], in cells
    std::vector<int32_t> sendProcList, receiveProcList;
    int32_t* sendIds = nullptr;
    int64_t nSendTotal = 0, nHaloTotal = 0;
    // work buffers
    double* vals = nullptr;
    size_t valsCap = 0;
    double* sendBuf = nullptr;
    size_t sendCap = 0;
    double* recvPatch = nullptr;
    double* thetaPatch = nullptr;
    double* sendPatch = nullptr;  // in-process peers: snapshot of the patch values to send
    size_t patchCap = 0;
    // wenoext_div: faces of a cell in ascending face index, device scratch
    int32_t* cellFaceStartD = nullptr;
    int32_t* cellFacesSortedD = nullptr;
    int32_t* neighbourD = nullptr;  // mesh.neighbour() [nInternalFaces]: upwind cell of a face in wenoext_div
    // wenoext_div with host buffers on a serial plan without coupled patches: the cells whose faces are final
    // after slab k (doneCells[k], a prefix) are integrated and copied out under the kernels of the slabs that
    // follow, instead of one launch and 8*nCells*nComp bytes of copy behind the last slab
    std::vector<int32_t> doneCells;
    bool divPipeOn = false, divPipeDone = false;
    wenoext::DivKernelArgs divPipeArgs{};
    double* divPipeHost = nullptr;
    std::vector<cudaEvent_t> evDiv;
    cudaStream_t sDiv = nullptr;
    bool skipD2H = false;           // host-buffer reconstruction whose face field stays on the device (wenoext_div)
    bool volValid = false;          // d_vol holds the cell volumes of an earlier wenoext_div call
    double *d_corr = nullptr, *d_bnd = nullptr, *d_vol = nullptr, *d_div = nullptr;
    size_t d_corrCap = 0, d_bndCap = 0, d_volCap = 0, d_divCap = 0;
    // wenoext_set_face_flux: the flux field of this time step (device); stepFluxOwned if uploaded here
    const double* stepFlux = nullptr;
    double* stepFluxOwned = nullptr;
    size_t stepFluxCap = 0;
    // staging for host-pointer calls
    double *s_vf = nullptr, *s_flux = nullptr, *s_pnf = nullptr, *s_corr = nullptr,
           *s_fv = nullptr;
    size_t s_vfCap = 0, s_fluxCap = 0, s_pnfCap = 0, s_corrCap = 0, s_fvCap = 0;
    // pipelined host-buffer path (serial plans): slab k = cells [slabEnd[k-1], slabEnd[k])
    // needs vf[0, needVf[k]) and flux[0, needFlux[k]); afterwards the internal faces
    // [0, doneFaces[k]) are final
    std::vector<int32_t> slabEnd, needVf, needFlux, doneFaces;
    // parallel plans: slabs run over the interior cell list; internal faces of boundary
    // cells are finalised last and travel compactly (fixFaces), scattered on the host
    std::vector<int32_t> fixFacesH;
    int32_t* fixFacesD = nullptr;
    double* d_fix = nullptr;
    size_t d_fixCap = 0;
    double* h_fix = nullptr;  // pinned
    size_t h_fixCap = 0;
    cudaStream_t sIn = nullptr, sOut = nullptr;
    cudaStream_t sAlt = nullptr;  // every second slab's kernel (overlaps the drain of the one before)
    cudaStream_t sPack = nullptr;  // packs of the slabs' cell values (high priority), so that the input copies run ahead
    std::vector<cudaEvent_t> evCopy;  // [slab] its input copies are done
    std::vector<cudaEvent_t> evSlab;  // [2 * slabs]: inputs arrived, slab computed
    // WENODict factors
    double p = 4.0, dm = 1000.0, eps = 1e-40;
    int bugCompat = 1;
    // in-process peers (single-process multi-rank runs, e.g. tests on one GPU)
    std::vector<wenoext_plan*> localPeers;
    wenoext_call call;
    // NCCL
    void* comm = nullptr;
    cudaStream_t commStream = nullptr;
    cudaEvent_t evA = nullptr, evB = nullptr, evC = nullptr;
    // statistics
    int lastLaunches = 0;
    bool timing = false;
    std::vector<cudaEvent_t> timeEvents;  // (start, stop) pairs of timed cell-kernel launches
    size_t timeUsed = 0;
    wenoext_plan_info info{};
    // launch geometry overrides (tuning): 0 = auto
    int cfgWarps = 0, cfgSlots = 0, cfgGrid = 0, cfgSmemKB = 0, cfgCells = 0;
};

namespace wenoext {
// plan.cu: enqueue the fused cell kernel over work items [workBase, workBase + nWork) of `list`
// (nullptr: cell = work item) with the call record P->call; coupled-patch epilogue
int launchCells(wenoext_plan* P, const int32_t* list, int nWork, int workBase = 0, int gateItem = -1,
                bool followUp = false, cudaStream_t st = nullptr, int counterSlot = 0);
int phase2(wenoext_plan* P);
// noalloc.cu: resolve the static scratch array's device address for `device` (current device)
void noAllocResolve(int device);
}  // namespace wenoext
