/* wenoext_b200 — C ABI of the B200-native WENOEXT reconstruction hot path.
 *
 * Drop-in boundary for the per-timestep work of
 *   WENOUpwindFit<Type>::correction()    libWENOEXT/WENOUpwindFit/WENOUpwindFit.C:212-290
 *   WENOUpwindFit01<Type>::correction()  libWENOEXT/WENOUpwindFit01/WENOUpwindFit01.C:215-292
 * and everything they call per step (WENOCoeff::getWENOPol/calcCoeff/calcWeight/
 * collectData, sumFlux, coupledRiemannSolver, swapData, calcLimiter).
 * All citations are relative to the reference tree (WENO-OF/WENOEXT @ 78907ff).
 *
 * Plain pointers and sizes only. Every function returns 0 on success and a
 * non-zero code otherwise; wenoext_last_error() gives the message. There is no
 * CPU fallback: calls that need the GPU fail when none is present.
 */
#ifndef WENOEXT_B200_H
#define WENOEXT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WENOEXT_OK 0
#define WENOEXT_ERR_INVALID 1   /* bad argument / inconsistent plan */
#define WENOEXT_ERR_CUDA 2      /* CUDA runtime error or no device */
#define WENOEXT_ERR_NCCL 3      /* NCCL unavailable or failed */
#define WENOEXT_ERR_SETUP 4     /* host setup (WENOBase stand-in) failed */
#define WENOEXT_ERR_TIMEOUT 5   /* a peer's halo did not arrive within WENOEXT_HALO_TIMEOUT_S */

/* WENOBase::Cell (WENOBase.H:62-67) */
#define WENOEXT_CELL_LOCAL (-1)
#define WENOEXT_CELL_DELETED (-4)
#define WENOEXT_CELL_EMPTY (-10)

/* fvPatch kinds the path distinguishes (WENOUpwindFit.C:433, :478, :493) */
#define WENOEXT_PATCH_GENERIC 0   /* not coupled: zero correction */
#define WENOEXT_PATCH_PROCESSOR 1 /* processorFvPatch, nbr = neighbour rank */
#define WENOEXT_PATCH_CYCLIC 2    /* cyclicFvPatch, nbr = neighbour patch index */
#define WENOEXT_PATCH_EMPTY 3     /* emptyFvPatch */

/* scheme selector: TypeName("WENOUpwindFit") WENOUpwindFit.H:156,
 *                  TypeName("WENOUpwindFit01") WENOUpwindFit01.H:149 */
#define WENOEXT_SCHEME_UPWINDFIT 0
#define WENOEXT_SCHEME_UPWINDFIT01 1

#define WENOEXT_MEM_HOST 0
#define WENOEXT_MEM_DEVICE 1

typedef struct wenoext_patch {
    int32_t type;  /* WENOEXT_PATCH_* */
    int32_t start; /* patch.start(): first face index */
    int32_t size;
    int32_t nbr;   /* neighbour rank (processor) / neighbour patch (cyclic) / -1 */
} wenoext_patch;

/* The immutable per-(mesh, order) data WENOBase owns, flattened to CSR.
 * Each member names the WENOBase accessor it replaces (WENOBase.H:319-397).
 * All pointers are HOST pointers; wenoext_plan_create copies what it needs. */
typedef struct wenoext_plan_desc {
    int32_t polOrder;       /* WENOUpwindFit.H:207 */
    int32_t nDvt;           /* degreesOfFreedom(), WENOBase.C:1130-1140 */
    int32_t dim[3];         /* dimList()[cell] (same for all cells, WENOBase.C:1147-1154) */
    int32_t nCells, nFaces, nInternalFaces;
    const int32_t* owner;         /* mesh.owner()      [nFaces] */
    const int32_t* neighbour;     /* mesh.neighbour()  [nInternalFaces] */
    const int32_t* cellFaceStart; /* mesh.cells() CSR  [nCells+1] */
    const int32_t* cellFaces;
    int32_t nPatches;
    const wenoext_patch* patches; /* mesh.boundary() */

    /* stencilsID()[cell][stencil][j] and cellToProcMap()[cell][stencil][j]:
     * cellStencilStart[cell] .. cellStencilStart[cell+1] are the cell's stencil
     * slots IN LIST ORDER (deleted ones included: one entry, value -4; an empty
     * cell has -10 in entry 0 of stencil 0, WENOCoeff.C:333).
     * Entry 0 of a live stencil is the cell itself (WENOBase.C:976-979). */
    int64_t nStencils;
    const int64_t* cellStencilStart;  /* [nCells+1] */
    const int64_t* stencilEntryStart; /* [nStencils+1] */
    const int32_t* stencilsID;
    const int32_t* cellToProcMap; /* may be NULL: every entry local */

    /* LSmatrix()[cell][stencil](): nDvt x (stencil size-1), row-major, UNPADDED.
     * stencilMatrix[s] indexes the de-duplicated pool (matrixDB.H:136-139);
     * -1 for deleted stencils. */
    const int32_t* stencilMatrix; /* [nStencils] */
    int32_t nUniqueMatrices;
    const int64_t* matrixStart; /* [nUniqueMatrices+1] offsets into matrixData */
    const int32_t* matrixCols;  /* [nUniqueMatrices] */
    const double* matrixData;

    const double* B;           /* B()[cell], row-major nDvt x nDvt   [nCells] */
    const double* intBasTrans; /* intBasTrans()[face][0|1](n,m,l), cubes of
                                  (polOrder+1)^3, List3D.H:238        [nFaces][2] */
    const double* refFacAr;    /* refFacAr()[face]                    [nFaces] */

    /* halo lists, indexed by processor id, length nProcs (WENOBase.H:122-138) */
    int32_t nProcs, myProc;
    const int32_t* sendProcList;    /* -1: nothing to send */
    const int32_t* receiveProcList; /* -1: nothing to receive */
    const int32_t* receiveHaloSize;
    const int64_t* sendHaloStart;   /* CSR over sendHaloCellIDList() [nProcs+1] */
    const int32_t* sendHaloCellID;

    /* optional: for each boundary face (index face - nInternalFaces) of a
     * coupled patch, where the cell on the other side lives: a local cell id
     * (cyclic) or nCells + haloOffset(proc) + haloId (processor); -1 otherwise.
     * Lets the bounded limiter form patchNeighbourField() (WENOUpwindFit.C:130)
     * from the halo buffer. NULL: pass patchNbrField to wenoext_correction. */
    const int32_t* patchNbrCell;
} wenoext_plan_desc;

/* The same descriptor for a WM_LABEL_SIZE=64 build of OpenFOAM (labels are int64 there,
 * libWENOEXT/CMakeLists.txt:216; the on-disk cache carries the width, WENOBase.C:1265-1312): every array the
 * 32-bit descriptor holds as OpenFOAM `label`s is int64 here, so the shim can pass its lists as they are.
 * Members have the meaning of their namesakes in wenoext_plan_desc. wenoext_plan_create_labels64 checks that
 * every value fits 32 bits (180 GB of HBM hold about 14 M order-2 cells; ids beyond 2^31 cannot occur in a
 * plan that fits one GPU), narrows into temporaries and creates the same plan. */
typedef struct wenoext_patch64 {
    int64_t type, start, size, nbr;
} wenoext_patch64;
typedef struct wenoext_plan_desc64 {
    int64_t polOrder, nDvt, dim[3];
    int64_t nCells, nFaces, nInternalFaces;
    const int64_t* owner;
    const int64_t* neighbour;
    const int64_t* cellFaceStart;
    const int64_t* cellFaces;
    int64_t nPatches;
    const wenoext_patch64* patches;
    int64_t nStencils;
    const int64_t* cellStencilStart;
    const int64_t* stencilEntryStart;
    const int64_t* stencilsID;
    const int64_t* cellToProcMap;
    const int64_t* stencilMatrix;
    int64_t nUniqueMatrices;
    const int64_t* matrixStart;
    const int64_t* matrixCols;
    const double* matrixData;
    const double* B;
    const double* intBasTrans;
    const double* refFacAr;
    int64_t nProcs, myProc;
    const int64_t* sendProcList;
    const int64_t* receiveProcList;
    const int64_t* receiveHaloSize;
    const int64_t* sendHaloStart;
    const int64_t* sendHaloCellID;
    const int64_t* patchNbrCell;
} wenoext_plan_desc64;

typedef struct wenoext_plan wenoext_plan;

typedef struct wenoext_plan_info {
    int64_t nCells, nFaces, nInternalFaces, nCoupledFaces;
    int64_t nLiveStencils, nHaloCells;
    int64_t nInteriorCells;     /* cells with no halo entry in any stencil */
    int64_t recordBytes;        /* device bytes of the per-cell records */
    int64_t maxRecordBytes;
    /* SURVEY.md 8(d) algorithmic bytes of one correction() call, split so the
     * caller can add the per-component terms:
     *   bytes(C) = algBytesFixed + C * algBytesPerComp                       */
    int64_t algBytesFixed;
    int64_t algBytesPerComp;
    int64_t algFlopsPerComp;
} wenoext_plan_info;

const char* wenoext_last_error(void);
/* number of CUDA devices visible; 0 without a GPU (never fails) */
int wenoext_device_count(void);

/* Flatten + upload (once). Replaces the WENOBase::instance() lookup done in
 * WENOCoeff's constructor (WENOCoeff.C:53-99, WENOBase.H:281-294). */
int wenoext_plan_create(const wenoext_plan_desc* desc, int device, wenoext_plan** out);
/* wenoext_plan_create for 64-bit labels (wenoext_plan_desc64 above); WENOEXT_ERR_INVALID with the name of
 * the array if a value does not fit 32 bits. */
int wenoext_plan_create_labels64(const wenoext_plan_desc64* desc, int device, wenoext_plan** out);
int wenoext_plan_destroy(wenoext_plan* plan);
int wenoext_plan_get_info(const wenoext_plan* plan, wenoext_plan_info* info);

/* WENODict run-time factors p, dm, epsilon (WENOCoeff.C:87-89; defaults 4, 1000,
 * 1e-40). p is truncated to unsigned as WENOCoeff.H:149 does. */
int wenoext_set_params(wenoext_plan* plan, double p, double dm, double epsilon);
/* 1 (default): reproduce WENOUpwindFit01.C:125 (intBasTrans()[cellI][1]); 0: use
 * the face index. */
int wenoext_set_fit01_bug_compat(wenoext_plan* plan, int on);

/* NCCL plumbing for the two exchanges (collectData, WENOCoeff.C:224-296 and
 * swapData, WENOUpwindFit.C:326-385). libnccl is dlopen'ed on first use. */
int wenoext_nccl_unique_id(char id[128]);
int wenoext_attach_nccl(wenoext_plan* plan, const char id[128], int rank, int nRanks);

/* Page-locked host memory for the buffers handed to the *_HOST calls (the library overlaps the
 * copies of a host-buffer call with its kernels slab by slab, which needs pinned memory; pageable
 * buffers work but serialise). What the OpenFOAM shim stages faceFlux / the result through. */
int wenoext_host_alloc(uint64_t bytes, void** out);
int wenoext_host_free(void* p);

/* faceFlux_ is a member of the scheme object and the same field for every correction() of a time
 * step (U, then each transported scalar: WENOUpwindFit.H:75, :195-211). Upload it once: later calls
 * that pass faceFlux == NULL use this copy (MEM_HOST: copied to the device here, the host buffer is
 * free again on return; MEM_DEVICE: the pointer is kept and must stay valid). Halves the host-to-device
 * bytes of a host-buffer call. faceFlux == NULL here forgets the field. */
int wenoext_set_face_flux(wenoext_plan* plan, const double* faceFlux, int memKind, void* stream);

/* correction(vf): one call = one reference correction() (halo exchange, fused
 * coefficient/weight/face kernel, coupled patches, limiter).
 *   vf            [nCells][nComp]            vf.internalField(), AoS
 *   faceFlux      [nFaces]                   faceFlux_ (internal faces, then patches); NULL: the
 *                 field given to wenoext_set_face_flux
 *   patchNbrField [nFaces-nInternalFaces][nComp] patchNeighbourField() values on
 *                 coupled patch faces; may be NULL if desc.patchNbrCell was given
 *                 or bounded == 0
 *   corr          [nFaces][nComp] out: tsfP incl. boundaryField (zero on
 *                 non-coupled patches)
 * memKind: WENOEXT_MEM_HOST (pageable or pinned; staged inside) or _DEVICE.
 * stream: cudaStream_t (NULL = default stream). Host buffers: returns after the
 * result is in `corr`. Device buffers: stream-ordered, returns immediately. */
int wenoext_correction(wenoext_plan* plan, int scheme, int bounded, int nComp,
                       const double* vf, const double* faceFlux, const double* patchNbrField,
                       double* corr, int memKind, void* stream);

/* Same, plus the interpolated face value psi_upwind + correction on internal
 * faces (what surfaceInterpolationScheme::interpolate forms from weights() and
 * correction()); faceValue [nInternalFaces][nComp], device or host like corr. */
int wenoext_interpolate(wenoext_plan* plan, int scheme, int bounded, int nComp,
                        const double* vf, const double* faceFlux, const double* patchNbrField,
                        double* corr, double* faceValue, int memKind, void* stream);

/* fvc::div(faceFlux, vf) with the WENO scheme, formed entirely on the device: what
 * gaussConvectionScheme<Type>::fvcDiv builds from this scheme (OpenFOAM side of
 * WENOUpwindFit.H:223-242), fvc::surfaceIntegrate(faceFlux * interpolate(vf)):
 *   face value = weights()*vf_P + (1-weights())*vf_N + correction()   (internal and coupled faces)
 *              = bndValue                                            (other patches)
 *   div[cell]  = (sum over the cell's faces, ascending face index as OpenFOAM's face loops:
 *                 +flux*value where the cell owns the face, -flux*value where it is the
 *                 neighbour) / cellVolumes[cell]
 * Only [nCells][nComp] returns to the host instead of [nFaces][nComp]; with host buffers
 * the reconstruction runs slab-pipelined like wenoext_correction (copies of the next slab under the kernel
 * of the current one), on serial and on parallel plans; on serial plans without coupled patches the divergence
 * itself is formed and copied out slab by slab under the kernels. faceFlux may be NULL after wenoext_set_face_flux.
 *   bndValue    [nFaces-nInternalFaces][nComp]  vf.boundaryField() on the patch faces (read on
 *               non-coupled patches only)
 *   cellVolumes [nCells]                        mesh.V(); host-buffer calls: NULL reuses the volumes the
 *               previous host-buffer call uploaded (the mesh does not change between time steps)
 *   div         [nCells][nComp] out
 * Coupled patches need the value across the face: patchNbrField, or desc.patchNbrCell. */
int wenoext_div(wenoext_plan* plan, int scheme, int bounded, int nComp, const double* vf,
                const double* faceFlux, const double* patchNbrField, const double* bndValue,
                const double* cellVolumes, double* div, int memKind, void* stream);

/* No counterpart in the reference.
 * correction() for callers that may not allocate device memory (CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc): same result as wenoext_correction with device buffers, but
 * this call performs no cudaMalloc / cudaFree / event creation. The cell kernel's scratch (the packed
 * value buffer [nCells][VS] doubles, VS = 1, 2 or 4 for nComp = 1, 2, 3, and its tile counters) lives
 * in a static __device__ array of WENOEXT_NOALLOC_MAX_VALUE_BYTES + 256 bytes that every GPU loading
 * the library holds until exit. Limits: serial plans (no processor patches), device pointers only,
 * nComp <= 3 (1 at order 4 in 3-D), nCells * VS * 8 <= WENOEXT_NOALLOC_MAX_VALUE_BYTES (64^3 cells of
 * a vector field); anything beyond is rejected with WENOEXT_ERR_INVALID and a message. All calls on
 * a device share the array, which keeps the last call's contents: calls that overlap on the GPU
 * (different streams, plans or host threads) race - serialise them. Stream-ordered, returns at once.
 *   vf [nCells][nComp], faceFlux [nFaces], corr [nFaces][nComp] out: DEVICE pointers. */
#define WENOEXT_NOALLOC_MAX_VALUE_BYTES (8u << 20)
int wenoext_correction_noalloc(wenoext_plan* plan, int scheme, int bounded, int nComp,
                               const double* vf, const double* faceFlux, double* corr, void* stream);

/* In-process ranks (several plans in one process, on one or more GPUs; used by
 * the single-GPU tests of the multi-rank path): plans[r] must have myProc == r.
 * Exchanges become device-to-device copies; drive them with the split-phase
 * calls: _begin on every rank, then _exchange on every rank, then _end on every
 * rank (device pointers, nComp <= 3). With NCCL attached, wenoext_correction
 * runs the same three phases back to back. */
int wenoext_attach_local(wenoext_plan** plans, int nRanks);
int wenoext_correction_begin(wenoext_plan* plan, int scheme, int bounded, int nComp,
                             const double* vf, const double* faceFlux,
                             const double* patchNbrField, double* corr, void* stream);
int wenoext_correction_exchange(wenoext_plan* plan);
int wenoext_correction_end(wenoext_plan* plan);

/* Number of this library's kernel launches issued by the last correction call. */
int wenoext_last_launch_count(const wenoext_plan* plan);
/* With timing on, every launch of the fused cell kernel is bracketed by a CUDA
 * event pair on the launching stream. wenoext_kernel_time waits for them and
 * returns the summed device time (ms) and the number of timed launches since
 * timing was enabled or last reset. */
int wenoext_enable_kernel_timing(wenoext_plan* plan, int on);
int wenoext_kernel_time(wenoext_plan* plan, double* msSum, int* nLaunches, int reset);

/* Batched truncated pseudo-inverse on the GPU: what WENOBase::calcMatrix gets
 * from OpenFOAM's SVD(A, minCondition) and VSinvUt() (WENOBase.C:527, :575), for nMat matrices at
 * once (one-sided Jacobi, one warp per matrix). A_i is rows[i] x n row-major at A + start[i]
 * (stencil cells x nDvt); Ainv_i is n x rows[i] row-major at Ainv + start[i]. nZeros[i] = number of
 * singular values <= minCondition * max, condition[i] = max/min over the others (may be NULL),
 * converged[i] = 0 if 60 sweeps did not suffice. Host pointers. */
int wenoext_pinv_batched(int32_t nMat, int32_t n, const int32_t* rows, const int64_t* start,
                         const double* A, double minCondition, double* Ainv, int32_t* nZeros,
                         double* condition, int32_t* converged, int device);

/* Least-squares matrices assembled on the GPU and pseudo-inverted there: the device
 * counterpart of the row loop of WENOBase::calcMatrix (WENOBase.C:410-597) -- geometryWENO::transformIntegral
 * (geometryWENO.C:222-325, divergence theorem over the triangulated cell surface with the 13-point rule of
 * geometryWENO.C:346-460) and WENOBase::calcGeom / addCoeffs (WENOBase.C:600-670) for every (target cell,
 * stencil cell) pair -- followed by the batched pseudo-inverse of wenoext_pinv_batched on the same device
 * buffers: only the mesh, the frames and index lists go up, only the pseudo-inverses come back.
 *   source mesh   points [nPoints][3]; triangles of cell c: triPoints[triStart[c] .. triStart[c+1])[3]
 *                 (fan of every face, in the cell's face order); cellCentres [nSrcCells][3]; cellVolumes
 *   frames        reference frame of a target cell (geometryWENO::initIntegrals, geometryWENO.C:42-219):
 *                 JInv [9] row-major, refPoint [3], refDet, frameCell = its cell in the source mesh,
 *                 volI = its own volume moments, cube of (polOrder+1)^3 (List3D.H:238)
 *   pairs         row p of the LS system: stencil cell pairCell[p] seen from frame pairFrame[p]
 *   matrices      matrix i has rows[i] rows of nDvt at start[i] (doubles, dense: start[i+1] = start[i] +
 *                 rows[i]*nDvt); its row r is pair rowPair[start[i]/nDvt + r]
 * Ainv_i is nDvt x rows[i] row-major at Ainv + start[i]; nZeros, condition (nullable), converged as in
 * wenoext_pinv_batched. rowsOut (nullable) [nPairs][nDvt] and Aout (nullable) return the assembled rows and
 * matrices (tests). Host pointers. The rows are bitwise those of the host setup path. */
typedef struct wenoext_ls_batch {
    int32_t polOrder, nDvt, dim[3];
    int32_t nPoints, nSrcCells;
    const double* points;
    const int64_t* triStart;
    const int32_t* triPoints;
    const double* cellCentres;
    const double* cellVolumes;
    int32_t nFrames;
    const double* JInv;
    const double* refPoint;
    const double* refDet;
    const int32_t* frameCell;
    const double* volI;
    int64_t nPairs;
    const int32_t* pairFrame;
    const int32_t* pairCell;
    int32_t nMat;
    const int32_t* rows;
    const int64_t* start;
    const int64_t* rowPair;
} wenoext_ls_batch;
int wenoext_ls_pinv_batched(const wenoext_ls_batch* batch, double minCondition, double* Ainv,
                            int32_t* nZeros, double* condition, int32_t* converged, double* rowsOut,
                            double* Aout, int device);

#ifdef __cplusplus
}
#endif
#endif /* WENOEXT_B200_H */
