"""wenoext_b200 — B200-native WENOEXT reconstruction hot path (host-side mirror).

Python mirror of the reference's plugin interface for this one path, on top of
the C ABI in include/wenoext_b200.h:

  reference (OpenFOAM plugin)                         here
  ------------------------------------------------    ---------------------------
  WENOBase::instance(mesh, polOrder)                  WENOBase.instance(case)
    libWENOEXT/WENOBase/WENOBase.H:281-294
  WENOUpwindFit<Type>(mesh, faceFlux, Istream)        WENOUpwindFit(case, faceFlux, "2 1")
    libWENOEXT/WENOUpwindFit/WENOUpwindFit.H:198-211
  weights(vf) / corrected() / correction(vf)          same names, same meaning
    WENOUpwindFit.H:223-242
  WENOUpwindFit01<Type>(mesh, faceFlux, Istream)      WENOUpwindFit01(case, faceFlux, "3")
    libWENOEXT/WENOUpwindFit01/WENOUpwindFit01.H:183-201
  system/WENODict: p, dm, epsilon                     case.WENODict
    libWENOEXT/WENOBase/WENOCoeff.C:87-89

All arithmetic happens in libwenoext_b200.so (hand-written sm_100a CUDA). Nothing
here computes a correction on the CPU and nothing imports oracle/.
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import (MEM_DEVICE, MEM_HOST, PATCH_CYCLIC, PATCH_EMPTY, PATCH_GENERIC,
                    PATCH_PROCESSOR, SCHEME_UPWINDFIT, SCHEME_UPWINDFIT01, WenoextError, check,
                    check_setup, lib)

__all__ = ["pinned_empty", "BoxCase", "CachedCase", "WENOBase", "Plan", "WENOUpwindFit", "WENOUpwindFit01", "WenoextError",
           "exchange_halo_lists", "exchange_halo_lists_distributed", "attach_local",
           "correction_local", "nccl_unique_id", "SCHEME_UPWINDFIT", "SCHEME_UPWINDFIT01",
           "PATCH_GENERIC", "PATCH_PROCESSOR", "PATCH_CYCLIC", "PATCH_EMPTY"]


def _np_view(ptr, shape, dtype):
    n = int(np.prod(shape))
    if n == 0 or not ptr:
        return np.zeros(shape, dtype=dtype)
    ctype = {np.int32: C.c_int32, np.int64: C.c_int64, np.float64: C.c_double}[dtype]
    arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ctype)), shape=(n,))
    return arr.reshape(shape)


class BoxCase:
    """A generated box case: mesh + the WENOBase preprocessing output.

    Stand-in for an OpenFOAM case directory (blockMesh box + constant/WENOBase<N>),
    because OpenFOAM is not available here. The preprocessing is the host C++
    restatement of WENOBase::WENOBase (wenoext_b200/csrc/setup/), run once.
    """

    def __init__(self, n, order, lo=(0.0, 0.0, 0.0), hi=(1.0, 1.0, 1.0),
                 sides=(0, 0, 0, 0, 0, 0), nProc=(1, 1, 1), myProc=0, tetSplit=0,
                 extendRatio=2.5, bestConditioned=False, maxCondition=1e-5, checkCondition=True,
                 nThreads=0, useClassCache=True, WENODict=None, gpuPinv=False, device=0):
        L = lib()
        box = _capi.BoxSpec()
        box.n[:] = list(n)
        box.lo[:] = list(lo)
        box.hi[:] = list(hi)
        box.sideType[:] = list(sides)
        box.tetSplit = tetSplit
        box.nProc[:] = list(nProc)
        prm = _capi.SetupParams(order, extendRatio, int(bestConditioned), maxCondition,
                                int(checkCondition), nThreads, int(useClassCache), int(gpuPinv), device)
        h = C.c_void_p()
        check_setup(L.wenoext_setup_box(C.byref(box), C.byref(prm), myProc, C.byref(h)))
        self._h = h
        self.order = order
        self.nProcs = int(np.prod(nProc))
        self.myProc = myProc
        # WENODict run-time factors and their defaults (WENOCoeff.C:41-48, :87-89)
        self.WENODict = {"p": 4.0, "dm": 1000.0, "epsilon": 1e-40}
        if WENODict:
            self.WENODict.update(WENODict)
        self._refresh()

    def _refresh(self):
        L = lib()
        self.desc = _capi.PlanDesc()
        check_setup(L.wenoext_setup_desc(self._h, C.byref(self.desc)))
        d = self.desc
        self.nCells, self.nFaces, self.nInternalFaces = d.nCells, d.nFaces, d.nInternalFaces
        self.nDvt = d.nDvt
        self.owner = _np_view(d.owner, (d.nFaces,), np.int32)
        self.neighbour = _np_view(d.neighbour, (d.nInternalFaces,), np.int32)
        self.patches = [(d.patches[i].type, d.patches[i].start, d.patches[i].size, d.patches[i].nbr)
                        for i in range(d.nPatches)]
        css = _np_view(d.cellStencilStart, (d.nCells + 1,), np.int64)
        ses = _np_view(d.stencilEntryStart, (d.nStencils + 1,), np.int64)
        nEnt = int(ses[-1])
        self.stencilsID = _np_view(d.stencilsID, (nEnt,), np.int32)
        # Cell::empty cells get zero correction (WENOCoeff.C:333)
        self.emptyCells = self.stencilsID[ses[css[:-1]]] == _capi.CELL_EMPTY
        cc, sf = _capi.f64p(), _capi.f64p()
        check_setup(L.wenoext_setup_geometry(self._h, C.byref(cc), C.byref(sf)))
        self.C = _np_view(cc, (d.nCells, 3), np.float64)       # mesh.C()
        self.Sf = _np_view(sf, (d.nFaces, 3), np.float64)      # mesh.Sf()

    def write_cache(self, directory, labelBytes=4, simdDoubles=8):
        """WENOBase::writeList (WENOBase.C:1265-1312): the eleven files of
        constant/WENOBase<order>/ in the reference's binary format."""
        check_setup(lib().wenoext_cache_write(self._h, str(directory).encode(), labelBytes, simdDoubles))

    def mesh_desc(self):
        """Mesh addressing (owner, neighbour, cells, patches) as a wenoext_mesh_desc."""
        m = _capi.MeshDesc()
        check_setup(lib().wenoext_setup_mesh(self._h, C.byref(m)))
        return m

    def halo_request(self, proc):
        ids, n = _capi.i32p(), C.c_int64()
        check_setup(lib().wenoext_setup_halo_request(self._h, proc, C.byref(ids), C.byref(n)))
        return _np_view(ids, (n.value,), np.int32).copy()

    def set_send_halo(self, proc, ids):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        check_setup(lib().wenoext_setup_set_send_halo(self._h, proc, ids.ctypes.data_as(_capi.i32p),
                                                ids.size))
        self._refresh()

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().wenoext_setup_destroy(self._h)
                self._h = None
        except Exception:
            pass


class CachedCase(BoxCase):
    """A case whose WENOBase data come from an on-disk cache constant/WENOBase<order>/
    (WENOBase::readList, WENOBase.C:1196-1262) instead of the preprocessing. The mesh
    addressing is not in the cache; pass it as a wenoext_mesh_desc (`mesh`, e.g.
    another case's mesh_desc()) or as arrays."""

    def __init__(self, directory, mesh=None, owner=None, neighbour=None, patches=(), myProc=0,
                 labelBytes=0, patchNbrCell=None, geometry=None, WENODict=None, caseOrder=None):
        keep = []
        if caseOrder is not None:
            # `directory` is an OpenFOAM case (or processor<N>) directory: the cache from
            # constant/WENOBase<order>/, the addressing from constant/polyMesh/
            h = C.c_void_p()
            check_setup(lib().wenoext_case_read(str(directory).encode(), caseOrder, myProc, labelBytes,
                                          C.byref(h)))
            self._finish(h, myProc, WENODict, geometry)
            return
        if mesh is None:
            owner = np.ascontiguousarray(owner, dtype=np.int32)
            neighbour = np.ascontiguousarray(neighbour, dtype=np.int32)
            parr = (_capi.Patch * max(1, len(patches)))(*[_capi.Patch(*p) for p in patches])
            mesh = _capi.MeshDesc()
            mesh.nCells = int(max(owner.max(), neighbour.max() if neighbour.size else -1)) + 1
            mesh.nFaces, mesh.nInternalFaces = owner.size, neighbour.size
            mesh.owner = owner.ctypes.data_as(_capi.i32p)
            mesh.neighbour = neighbour.ctypes.data_as(_capi.i32p)
            mesh.nPatches, mesh.patches = len(patches), parr
            if patchNbrCell is not None:
                pn = np.ascontiguousarray(patchNbrCell, dtype=np.int32)
                mesh.patchNbrCell = pn.ctypes.data_as(_capi.i32p)
                keep.append(pn)
            keep += [owner, neighbour, parr]
        h = C.c_void_p()
        check_setup(lib().wenoext_cache_read(str(directory).encode(), C.byref(mesh), myProc, labelBytes,
                                       C.byref(h)))
        self._finish(h, myProc, WENODict, geometry)

    @classmethod
    def from_case(cls, caseDir, order, myProc=0, **kw):
        """`caseDir/constant/WENOBase<order>/` + `caseDir/constant/polyMesh/`."""
        return cls(caseDir, caseOrder=order, myProc=myProc, **kw)

    def _finish(self, h, myProc, WENODict, geometry):
        self._h = h
        self.myProc = myProc
        self.WENODict = {"p": 4.0, "dm": 1000.0, "epsilon": 1e-40}
        if WENODict:
            self.WENODict.update(WENODict)
        self._refresh()
        self.order = self.desc.polOrder
        self.nProcs = self.desc.nProcs
        if geometry is not None:  # (C, Sf) for synthetic fields; not read by the hot path
            self.C, self.Sf = geometry

    def _refresh(self):
        super()._refresh()
        if self.C.size and not self.C.any():
            self.C = None
            self.Sf = None


def exchange_halo_lists(cases):
    """distributeStencils (WENOBase.C:334-407) for in-process ranks: hand every
    rank's request list to the owner rank."""
    for r, cr in enumerate(cases):
        for p, cp in enumerate(cases):
            if p != r:
                cp.set_send_halo(r, cr.halo_request(p))


def exchange_halo_lists_distributed(case):
    """Same, one process per rank, lists shipped with torch.distributed."""
    import torch.distributed as dist

    req = [case.halo_request(p).tolist() for p in range(case.nProcs)]
    allreq = [None] * case.nProcs
    dist.all_gather_object(allreq, req)
    for p in range(case.nProcs):
        if p != case.myProc:
            case.set_send_halo(p, np.asarray(allreq[p][case.myProc], dtype=np.int32))


class _PinnedArray(np.ndarray):
    """ndarray over wenoext_host_alloc memory; `_owner` frees it when the last view is gone"""
    _owner = None


class _PinnedOwner:
    def __init__(self, addr):
        self.addr = addr

    def __del__(self):
        try:
            lib().wenoext_host_free(self.addr)
        except Exception:
            pass


def pinned_empty(shape, dtype=np.float64):
    """numpy array in page-locked host memory from wenoext_host_alloc (freed with the array): what a
    host-side caller stages its fields through so that the library can overlap copies and kernels."""
    count = int(np.prod(shape))
    n = count * np.dtype(dtype).itemsize
    p = C.c_void_p()
    check(lib().wenoext_host_alloc(n, C.byref(p)))
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=count).reshape(shape).view(_PinnedArray)
    arr._owner = _PinnedOwner(p.value)
    return arr


def labels64(desc):
    """The descriptor as a WM_LABEL_SIZE=64 build of OpenFOAM would hand it over (wenoext_plan_desc64: every
    label array int64). Returns (PlanDesc64, keepalive): the arrays live in `keepalive`."""
    import numpy as np
    d64, keep = _capi.PlanDesc64(), []

    def wide(ptr, n):
        if not ptr or n <= 0:
            return None
        a = np.ctypeslib.as_array(ptr, shape=(int(n),)).astype(np.int64)
        keep.append(a)
        return a.ctypes.data_as(_capi.i64p)

    for f in ("polOrder", "nDvt", "nCells", "nFaces", "nInternalFaces", "nPatches", "nStencils",
              "nUniqueMatrices", "nProcs", "myProc"):
        setattr(d64, f, getattr(desc, f))
    for k in range(3):
        d64.dim[k] = desc.dim[k]
    nC, nF, nI, nS = desc.nCells, desc.nFaces, desc.nInternalFaces, desc.nStencils
    d64.owner, d64.neighbour = wide(desc.owner, nF), wide(desc.neighbour, nI)
    d64.cellFaceStart = wide(desc.cellFaceStart, nC + 1)
    d64.cellFaces = wide(desc.cellFaces, desc.cellFaceStart[nC])
    if desc.nPatches > 0:
        pa = (_capi.Patch64 * desc.nPatches)()
        for i in range(desc.nPatches):
            q = desc.patches[i]
            pa[i].type, pa[i].start, pa[i].size, pa[i].nbr = q.type, q.start, q.size, q.nbr
        keep.append(pa)
        d64.patches = pa
    d64.cellStencilStart, d64.stencilEntryStart = desc.cellStencilStart, desc.stencilEntryStart
    nE = desc.stencilEntryStart[nS]
    d64.stencilsID, d64.cellToProcMap = wide(desc.stencilsID, nE), wide(desc.cellToProcMap, nE)
    d64.stencilMatrix = wide(desc.stencilMatrix, nS)
    d64.matrixStart, d64.matrixCols = desc.matrixStart, wide(desc.matrixCols, desc.nUniqueMatrices)
    d64.matrixData, d64.B, d64.intBasTrans, d64.refFacAr = desc.matrixData, desc.B, desc.intBasTrans, desc.refFacAr
    P = desc.nProcs
    d64.sendProcList, d64.receiveProcList = wide(desc.sendProcList, P), wide(desc.receiveProcList, P)
    d64.receiveHaloSize, d64.sendHaloStart = wide(desc.receiveHaloSize, P), desc.sendHaloStart
    if desc.sendHaloStart and P > 0:
        d64.sendHaloCellID = wide(desc.sendHaloCellID, desc.sendHaloStart[P])
    d64.patchNbrCell = wide(desc.patchNbrCell, nF - nI)
    keep.append(desc)
    return d64, keep


class Plan:
    """Device plan (wenoext_plan): owns all device memory of one (mesh, order)."""

    def __init__(self, desc, device=0):
        self._h = C.c_void_p()
        self.device = device
        if isinstance(desc, _capi.PlanDesc64):  # descriptor of a WM_LABEL_SIZE=64 build
            check(lib().wenoext_plan_create_labels64(C.byref(desc), device, C.byref(self._h)))
        else:
            check(lib().wenoext_plan_create(C.byref(desc), device, C.byref(self._h)))
        self.info = _capi.PlanInfo()
        check(lib().wenoext_plan_get_info(self._h, C.byref(self.info)))
        self.nCells, self.nFaces = int(self.info.nCells), int(self.info.nFaces)
        self.nInternalFaces = int(self.info.nInternalFaces)

    def set_params(self, p=4.0, dm=1000.0, epsilon=1e-40):
        check(lib().wenoext_set_params(self._h, p, dm, epsilon))

    def set_fit01_bug_compat(self, on):
        check(lib().wenoext_set_fit01_bug_compat(self._h, int(on)))

    def attach_nccl(self, uid, rank, nRanks):
        check(lib().wenoext_attach_nccl(self._h, uid, rank, nRanks))

    def set_face_flux(self, faceFlux, stream=None):
        """wenoext_set_face_flux: upload the flux field once per time step; later calls may pass
        faceFlux=None (WENOUpwindFit.H:75: faceFlux_ is the same field for every correction() of a step)."""
        if faceFlux is None:
            check(lib().wenoext_set_face_flux(self._h, None, MEM_HOST, None))
            self._step_flux = None
            return
        pf, kind, keep = self._ptr(faceFlux)
        st = 0
        if kind == MEM_DEVICE:
            import torch
            st = (stream if stream is not None else torch.cuda.current_stream(faceFlux.device)).cuda_stream
        check(lib().wenoext_set_face_flux(self._h, pf, kind, st))
        self._step_flux = keep if kind == MEM_DEVICE else None  # a device pointer is kept by the library

    def alg_bytes(self, nComp):
        return int(self.info.algBytesFixed + nComp * self.info.algBytesPerComp)

    def enable_kernel_timing(self, on=True):
        check(lib().wenoext_enable_kernel_timing(self._h, int(on)))

    def kernel_time(self, reset=True):
        """(summed device ms, launches) of the fused cell kernel since timing was
        enabled / last reset (CUDA events on the launching stream)."""
        ms, n = C.c_double(), C.c_int()
        check(lib().wenoext_kernel_time(self._h, C.byref(ms), C.byref(n), int(reset)))
        return ms.value, n.value

    def last_launch_count(self):
        return lib().wenoext_last_launch_count(self._h)

    @staticmethod
    def _ptr(x):
        """(pointer, memKind, keepalive) of a numpy array or a torch tensor."""
        if x is None:
            return None, None, None
        if isinstance(x, np.ndarray):
            a = np.ascontiguousarray(x, dtype=np.float64)
            return a.ctypes.data, MEM_HOST, a
        # torch tensor
        t = x.contiguous()
        if t.dtype.is_floating_point and t.element_size() != 8:
            raise WenoextError("fields must be float64 (the path computes in fp64)")
        return t.data_ptr(), (MEM_DEVICE if t.is_cuda else MEM_HOST), t

    def correction(self, scheme, bounded, vf, faceFlux, patchNbrField=None, out=None,
                   faceValue=None, stream=None):
        """wenoext_correction / wenoext_interpolate through the C ABI.

        numpy arrays are HOST buffers (copies happen inside the call); torch CUDA
        tensors are used in place on `stream` (default: torch's current stream)."""
        nComp = 1 if vf.ndim == 1 else int(vf.shape[1])
        pv, kind, kv = self._ptr(vf)
        pf, kindf, kf = self._ptr(faceFlux)  # None: the field given to set_face_flux
        pn, _, kn = self._ptr(patchNbrField)
        if faceFlux is not None and kind != kindf:
            raise WenoextError("vf and faceFlux must live in the same memory space")
        shape = (self.nFaces,) if vf.ndim == 1 else (self.nFaces, nComp)
        if out is None:
            if kind == MEM_HOST and isinstance(vf, np.ndarray):
                out = np.empty(shape, dtype=np.float64)
            else:
                import torch
                out = torch.empty(shape, dtype=torch.float64, device=vf.device)
        po, kindo, ko = self._ptr(out)
        if kindo != kind:
            raise WenoextError("out must live in the same memory space as vf")
        st = 0
        if kind == MEM_DEVICE:
            import torch
            st = (stream if stream is not None else torch.cuda.current_stream(vf.device)).cuda_stream
        if faceValue is not None:
            pfv, _, kfv = self._ptr(faceValue)
            check(lib().wenoext_interpolate(self._h, scheme, int(bounded), nComp, pv, pf, pn, po,
                                            pfv, kind, st))
        else:
            check(lib().wenoext_correction(self._h, scheme, int(bounded), nComp, pv, pf, pn, po,
                                           kind, st))
        return out

    def div(self, scheme, bounded, vf, faceFlux, bndValue, cellVolumes, patchNbrField=None, out=None,
            stream=None):
        """wenoext_div: fvc::div(faceFlux, vf) with the WENO scheme, formed on the device."""
        nComp = 1 if vf.ndim == 1 else int(vf.shape[1])
        pv, kind, kv = self._ptr(vf)
        pf, _, kf = self._ptr(faceFlux)
        pn, _, kn = self._ptr(patchNbrField)
        pb, _, kb = self._ptr(bndValue)
        pvol, _, kvol = self._ptr(cellVolumes)
        if out is None:
            if isinstance(vf, np.ndarray):
                out = np.empty(vf.shape, dtype=np.float64)
            else:
                import torch
                out = torch.empty(vf.shape, dtype=torch.float64, device=vf.device)
        po, _, ko = self._ptr(out)
        st = 0
        if kind == MEM_DEVICE:
            import torch
            st = (stream if stream is not None else torch.cuda.current_stream(vf.device)).cuda_stream
        check(lib().wenoext_div(self._h, scheme, int(bounded), nComp, pv, pf, pn, pb, pvol, po, kind, st))
        return out

    def destroy(self):
        if self._h:
            lib().wenoext_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


def nccl_unique_id():
    """ncclGetUniqueId through the C ABI (rank 0 creates it, the host side ships it)."""
    buf = C.create_string_buffer(128)
    check(lib().wenoext_nccl_unique_id(buf))
    return buf.raw


def attach_local(plans):
    """wenoext_attach_local: all ranks of a decomposed case as plans in ONE
    process (exchanges become device-to-device copies)."""
    arr = (C.c_void_p * len(plans))(*[p._h for p in plans])
    check(lib().wenoext_attach_local(arr, len(plans)))


def correction_local(plans, scheme, bounded, vfs, fluxes, patchNbrFields=None, outs=None):
    """Split-phase correction() of in-process ranks (torch CUDA tensors): phase k
    runs on every rank before phase k+1, like the reference's MPI ranks do between
    collectData / swapData (WENOCoeff.C:224-296, WENOUpwindFit.C:326-385)."""
    import torch
    L = lib()
    R = len(plans)
    keep = []
    if outs is None:
        outs = []
        for r in range(R):
            nComp = 1 if vfs[r].ndim == 1 else int(vfs[r].shape[1])
            shape = (plans[r].nFaces,) if vfs[r].ndim == 1 else (plans[r].nFaces, nComp)
            outs.append(torch.empty(shape, dtype=torch.float64, device=vfs[r].device))
    for r in range(R):
        vf, fl = vfs[r].contiguous(), fluxes[r].contiguous()
        pn = None if patchNbrFields is None else patchNbrFields[r].contiguous()
        keep += [vf, fl, pn]
        nComp = 1 if vf.ndim == 1 else int(vf.shape[1])
        st = torch.cuda.current_stream(vf.device).cuda_stream
        check(L.wenoext_correction_begin(plans[r]._h, scheme, int(bounded), nComp, vf.data_ptr(),
                                         fl.data_ptr(), None if pn is None else pn.data_ptr(),
                                         outs[r].data_ptr(), st))
    for r in range(R):
        check(L.wenoext_correction_exchange(plans[r]._h))
    for r in range(R):
        check(L.wenoext_correction_end(plans[r]._h))
    return outs


class WENOBase:
    """Process-wide registry of device plans, the counterpart of the reference's
    singleton `WENOBase::instance(mesh, polOrder)` (WENOBase.H:281-294). OpenFOAM
    constructs a new scheme object on every fvm::div call, so the expensive state
    must not live in the scheme. Keyed by (case, order, device)
    instead of "first caller wins"."""

    _plans = {}

    @classmethod
    def instance(cls, case, device=0):
        key = (id(case), case.order, device)
        if key not in cls._plans:
            plan = Plan(case.desc, device)
            cls._plans[key] = (plan, case)
        return cls._plans[key][0]

    @classmethod
    def clear(cls):
        for plan, _ in cls._plans.values():
            plan.destroy()
        cls._plans.clear()


def _tokens(spec):
    if isinstance(spec, str):
        return spec.split()
    if isinstance(spec, (list, tuple)):
        return [str(s) for s in spec]
    return [str(spec)]


def _read_bool(tok):
    # OpenFOAM readBool accepts 0/1, true/false, on/off, yes/no
    t = tok.lower()
    if t in ("1", "true", "on", "yes", "y", "t"):
        return True
    if t in ("0", "false", "off", "no", "n", "f", "none"):
        return False
    raise WenoextError(f"bad bool token '{tok}'")


class WENOUpwindFit:
    """`Gauss WENOUpwindFit <order> <bounded>` (README.md:93-99 of the reference).

    Mirrors WENOUpwindFit<Type> (WENOUpwindFit.H:57-243): constructed from the mesh,
    the face flux and the scheme tokens; weights() is pure upwind, corrected() is
    true, correction(vf) is the explicit high-order correction per face."""

    typeName = "WENOUpwindFit"
    _scheme = SCHEME_UPWINDFIT

    def __init__(self, case, faceFlux, spec, device=0):
        tok = _tokens(spec)
        self.polOrder_ = int(float(tok[0]))            # readScalar(is), WENOUpwindFit.H:207
        self.limFac_ = self._read_limiter(tok)         # readBool(is),   WENOUpwindFit.H:208
        if self.polOrder_ != case.order:
            raise WenoextError(f"case was preprocessed for order {case.order}, "
                               f"scheme asks for {self.polOrder_}")
        self.mesh_ = case
        self.faceFlux_ = faceFlux
        self.plan_ = WENOBase.instance(case, device)
        wd = case.WENODict
        self.plan_.set_params(wd["p"], wd["dm"], wd["epsilon"])

    @staticmethod
    def _read_limiter(tok):
        if len(tok) < 2:
            raise WenoextError("WENOUpwindFit needs '<order> <bounded>'")
        return _read_bool(tok[1])

    def weights(self, vf=None):
        """pos(faceFlux_) (WENOUpwindFit.H:223-229); pos(0) == 1 in OpenFOAM."""
        f = self.faceFlux_
        if isinstance(f, np.ndarray):
            return (f >= 0).astype(np.float64)
        return (f >= 0).to(f.dtype)

    def corrected(self):
        return True

    def correction(self, vf, patchNbrField=None, out=None):
        return self.plan_.correction(self._scheme, self.limFac_, vf, self.faceFlux_,
                                     patchNbrField, out=out)

    def interpolate(self, vf, patchNbrField=None):
        """Internal-face values weights*vf_P + (1-weights)*vf_N + correction, formed
        on the device (what surfaceInterpolationScheme::interpolate returns)."""
        nComp = 1 if vf.ndim == 1 else int(vf.shape[1])
        nI = self.mesh_.nInternalFaces
        shape = (nI,) if vf.ndim == 1 else (nI, nComp)
        if isinstance(vf, np.ndarray):
            fv = np.empty(shape, dtype=np.float64)
        else:
            import torch
            fv = torch.empty(shape, dtype=torch.float64, device=vf.device)
        corr = self.plan_.correction(self._scheme, self.limFac_, vf, self.faceFlux_,
                                     patchNbrField, faceValue=fv)
        return fv, corr


class WENOUpwindFit01(WENOUpwindFit):
    """`Gauss WENOUpwindFit01 <order>`: always limited to [0, 1]
    (WENOUpwindFit01.H:169-201, WENOUpwindFit01.C:37-212)."""

    typeName = "WENOUpwindFit01"
    _scheme = SCHEME_UPWINDFIT01

    @staticmethod
    def _read_limiter(tok):
        return True
