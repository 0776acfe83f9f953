"""ctypes binding of the CPU oracle (oracle/). Test infrastructure: only tests/,
__graft_entry__.smoke() and bench.py's CPU legs may import this."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libweno_oracle.so")
BLAZE_SO = os.path.join(ROOT, "oracle", "_ref", "libweno_oracle_blaze.so")

vp = C.c_void_p
_dp = lambda a: a.ctypes.data_as(vp) if a is not None else None  # noqa: E731


def _f64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


class Oracle:
    """variant: 'c' (plain C, double), 'ld' (long double accumulation), 'blaze'
    (reference's vendored Blaze for the three matrix products)."""

    def __init__(self, variant="c"):
        self.variant = variant
        self.L = C.CDLL(ORACLE_SO)
        pre = "weno_oracle_ld" if variant == "ld" else "weno_oracle"
        self._pol = getattr(self.L, pre + "_getWENOPol")
        self._A = getattr(self.L, pre + "_phaseA")
        self._B = getattr(self.L, pre + "_phaseB")
        self._pol.argtypes = [vp, C.c_int, C.c_double, C.c_double, C.c_double, vp, vp, vp]
        self._polr = getattr(self.L, pre + "_getWENOPol_range")
        self._polr.argtypes = [vp, C.c_int, C.c_double, C.c_double, C.c_double, vp, vp, vp,
                               C.c_int, C.c_int]
        self._facesr = getattr(self.L, pre + "_internalFaces_range")
        self._facesr.argtypes = [vp, C.c_int, vp, vp, vp, C.c_int, C.c_int]
        self._limr = getattr(self.L, pre + "_limiterFitInternal_range")
        self._limr.argtypes = [vp, C.c_int, vp, vp, vp, C.c_int, C.c_int]
        self._A.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        self._B.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp]
        self.BL = None
        if variant == "blaze":
            self.BL = C.CDLL(BLAZE_SO)
            self.BL.weno_oracle_blaze_create.restype = vp
            self.BL.weno_oracle_blaze_create.argtypes = [vp]
            self.BL.weno_oracle_blaze_destroy.argtypes = [vp]
            self.BL.weno_oracle_blaze_getWENOPol.argtypes = [vp, C.c_int, C.c_double, C.c_double,
                                                             C.c_double, vp, vp, vp]
            self.BL.weno_oracle_blaze_getWENOPol_range.argtypes = [
                vp, C.c_int, C.c_double, C.c_double, C.c_double, vp, vp, vp, C.c_int, C.c_int]
        self._handles = {}

    def _blaze_handle(self, case):
        k = id(case.desc)
        if k not in self._handles:
            self._handles[k] = (self.BL.weno_oracle_blaze_create(C.addressof(case.desc)), case)
        return self._handles[k][0]

    def getWENOPol(self, case, vf, halo=None, p=4.0, dm=1000.0, eps=1e-40):
        vf = _f64(vf)
        nComp = 1 if vf.ndim == 1 else vf.shape[1]
        halo = _f64(halo)
        cw = np.zeros((case.nCells, case.nDvt, nComp))
        if self.variant == "blaze":
            rc = self.BL.weno_oracle_blaze_getWENOPol(self._blaze_handle(case), nComp, p, dm, eps,
                                                      _dp(vf), _dp(halo), _dp(cw))
        else:
            rc = self._pol(C.addressof(case.desc), nComp, p, dm, eps, _dp(vf), _dp(halo), _dp(cw))
        assert rc == 0
        return cw

    def phaseA(self, case, faceFlux, cw):
        nComp = cw.shape[2]
        corr = np.zeros((case.nFaces, nComp))
        ud = np.zeros((max(case.nFaces - case.nInternalFaces, 1), nComp))
        rc = self._A(C.addressof(case.desc), nComp, _dp(_f64(faceFlux)), _dp(cw), _dp(corr), _dp(ud))
        assert rc == 0
        return corr, ud

    def phaseB(self, case, scheme, bounded, vf, faceFlux, cw, corr, pnf=None, recv=None,
               bugCompat=1):
        nComp = cw.shape[2]
        vf = _f64(vf)
        rc = self._B(C.addressof(case.desc), scheme, int(bounded), nComp, bugCompat, _dp(vf),
                     _dp(_f64(faceFlux)), _dp(cw), _dp(_f64(pnf)), _dp(_f64(recv)), _dp(corr))
        assert rc == 0
        return corr

    def correction(self, case, scheme, bounded, vf, faceFlux, p=4.0, dm=1000.0, eps=1e-40,
                   bugCompat=1):
        """Serial correction(); returns [nFaces] or [nFaces][nComp] like vf."""
        vf = _f64(vf)
        cw = self.getWENOPol(case, vf, None, p, dm, eps)
        corr, ud = self.phaseA(case, faceFlux, cw)
        self.phaseB(case, scheme, bounded, vf, faceFlux, cw, corr, None, ud, bugCompat)
        return corr[:, 0] if vf.ndim == 1 else corr


def correction_threaded(orc, case, bounded, vf, faceFlux, nThreads, p=4.0, dm=1000.0, eps=1e-40):
    """WENOUpwindFit correction() of a SERIAL case with the cell loop
    (WENOCoeff.C:326-351), the internal-face loop (WENOUpwindFit.C:256-282) and the
    internal-face limiter (:81-108,:172-179) split over `nThreads` host threads
    (ctypes releases the GIL). Stand-in for `mpirun -np <cores>` of the reference on
    a box without MPI; used by bench.py's CPU legs. Coupled patches are not handled
    (serial box cases have none)."""
    from concurrent.futures import ThreadPoolExecutor

    vf = _f64(vf)
    flux = _f64(faceFlux)
    nComp = 1 if vf.ndim == 1 else vf.shape[1]
    d = C.addressof(case.desc)
    cw = np.empty((case.nCells, case.nDvt, nComp))
    corr = np.zeros((case.nFaces, nComp))
    blaze = orc.variant == "blaze"
    h = orc._blaze_handle(case) if blaze else None

    def cells(r):
        a, b = r
        if blaze:
            return orc.BL.weno_oracle_blaze_getWENOPol_range(h, nComp, p, dm, eps, _dp(vf), None,
                                                             _dp(cw), a, b)
        return orc._polr(d, nComp, p, dm, eps, _dp(vf), None, _dp(cw), a, b)

    def faces(r):
        a, b = r
        rc = orc._facesr(d, nComp, _dp(flux), _dp(cw), _dp(corr), a, b)
        if bounded and not rc:
            rc = orc._limr(d, nComp, _dp(vf), _dp(flux), _dp(corr), a, b)
        return rc

    def split(n):
        e = np.linspace(0, n, nThreads + 1).astype(int)
        return [(int(e[i]), int(e[i + 1])) for i in range(nThreads)]

    if nThreads <= 1:
        assert cells((0, case.nCells)) == 0 and faces((0, case.nInternalFaces)) == 0
    else:
        with ThreadPoolExecutor(nThreads) as ex:
            assert all(rc == 0 for rc in ex.map(cells, split(case.nCells)))
            assert all(rc == 0 for rc in ex.map(faces, split(case.nInternalFaces)))
    return corr[:, 0] if vf.ndim == 1 else corr


def correction_multirank(orc, cases, scheme, bounded, vfs, fluxes, p=4.0, dm=1000.0, eps=1e-40,
                         bugCompat=1):
    """In-process emulation of an MPI run of the reference: collectData
    (WENOCoeff.C:224-296), per-rank getWENOPol, swapData (WENOUpwindFit.C:326-385),
    per-rank second half + limiter. `cases` already had exchange_halo_lists()."""
    R = len(cases)
    vfs = [_f64(v).reshape(c.nCells, -1) for v, c in zip(vfs, cases)]
    nComp = vfs[0].shape[1]
    halos = []
    for r, c in enumerate(cases):
        segs = []
        for q in range(R):
            n = c.desc.receiveHaloSize[q]
            if n:
                send_ids = _send_list(cases[q], r)
                assert len(send_ids) == n
                segs.append(vfs[q][send_ids])
        halos.append(np.concatenate(segs) if segs else np.zeros((0, nComp)))
    cws, corrs, uds = [], [], []
    for r, c in enumerate(cases):
        cw = orc.getWENOPol(c, vfs[r], halos[r], p, dm, eps)
        corr, ud = orc.phaseA(c, fluxes[r], cw)
        cws.append(cw), corrs.append(corr), uds.append(ud)
    outs = []
    for r, c in enumerate(cases):
        nI = c.nInternalFaces
        recv = np.zeros_like(uds[r])
        pnf = np.zeros_like(uds[r])
        for (ptype, start, size, nbr) in c.patches:
            if ptype != 1 or size == 0:
                continue
            q = cases[nbr]
            (qs,) = [ps for (pt, ps, pz, pn) in q.patches if pt == 1 and pn == r]
            recv[start - nI:start - nI + size] = uds[nbr][qs - q.nInternalFaces:qs - q.nInternalFaces + size]
            pnf[start - nI:start - nI + size] = vfs[nbr][q.owner[qs:qs + size]]
        outs.append(orc.phaseB(c, scheme, bounded, vfs[r], fluxes[r], cws[r], corrs[r], pnf, recv,
                               bugCompat))
    return outs


def _send_list(case, to):
    d = case.desc
    a, b = d.sendHaloStart[to], d.sendHaloStart[to + 1]
    return np.array([d.sendHaloCellID[i] for i in range(a, b)], dtype=np.int64)


def correction_distributed(orc, case, scheme, bounded, vf, flux, p=4.0, dm=1000.0, eps=1e-40,
                           bugCompat=1, group=None):
    """One process per rank (torch.distributed, any backend with send/recv): the same
    two exchange steps as the reference's MPI run - collectData (WENOCoeff.C:224-296)
    and swapData (WENOUpwindFit.C:326-385) - around the per-rank oracle calls.
    `case` already had exchange_halo_lists_distributed(). `group`: a process group whose backend
    moves CPU tensors (gloo) when the default group is NCCL."""
    import torch
    import torch.distributed as dist

    me, R = case.myProc, case.nProcs
    vf = _f64(vf).reshape(case.nCells, -1)
    nComp = vf.shape[1]
    # exchange #1: cell values of the halo lists
    reqs, recv_bufs = [], {}
    for q in range(R):
        n = case.desc.receiveHaloSize[q]
        if q != me and n:
            recv_bufs[q] = torch.empty((n, nComp), dtype=torch.float64)
            reqs.append(dist.irecv(recv_bufs[q], src=q, group=group))
    for q in range(R):
        ids = _send_list(case, q)
        if q != me and len(ids):
            reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(vf[ids])), dst=q, group=group))
    for r in reqs:
        r.wait()
    segs = [recv_bufs[q].numpy() for q in range(R) if q in recv_bufs]
    halo = np.concatenate(segs) if segs else np.zeros((0, nComp))
    cw = orc.getWENOPol(case, vf, halo, p, dm, eps)
    corr, ud = orc.phaseA(case, flux, cw)
    # exchange #2: processor-patch face values and the patch-neighbour cell values
    nI = case.nInternalFaces
    recv, pnf = np.zeros_like(ud), np.zeros_like(ud)
    reqs, bufs = [], []
    for (ptype, start, size, nbr) in case.patches:
        if ptype != 1 or size == 0:
            continue
        mine = np.concatenate([ud[start - nI:start - nI + size], vf[case.owner[start:start + size]]])
        theirs = torch.empty(mine.shape, dtype=torch.float64)
        bufs.append((start, size, theirs))
        reqs.append(dist.irecv(theirs, src=nbr, group=group))
        reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(mine)), dst=nbr, group=group))
    for r in reqs:
        r.wait()
    for start, size, theirs in bufs:
        t = theirs.numpy()
        recv[start - nI:start - nI + size] = t[:size]
        pnf[start - nI:start - nI + size] = t[size:]
    return orc.phaseB(case, scheme, bounded, vf, flux, cw, corr, pnf, recv, bugCompat)


# ----------------------------------------------------------------------------------------------------
# The reference's OWN hot-path code (oracle/_ref/libweno_ref_<isa>.so, built by oracle/Makefile from the
# unmodified translation units under /root/reference against oracle/of_stub). Parity pin + CPU baseline.
def _ref_so_path():
    """AVX-512 build when the host has it (what the reference's -march=native gives), else AVX2."""
    flags = ""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    flags = line
                    break
    except OSError:
        pass
    need = ("avx512f", "avx512dq", "avx512cd", "avx512bw", "avx512vl")
    isa = "v4" if all(x in flags.split() for x in need) else "v3"
    if os.environ.get("WENO_REF_ISA"):
        isa = os.environ["WENO_REF_ISA"]
    return os.path.join(ROOT, "oracle", "_ref", f"libweno_ref_{isa}.so"), isa


def reference_build_available():
    return os.path.exists(_ref_so_path()[0])


_TYPE_OF_NCOMP = {1: "scalar", 3: "vector", 6: "symmTensor", 9: "tensor"}


class ReferenceBuild:
    """Runs WENOUpwindFit<Type>::correction / WENOUpwindFit01<Type>::correction of the reference
    itself on one case (serial) or a list of per-rank cases (one thread per MPI rank)."""

    def __init__(self, cases, lib_path=None):
        path, self.isa = _ref_so_path()
        if lib_path:
            path, self.isa = lib_path, "shim"
        self.L = C.CDLL(path)
        self.L.weno_ref_create.restype = vp
        self.L.weno_ref_create.argtypes = [C.c_int, vp]
        self.L.weno_ref_destroy.argtypes = [vp]
        self.L.weno_ref_last_error.restype = C.c_char_p
        self.L.weno_ref_provenance.restype = C.c_char_p
        self.L.weno_ref_set_dict.argtypes = [vp, C.c_int, C.c_double, C.c_int, C.c_double, C.c_int,
                                             C.c_double]
        self.L.weno_ref_correction.argtypes = [vp, C.c_char_p, C.c_char_p, vp, vp, vp, vp, C.c_int, vp]
        self.cases = list(cases) if isinstance(cases, (list, tuple)) else [cases]
        self.serial_input = not isinstance(cases, (list, tuple))
        R = len(self.cases)
        self._descs = (vp * R)(*[C.addressof(c.desc) for c in self.cases])
        self.h = self.L.weno_ref_create(R, self._descs)
        if not self.h:
            raise RuntimeError(self.L.weno_ref_last_error().decode())
        self.provenance = self.L.weno_ref_provenance().decode()
        self.last_seconds = None

    def __del__(self):
        if getattr(self, "h", None):
            self.L.weno_ref_destroy(self.h)
            self.h = None

    def set_dict(self, p=None, dm=None, epsilon=None):
        """system/WENODict (WENOCoeff.C:87-89); None keeps the reference's default."""
        self.L.weno_ref_set_dict(self.h, p is not None, p or 0.0, dm is not None, dm or 0.0,
                                 epsilon is not None, epsilon or 0.0)

    def correction(self, scheme, order, bounded, vfs, fluxes, typeName=None, repeat=1,
                   want_weights=False):
        """scheme: 0 / 'WENOUpwindFit' or 1 / 'WENOUpwindFit01'. Returns corr per rank
        ([nFaces] or [nFaces][nComp], like vf), or the single array for a serial case."""
        name = {0: "WENOUpwindFit", 1: "WENOUpwindFit01"}.get(scheme, scheme)
        spec = f"{name} {order}" + (f" {int(bounded)}" if name == "WENOUpwindFit" else "")
        if self.serial_input:
            vfs, fluxes = [vfs], [fluxes]
        R = len(self.cases)
        vfs = [_f64(v) for v in vfs]
        flat = vfs[0].ndim == 1
        vfs2 = [v.reshape(c.nCells, -1) for v, c in zip(vfs, self.cases)]
        nComp = vfs2[0].shape[1]
        typeName = typeName or _TYPE_OF_NCOMP[nComp]
        fl = [_f64(f) for f in fluxes]
        corr = [np.zeros((c.nFaces, nComp)) for c in self.cases]
        wts = [np.zeros(c.nFaces) for c in self.cases] if want_weights else None
        arr = lambda xs: (vp * R)(*[x.ctypes.data for x in xs])  # noqa: E731
        secs = C.c_double(0.0)
        rc = self.L.weno_ref_correction(self.h, spec.encode(), typeName.encode(), arr(vfs2), arr(fl),
                                        arr(corr), arr(wts) if wts else None, repeat, C.byref(secs))
        if rc != 0:
            raise RuntimeError(self.L.weno_ref_last_error().decode())
        self.last_seconds = secs.value
        out = [c[:, 0] if flat else c for c in corr]
        if want_weights:
            return (out[0], wts[0]) if self.serial_input else (out, wts)
        return out[0] if self.serial_input else out


class ExpandedCase:
    """The same case with one private copy of the pseudo-inverse per (cell, stencil) instead of the
    matrixDB-style de-duplicated pool: the CPU then streams the same matrix bytes as the GPU's
    per-cell records ("expanded per-cell matrices"). Only the matrix pool of the
    descriptor changes; everything else is shared with `case` (which must stay alive)."""

    def __init__(self, case):
        from wenoext_b200 import _capi
        d = case.desc
        nS = int(d.nStencils)
        sm = np.ctypeslib.as_array(d.stencilMatrix, shape=(nS,))
        nU = int(d.nUniqueMatrices)
        ms = np.ctypeslib.as_array(d.matrixStart, shape=(nU + 1,))
        mc = np.ctypeslib.as_array(d.matrixCols, shape=(nU,))
        md = np.ctypeslib.as_array(d.matrixData, shape=(int(ms[-1]),))
        live = np.where(sm >= 0)[0]
        sizes = (ms[1:] - ms[:-1])[sm[live]]
        self._start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        self._cols = np.ascontiguousarray(mc[sm[live]], dtype=np.int32)
        # gather: element k of copy i comes from ms[sm[live[i]]] + k
        src0 = np.repeat(ms[sm[live]], sizes)
        within = np.arange(int(self._start[-1]), dtype=np.int64) - np.repeat(self._start[:-1], sizes)
        self._data = np.ascontiguousarray(md[src0 + within])
        self._sm = np.full(nS, -1, dtype=np.int32)
        self._sm[live] = np.arange(len(live), dtype=np.int32)
        self.desc = _capi.PlanDesc()
        C.memmove(C.addressof(self.desc), C.addressof(d), C.sizeof(_capi.PlanDesc))
        self.desc.stencilMatrix = self._sm.ctypes.data_as(_capi.i32p)
        self.desc.nUniqueMatrices = len(live)
        self.desc.matrixStart = self._start.ctypes.data_as(_capi.i64p)
        self.desc.matrixCols = self._cols.ctypes.data_as(_capi.i32p)
        self.desc.matrixData = self._data.ctypes.data_as(_capi.f64p)
        self._base = case
        for k in ("nCells", "nFaces", "nInternalFaces", "nDvt", "owner", "neighbour", "patches", "C", "Sf",
                  "myProc", "nProcs"):
            if hasattr(case, k):
                setattr(self, k, getattr(case, k))
        self.matrix_bytes = int(self._data.nbytes)


SHIM_SO = os.path.join(ROOT, "shim", "_build", "libweno_shim_check.so")


class ShimBuild(ReferenceBuild):
    """The same harness around the B200 shim classes (shim/openfoam/*.C compiled against oracle/of_stub
    by `make -C shim check`): WENOUpwindFit<Type> / WENOUpwindFit01<Type> selected at run time, with
    correction() going through libwenoext_b200.so to the GPU. Serial cases."""

    def __init__(self, case):
        super().__init__(case, lib_path=SHIM_SO)
