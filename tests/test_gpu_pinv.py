"""Batched truncated pseudo-inverse on the GPU (wenoext_pinv_batched): the
device counterpart of the SVD(A, minCondition) / VSinvUt() pair WENOBase::calcMatrix uses
(libWENOEXT/WENOBase/WENOBase.C:527, :575). Checked against numpy's SVD-based pinv with the
same cut-off rule, and through the setup (device batch vs per-stencil host SVD)."""
import ctypes as C

import numpy as np
import pytest

import wenoext_b200 as wx
from tests import helpers as H
from tests.oracle_binding import Oracle
from tests.test_box_fast import _arrays
from wenoext_b200 import _capi

pytestmark = pytest.mark.gpu


def pinv_batched(mats, minCondition):
    n = mats[0].shape[1]
    rows = np.array([m.shape[0] for m in mats], dtype=np.int32)
    start = np.zeros(len(mats) + 1, dtype=np.int64)
    start[1:] = np.cumsum(rows.astype(np.int64) * n)
    A = np.concatenate([np.ascontiguousarray(m, dtype=np.float64).ravel() for m in mats])
    P = np.empty_like(A)
    nz = np.empty(len(mats), dtype=np.int32)
    conv = np.empty(len(mats), dtype=np.int32)
    cond = np.empty(len(mats), dtype=np.float64)
    wx.check(wx.lib().wenoext_pinv_batched(
        len(mats), n, rows.ctypes.data_as(_capi.i32p), start.ctypes.data_as(_capi.i64p),
        A.ctypes.data_as(_capi.f64p), minCondition, P.ctypes.data_as(_capi.f64p),
        nz.ctypes.data_as(_capi.i32p), cond.ctypes.data_as(_capi.f64p), conv.ctypes.data_as(_capi.i32p), 0))
    out = [P[start[i]:start[i + 1]].reshape(n, rows[i]) for i in range(len(mats))]
    return out, nz, cond, conv


@pytest.mark.parametrize("m,n", [(21, 9), (46, 19), (11, 5), (84, 34), (9, 9)])
def test_against_numpy_pinv(m, n):
    rng = np.random.default_rng(m * 100 + n)
    mats = [rng.standard_normal((m, n)) * 10.0 ** rng.uniform(-3, 3) for _ in range(300)]
    # ragged batch: some matrices with fewer rows (bestConditioned-like stencil sizes)
    mats += [rng.standard_normal((max(n, m - k), n)) for k in range(1, 6)]
    got, nz, cond, conv = pinv_batched(mats, 1e-5)
    assert conv.all() and not nz.any()
    for A, P, c in zip(mats, got, cond):
        ref = np.linalg.pinv(A, rcond=1e-5)
        assert np.abs(P - ref).max() <= 1e-10 * np.abs(ref).max()
        s = np.linalg.svd(A, compute_uv=False)
        assert abs(c - s[0] / s[-1]) <= 1e-8 * c


def test_rank_deficient_is_truncated():
    rng = np.random.default_rng(5)
    mats = []
    for k in range(50):
        A = rng.standard_normal((21, 9))
        A[:, 7] = A[:, 2] * (1.0 + 1e-9 * k)  # (nearly) dependent column: below the cut-off
        if k % 2:
            A[:, 4] = 0.0
        mats.append(A)
    got, nz, cond, conv = pinv_batched(mats, 1e-5)
    assert conv.all()
    for k, (A, P) in enumerate(zip(mats, got)):
        assert nz[k] == (2 if k % 2 else 1)
        ref = np.linalg.pinv(A, rcond=1e-5)
        assert np.abs(P - ref).max() <= 1e-8 * np.abs(ref).max()


def test_bad_arguments_fail_loudly():
    rows = np.array([5], dtype=np.int32)
    start = np.array([0, 44], dtype=np.int64)  # != 5 * 9
    buf = np.zeros(64)
    z = np.zeros(1, dtype=np.int32)
    rc = wx.lib().wenoext_pinv_batched(1, 9, rows.ctypes.data_as(_capi.i32p), start.ctypes.data_as(_capi.i64p),
                                       buf.ctypes.data_as(_capi.f64p), 1e-5, buf.ctypes.data_as(_capi.f64p),
                                       z.ctypes.data_as(_capi.i32p), None, z.ctypes.data_as(_capi.i32p), 0)
    assert rc != 0 and b"start" in wx.lib().wenoext_last_error()


@pytest.mark.parametrize("kw", [dict(n=(5, 5, 5), order=2, tetSplit=6), dict(n=(9, 8, 7), order=2),
                                dict(n=(8, 8, 1), order=3, hi=(0.1, 0.1, 0.01), sides=(0, 0, 0, 0, 3, 3))])
def test_setup_with_device_batch_equals_host_svd(kw):
    host = wx.BoxCase(useClassCache=False, **kw)
    dev = wx.BoxCase(useClassCache=False, gpuPinv=True, device=0, **kw)
    a, b = _arrays(host), _arrays(dev)
    for k in ("css", "ses", "sid", "B", "ibt", "rfa"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["sm"] < 0, b["sm"] < 0)
    worst = 0.0
    for s in np.where(a["sm"] >= 0)[0]:
        ma = a["md"][a["ms"][a["sm"][s]]:a["ms"][a["sm"][s] + 1]]
        mb = b["md"][b["ms"][b["sm"][s]]:b["ms"][b["sm"][s] + 1]]
        assert ma.shape == mb.shape
        worst = max(worst, np.abs(ma - mb).max() / np.abs(ma).max())
    assert worst <= 2e-9  # matrixDB's own de-duplication tolerance bounds the difference
    # and the hot path on the device-built plan matches the oracle run on the same plan
    Cn = dev.C / dev.C.max(axis=0)
    vf = H.vector_field(Cn)
    phi = H.rotating_flux(dev)
    s = wx.WENOUpwindFit(dev, phi, f"{dev.order} 1")
    got = s.correction(vf)
    want = Oracle("c").correction(dev, 0, True, vf, phi)
    assert H.normwise_err(got, want) < 1e-12
    wx.WENOBase.clear()


# ---- least-squares rows and matrices assembled on the device (wenoext_ls_pinv_batched, csrc/ls_rows.cu) ----
import os

_HERE = os.path.dirname(os.path.abspath(__file__))


def _hooks():
    L = C.CDLL(os.path.join(_HERE, "native", "libsetup_hooks.so"))
    dp, ip, lp = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_int64)
    L.hook_mesh_create.restype = C.c_void_p
    L.hook_mesh_create.argtypes = [ip, dp, dp, C.c_int]
    L.hook_mesh_destroy.argtypes = [C.c_void_p]
    L.hook_mesh_sizes.argtypes = [C.c_void_p, ip, ip, lp]
    L.hook_mesh_copy.argtypes = [C.c_void_p, dp, lp, ip, dp, dp]
    L.hook_pair_moments.restype = C.c_int
    L.hook_pair_moments.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, ip, dp, dp, dp, dp, dp, dp, dp]
    return L


def _p(a, t):
    return a.ctypes.data_as(t)


class HookMesh:
    def __init__(self, n, lo, hi, tetSplit):
        self.L = _hooks()
        n3, lo3, hi3 = np.array(n, dtype=np.int32), np.array(lo, dtype=np.float64), np.array(hi, dtype=np.float64)
        self.h = self.L.hook_mesh_create(_p(n3, _capi.i32p), _p(lo3, _capi.f64p), _p(hi3, _capi.f64p), tetSplit)
        nP, nC, nT = C.c_int32(), C.c_int32(), C.c_int64()
        self.L.hook_mesh_sizes(self.h, C.byref(nP), C.byref(nC), C.byref(nT))
        self.nPoints, self.nCells, self.nTri = nP.value, nC.value, nT.value
        self.points = np.empty((self.nPoints, 3))
        self.triStart = np.empty(self.nCells + 1, dtype=np.int64)
        self.triPoints = np.empty((self.nTri, 3), dtype=np.int32)
        self.centres = np.empty((self.nCells, 3))
        self.volumes = np.empty(self.nCells)
        self.L.hook_mesh_copy(self.h, _p(self.points, _capi.f64p), _p(self.triStart, _capi.i64p),
                              _p(self.triPoints, _capi.i32p), _p(self.centres, _capi.f64p), _p(self.volumes, _capi.f64p))

    def close(self):
        self.L.hook_mesh_destroy(self.h)

    def frame(self, order, cellI, cellsJ):
        """host initIntegrals of cell I and transformIntegral of every J in its frame"""
        cube = (order + 1) ** 3
        cj = np.array(cellsJ, dtype=np.int32)
        out = dict(volI=np.empty(cube), JInv=np.empty(9), refPoint=np.empty(3), refDet=np.empty(1), tI=np.empty(3),
                   tJ=np.empty((len(cj), 3)), mom=np.empty((len(cj), cube)))
        rc = self.L.hook_pair_moments(self.h, order, cellI, len(cj), _p(cj, _capi.i32p), *[
            _p(out[k], _capi.f64p) for k in ("volI", "JInv", "refPoint", "refDet", "tI", "tJ", "mom")])
        assert rc == 0
        return out


def _host_row(order, dim, x, mom, volI):
    """WENOBase::calcGeom / addCoeffs (WENOBase.C:600-670) on the host's moments, in the host's operation order"""
    from math import factorial
    s1 = order + 1
    M = mom.reshape(s1, s1, s1)
    V = volI.reshape(s1, s1, s1)

    def ipow(b, e):
        t = 1.0
        for _ in range(e):
            t *= b
        return t
    row = []
    for n in range(dim[0] + 1):
        for m in range(dim[1] + 1):
            for l in range(dim[2] + 1):
                if n + m + l > order or n + m + l == 0:
                    continue
                g = 0.0
                for k in range(n + 1):
                    for q in range(m + 1):
                        for j in range(l + 1):
                            g += (float(factorial(n)) / (float(factorial(k)) * float(factorial(n - k))) * float(factorial(m))
                                  / (float(factorial(q)) * float(factorial(m - q))) * float(factorial(l))
                                  / (float(factorial(j)) * float(factorial(l - j))) * ipow(x[0], k) * ipow(x[1], q)
                                  * ipow(x[2], j) * M[n - k, m - q, l - j])
                v = g - V[n, m, l]
                row.append(0.0 if abs(v) < 1e-9 else v)
    return row


def _ls_batch(mesh, order, dim, K, frames, pairs, mats):
    """frames: list of (cellI, host frame dict); pairs: list of (frame index, cellJ); mats: list of lists of pair
    indices. Returns (rows, A, Ainv list, nZeros, converged)."""
    b = _capi.LsBatch()
    b.polOrder, b.nDvt = order, K
    for k in range(3):
        b.dim[k] = dim[k]
    b.nPoints, b.nSrcCells = mesh.nPoints, mesh.nCells
    keep = [mesh.points, mesh.triStart, mesh.triPoints, mesh.centres, mesh.volumes]
    b.points, b.triStart, b.triPoints = _p(mesh.points, _capi.f64p), _p(mesh.triStart, _capi.i64p), _p(mesh.triPoints, _capi.i32p)
    b.cellCentres, b.cellVolumes = _p(mesh.centres, _capi.f64p), _p(mesh.volumes, _capi.f64p)
    JInv = np.concatenate([f["JInv"] for _, f in frames])
    ref = np.concatenate([f["refPoint"] for _, f in frames])
    det = np.concatenate([f["refDet"] for _, f in frames])
    volI = np.concatenate([f["volI"] for _, f in frames])
    fc = np.array([c for c, _ in frames], dtype=np.int32)
    pf = np.array([p[0] for p in pairs], dtype=np.int32)
    pc = np.array([p[1] for p in pairs], dtype=np.int32)
    rows = np.array([len(m) for m in mats], dtype=np.int32)
    start = np.zeros(len(mats) + 1, dtype=np.int64)
    start[1:] = np.cumsum(rows.astype(np.int64) * K)
    rp = np.array([i for m in mats for i in m], dtype=np.int64)
    keep += [JInv, ref, det, volI, fc, pf, pc, rows, start, rp]
    b.nFrames, b.JInv, b.refPoint, b.refDet = len(frames), _p(JInv, _capi.f64p), _p(ref, _capi.f64p), _p(det, _capi.f64p)
    b.frameCell, b.volI = _p(fc, _capi.i32p), _p(volI, _capi.f64p)
    b.nPairs, b.pairFrame, b.pairCell = len(pairs), _p(pf, _capi.i32p), _p(pc, _capi.i32p)
    b.nMat, b.rows, b.start, b.rowPair = len(mats), _p(rows, _capi.i32p), _p(start, _capi.i64p), _p(rp, _capi.i64p)
    out_rows = np.empty((len(pairs), K))
    A = np.empty(int(start[-1]))
    P = np.empty(int(start[-1]))
    nz = np.empty(len(mats), dtype=np.int32)
    conv = np.empty(len(mats), dtype=np.int32)
    rc = wx.lib().wenoext_ls_pinv_batched(C.byref(b), 1e-5, _p(P, _capi.f64p), _p(nz, _capi.i32p), None,
                                          _p(conv, _capi.i32p), _p(out_rows, _capi.f64p), _p(A, _capi.f64p), 0)
    del keep
    return rc, out_rows, A, [P[start[i]:start[i + 1]].reshape(K, rows[i]) for i in range(len(mats))], nz, conv, start


@pytest.mark.parametrize("n,hi,tets,order,dim,K", [
    ((6, 5, 4), (1.0, 0.7, 0.45), 0, 2, (2, 2, 2), 9),
    ((4, 4, 4), (1.0, 1.0, 1.0), 6, 2, (2, 2, 2), 9),
    ((5, 5, 5), (0.3, 0.5, 1.1), 0, 3, (3, 3, 3), 19),
    ((7, 6, 1), (0.1, 0.1, 0.01), 0, 2, (2, 2, 0), 5),
    ((5, 4, 4), (1.0, 1.0, 1.0), 0, 4, (4, 4, 4), 34),
])
def test_ls_rows_on_device_equal_host_rows(n, hi, tets, order, dim, K):
    """Rows of the LS matrices from the device (transformIntegral + calcGeom per (I, J) pair) against the host's
    geometry code through the test hooks: bitwise (the kernel repeats the host arithmetic without FMA contraction),
    and the assembled matrices and their pseudo-inverses against numpy."""
    mesh = HookMesh(n, (0.0, 0.0, 0.0), hi, tets)
    try:
        rng = np.random.default_rng(7)
        cellsI = rng.choice(mesh.nCells, size=min(6, mesh.nCells), replace=False)
        frames, pairs, mats, want = [], [], [], []
        for fi, ci in enumerate(cellsI):
            # "stencil cells": a random set around the mesh (the row does not care whether J neighbours I)
            cj = [int(c) for c in rng.choice(mesh.nCells, size=min(2 * K + 3, mesh.nCells - 1), replace=False) if c != ci]
            f = mesh.frame(order, int(ci), cj)
            frames.append((int(ci), f))
            base = len(pairs)
            for k, c in enumerate(cj):
                pairs.append((fi, c))
                want.append(_host_row(order, dim, f["tJ"][k] - f["tI"], f["mom"][k], f["volI"]))
            mats.append(list(range(base, base + len(cj))))               # all rows of the frame
            mats.append(list(range(base + len(cj) - 1, base + 1, -2)))   # a shared subset, other order
        rc, rows, A, Pinv, nz, conv, start = _ls_batch(mesh, order, dim, K, frames, pairs, mats)
        assert rc == 0, wx.lib().wenoext_last_error()
        want = np.array(want)
        assert want.shape == rows.shape
        assert np.array_equal(rows, want), np.abs(rows - want).max()
        for i, m in enumerate(mats):
            Ai = A[start[i]:start[i + 1]].reshape(len(m), K)
            assert np.array_equal(Ai, rows[m])
            if len(m) >= K and conv[i] and nz[i] == 0:
                ref = np.linalg.pinv(Ai, rcond=1e-5)
                assert np.abs(Pinv[i] - ref).max() <= 1e-9 * np.abs(ref).max()
        assert conv.all()
    finally:
        mesh.close()


def test_ls_rows_known_answer_on_unit_cubes():
    """Cubes of edge h: in the frame of cell I the moments of cell J are those of a unit cube centred at the offset
    (dx, dy, dz) in cell units, so the row entry (n, m, l) is prod_a int_{d_a-1/2}^{d_a+1/2} t^{e_a} dt minus the same
    with d = 0 (independent of the host code; the 13-point rule is exact to degree 7)."""
    n = (5, 5, 5)
    mesh = HookMesh(n, (0.0, 0.0, 0.0), (1.0, 1.0, 1.0), 0)
    try:
        ci = 2 + 5 * 2 + 25 * 2
        cj = [c for c in range(mesh.nCells) if c != ci][::3]
        f = mesh.frame(2, ci, cj)
        rc, rows, *_ = _ls_batch(mesh, 2, (2, 2, 2), 9, [(ci, f)], [(0, c) for c in cj], [])
        assert rc == 0
        J = f["JInv"].reshape(3, 3)

        def mono(d, e):  # int over [d-1/2, d+1/2] of t^e
            return ((d + 0.5) ** (e + 1) - (d - 0.5) ** (e + 1)) / (e + 1)
        for k, c in enumerate(cj):
            d = J @ (mesh.centres[c] - mesh.centres[ci])  # offset in the frame's units (+-1 per cell)
            col = 0
            for a in range(3):
                for b in range(3):
                    for e in range(3):
                        if a + b + e > 2 or a + b + e == 0:
                            continue
                        want = mono(d[0], a) * mono(d[1], b) * mono(d[2], e) - mono(0, a) * mono(0, b) * mono(0, e)
                        want = 0.0 if abs(want) < 1e-9 else want  # addCoeffs flushes small entries
                        assert abs(rows[k, col] - want) <= 1e-12 * max(1.0, abs(want)), (c, a, b, e)
                        col += 1
    finally:
        mesh.close()


def test_ls_batch_bad_arguments_fail_loudly():
    mesh = HookMesh((3, 3, 3), (0.0, 0.0, 0.0), (1.0, 1.0, 1.0), 0)
    try:
        f = mesh.frame(2, 13, [0, 1])
        rc, *_ = _ls_batch(mesh, 2, (2, 2, 2), 9, [(13, f)], [(0, 0), (0, mesh.nCells)], [])
        assert rc != 0 and b"pair out of range" in wx.lib().wenoext_last_error()
        rc, *_ = _ls_batch(mesh, 7, (2, 2, 2), 9, [(13, f)], [(0, 0)], [])
        assert rc != 0 and b"order" in wx.lib().wenoext_last_error()
    finally:
        mesh.close()
