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
