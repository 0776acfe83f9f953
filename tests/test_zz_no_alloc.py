"""wenoext_correction_noalloc (include/wenoext_b200.h): the entry point that allocates nothing, with the
cell kernel's scratch in a static __device__ array. No counterpart in the reference.
Checked through the C ABI against the ordinary entry point
(wenoext_correction, device buffers) on two calls in a row with different inputs, and against the
oracle; limits are rejected with a message. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import wenoext_b200 as wx
from tests import helpers as H
from tests.oracle_binding import Oracle

pytestmark = pytest.mark.gpu
TOL = 1e-12
FIT, FIT01 = 0, 1


def _noalloc(plan, scheme, bounded, dvf, dflux, dout):
    import torch
    nComp = 1 if dvf.ndim == 1 else int(dvf.shape[1])
    st = torch.cuda.current_stream(dvf.device).cuda_stream
    return wx.lib().wenoext_correction_noalloc(plan._h, scheme, int(bounded), nComp, dvf.data_ptr(),
                                               dflux.data_ptr(), dout.data_ptr(), st)


@pytest.mark.parametrize("sides", [(0,) * 6, (2, 2, 0, 0, 0, 0)], ids=["walls", "cyclic_x"])
def test_noalloc_equals_ordinary_entry_point_two_calls_in_a_row(sides):
    import torch
    case = wx.BoxCase((10, 8, 8), 2, sides=sides, useClassCache=False)
    plan = wx.Plan(case.desc)
    orc = Oracle("c")
    try:
        X = case.C / case.C.max(axis=0)
        inputs = [
            (FIT, 1, H.vector_field(X, seed=11), H.rotating_flux(case)),
            (FIT01, 1, H.zalesak_like(X) + 0.0, H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.03)),
            (FIT, 0, H.vector_field(X, seed=12)[:, :2].copy(), H.face_flux(case.Sf)),
        ]
        for scheme, bounded, vf, flux in inputs:  # back to back: the scratch keeps the last call's contents
            dvf = torch.from_numpy(np.ascontiguousarray(vf)).cuda()
            dfl = torch.from_numpy(np.ascontiguousarray(flux)).cuda()
            shape = (case.nFaces,) if vf.ndim == 1 else (case.nFaces, vf.shape[1])
            out = torch.full(shape, float("nan"), dtype=torch.float64, device="cuda")
            rc = _noalloc(plan, scheme, bounded, dvf, dfl, out)
            assert rc == 0, wx.lib().wenoext_last_error().decode()
            assert plan.last_launch_count() >= 2
            torch.cuda.synchronize()
            got = out.cpu().numpy()
            ref = np.asarray(plan.correction(scheme, bounded, dvf, dfl).cpu().numpy())
            torch.cuda.synchronize()
            assert np.isfinite(got).all()
            assert H.normwise_err(got, ref.reshape(got.shape)) <= TOL
            want = orc.correction(case, scheme, bounded, vf, flux)
            assert H.normwise_err(got, np.asarray(want).reshape(got.shape)) <= TOL
            # and once more with the same inputs: same bits
            out2 = torch.full(shape, float("nan"), dtype=torch.float64, device="cuda")
            assert _noalloc(plan, scheme, bounded, dvf, dfl, out2) == 0
            torch.cuda.synchronize()
            np.testing.assert_array_equal(out2.cpu().numpy(), got)
    finally:
        plan.destroy()


def test_noalloc_does_not_allocate_device_memory():
    """Free device memory is the same before and after a burst of calls (the ordinary entry point was
    never called on this plan, so nothing is pre-allocated for it either)."""
    import torch
    case = wx.BoxCase((12, 12, 12), 2)
    plan = wx.Plan(case.desc)
    try:
        X = case.C / case.C.max(axis=0)
        dvf = torch.from_numpy(H.vector_field(X, seed=1)).cuda()
        dfl = torch.from_numpy(H.rotating_flux(case)).cuda()
        out = torch.empty((case.nFaces, 3), dtype=torch.float64, device="cuda")
        assert _noalloc(plan, FIT, 1, dvf, dfl, out) == 0  # first launch: module/attribute set-up
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        for _ in range(20):
            assert _noalloc(plan, FIT, 1, dvf, dfl, out) == 0
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0
        want = Oracle("c").correction(case, FIT, 1, dvf.cpu().numpy(), dfl.cpu().numpy())
        assert H.normwise_err(out.cpu().numpy(), want) <= TOL
    finally:
        plan.destroy()


def test_noalloc_rejects_what_it_cannot_do():
    import torch
    L = wx.lib()
    # more packed values than the static scratch holds: 66^3 cells x 4 doubles x 8 B > 8 MiB
    big = wx.BoxCase((66, 66, 66), 2)
    plan = wx.Plan(big.desc)
    try:
        dvf = torch.zeros((big.nCells, 3), dtype=torch.float64, device="cuda")
        dfl = torch.zeros(big.nFaces, dtype=torch.float64, device="cuda")
        out = torch.zeros((big.nFaces, 3), dtype=torch.float64, device="cuda")
        assert _noalloc(plan, FIT, 1, dvf, dfl, out) == 1
        msg = L.wenoext_last_error().decode()
        assert "WENOEXT_NOALLOC_MAX_VALUE_BYTES" in msg and str(big.nCells) in msg
        # the same mesh with a scalar field fits (VS = 1)
        dsc = torch.zeros(big.nCells, dtype=torch.float64, device="cuda")
        osc = torch.zeros(big.nFaces, dtype=torch.float64, device="cuda")
        assert _noalloc(plan, FIT, 1, dsc, dfl, osc) == 0
        torch.cuda.synchronize()
        # more components than one batch
        d6 = torch.zeros((big.nCells, 6), dtype=torch.float64, device="cuda")
        o6 = torch.zeros((big.nFaces, 6), dtype=torch.float64, device="cuda")
        assert _noalloc(plan, FIT, 1, d6, dfl, o6) == 1
        assert "nComp" in L.wenoext_last_error().decode()
        assert L.wenoext_correction_noalloc(plan._h, FIT, 1, 1, None, dfl.data_ptr(), osc.data_ptr(), None) == 1
    finally:
        plan.destroy()
    # decomposed plans use plan-owned exchange buffers: refused
    cases = [wx.BoxCase((8, 6, 6), 2, nProc=(2, 1, 1), myProc=r) for r in range(2)]
    wx.exchange_halo_lists(cases)
    p0 = wx.Plan(cases[0].desc)
    try:
        dvf = torch.zeros(cases[0].nCells, dtype=torch.float64, device="cuda")
        dfl = torch.zeros(cases[0].nFaces, dtype=torch.float64, device="cuda")
        out = torch.zeros(cases[0].nFaces, dtype=torch.float64, device="cuda")
        assert _noalloc(p0, FIT, 1, dvf, dfl, out) == 1
        assert "serial plans only" in L.wenoext_last_error().decode()
    finally:
        p0.destroy()
