"""GPU parity tests: the CUDA path, called through the C ABI
(include/wenoext_b200.h), against the CPU oracle on the same plan data and the
same seeded inputs. Tolerance: 1e-12 relative, norm-wise per component
(BASELINE.json north_star "to within 1e-12 relative")."""
import ctypes as C

import numpy as np
import pytest

import wenoext_b200 as wx
from tests import helpers as H
from tests.oracle_binding import Oracle, correction_multirank

pytestmark = pytest.mark.gpu

FIT, FIT01 = 0, 1
TOL = 1e-12
SCHEMES = [(FIT, 0), (FIT, 1), (FIT01, 1)]


def _mk(name):
    if name == "hex8":
        return wx.BoxCase((8, 8, 8), 2, useClassCache=False)
    if name == "cavity":  # BASELINE config 1: 20x20x1, front/back empty -> 2-D, K=5
        return wx.BoxCase((20, 20, 1), 2, hi=(0.1, 0.1, 0.01), sides=(0, 0, 0, 0, 3, 3),
                          useClassCache=False)
    if name == "tets":  # irregular stencil counts
        return wx.BoxCase((4, 4, 4), 2, tetSplit=6, useClassCache=False)
    if name == "tets_bc":  # irregular stencil sizes (bestConditioned)
        return wx.BoxCase((3, 3, 3), 2, tetSplit=6, bestConditioned=True, useClassCache=False)
    if name == "hex_o3":  # WENOUpwindFit 3 x: K=19, N=46
        return wx.BoxCase((7, 7, 7), 3, useClassCache=False)
    if name == "hex_o1":  # order 1: K=3
        return wx.BoxCase((8, 8, 8), 1, useClassCache=False)
    if name == "hex_o4":  # order 4 in 3-D: K=34, one component per matrix read
        return wx.BoxCase((10, 10, 10), 4, useClassCache=False)
    if name == "quad_o1":  # 2-D order 1: K=2
        return wx.BoxCase((12, 12, 1), 1, sides=(0, 0, 0, 0, 3, 3), useClassCache=False)
    if name == "quad_o4":  # 2-D order 4: K=14
        return wx.BoxCase((14, 14, 1), 4, sides=(0, 0, 0, 0, 3, 3), useClassCache=False)
    if name == "quad_o5":  # 2-D order 5: K=20
        return wx.BoxCase((14, 14, 1), 5, sides=(0, 0, 0, 0, 3, 3), useClassCache=False)
    if name == "cyclic":  # 3DMesh-cyclic-like: x cyclic
        return wx.BoxCase((8, 6, 6), 2, sides=(2, 2, 0, 0, 0, 0), useClassCache=False)
    if name == "cyclic2d":  # 2DMesh-cyclic-like
        return wx.BoxCase((12, 12, 1), 2, sides=(2, 2, 2, 2, 3, 3), useClassCache=False)
    raise KeyError(name)


_cache = {}


def case_of(name):
    if name not in _cache:
        _cache[name] = _mk(name)
    return _cache[name]


def _fields(case, nComp, seed=1):
    Cn = (case.C - case.C.min(axis=0)) / (case.C.max(axis=0) - case.C.min(axis=0) + 1e-300)
    v3 = H.vector_field(Cn, seed=seed)
    if nComp == 1:
        return v3[:, 0].copy()
    reps = [v3 * (1.0 + 0.37 * i) + 0.11 * i for i in range((nComp + 2) // 3)]
    return np.ascontiguousarray(np.concatenate(reps, axis=1)[:, :nComp])


def _plan(case):
    return wx.WENOBase.instance(case)


@pytest.mark.parametrize("name", ["hex8", "cavity", "tets", "tets_bc", "hex_o3", "cyclic",
                                  "cyclic2d", "hex_o1", "hex_o4", "quad_o1", "quad_o4", "quad_o5"])
@pytest.mark.parametrize("scheme,bounded", SCHEMES)
@pytest.mark.parametrize("nComp", [1, 3])
def test_parity_host_buffers(name, scheme, bounded, nComp):
    case = case_of(name)
    vf = _fields(case, nComp)
    phi = H.rotating_flux(case)
    phi[::17] = 0.0  # third branch of WENOUpwindFit.C:258-281
    want = Oracle("c").correction(case, scheme, bounded, vf, phi)
    got = _plan(case).correction(scheme, bounded, vf, phi)
    assert got.shape == want.shape
    assert H.normwise_err(got, want) < TOL
    # interpolated values: upwind value + correction
    nI = case.nInternalFaces
    up = np.where(phi[:nI] >= 0, case.owner[:nI], case.neighbour)
    fv = np.empty((nI,) + vf.shape[1:])
    _plan(case).correction(scheme, bounded, vf, phi, faceValue=fv)
    # flux == 0: weights() = pos(0) = 1 -> owner value, zero correction
    assert H.normwise_err(fv, vf[up] + want[:nI]) < TOL


@pytest.mark.parametrize("nComp", [2, 6, 9])  # symmTensor / tensor: several batches
def test_parity_more_components(nComp):
    case = case_of("hex8")
    vf = _fields(case, nComp)
    phi = H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.02)
    for scheme, bounded in SCHEMES:
        want = Oracle("c").correction(case, scheme, bounded, vf, phi)
        got = _plan(case).correction(scheme, bounded, vf, phi)
        assert H.normwise_err(got, want) < TOL


def test_parity_device_buffers_and_stream():
    import torch
    case = case_of("hex8")
    vf = _fields(case, 3)
    phi = H.rotating_flux(case)
    want = Oracle("c").correction(case, FIT, 1, vf, phi)
    dv = torch.from_numpy(vf).cuda()
    dp = torch.from_numpy(phi).cuda()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        out = _plan(case).correction(FIT, 1, dv, dp, stream=st)
    st.synchronize()
    assert H.normwise_err(out.cpu().numpy(), want) < TOL
    assert _plan(case).last_launch_count() >= 2


def test_against_blaze_and_long_double():
    """the same GPU result against the reference's own arithmetic library (Blaze)
    and against long-double accumulation"""
    case = case_of("hex8")
    vf = _fields(case, 3)
    phi = H.rotating_flux(case)
    got = _plan(case).correction(FIT, 0, vf, phi)
    for variant in ("blaze", "ld"):
        want = Oracle(variant).correction(case, FIT, 0, vf, phi)
        assert H.normwise_err(got, want) < TOL


def test_constant_and_zero_fields_exact():
    case = case_of("hex8")
    phi = H.rotating_flux(case)
    for val in (0.0, 3.25):
        vf = np.full((case.nCells, 3), val)
        for scheme, bounded in SCHEMES:
            got = _plan(case).correction(scheme, bounded, vf, phi)
            assert np.all(got == 0.0)


def test_wenodict_params():
    case = case_of("hex8")
    vf = _fields(case, 3)
    phi = H.rotating_flux(case)
    plan = _plan(case)
    try:
        for p, dm, eps in ((1.0, 1.0, 1e-6), (2.7, 10.0, 1e-20), (0.0, 1000.0, 1e-40)):
            plan.set_params(p, dm, eps)
            want = Oracle("c").correction(case, FIT, 0, vf, phi, p=p, dm=dm, eps=eps)
            got = plan.correction(FIT, 0, vf, phi)
            assert H.normwise_err(got, want) < TOL
    finally:
        plan.set_params()


def test_fit01_bug_compat_switch():
    case = case_of("cavity")
    vf = H.zalesak_like(case.C / 0.1)
    phi = H.rotating_flux(case)
    plan = _plan(case)
    try:
        for bc in (1, 0):
            plan.set_fit01_bug_compat(bc)
            want = Oracle("c").correction(case, FIT01, 1, vf, phi, bugCompat=bc)
            got = plan.correction(FIT01, 1, vf, phi)
            assert H.normwise_err(got, want) < TOL
            nI = case.nInternalFaces
            up = np.where(phi[:nI] > 0, case.owner[:nI], case.neighbour)
            face = vf[up] + got[:nI]
            if bc == 0:
                assert face.min() > -1e-9 and face.max() < 1 + 1e-9
    finally:
        plan.set_fit01_bug_compat(1)


def test_empty_cells_get_zero():
    case = case_of("hex8")
    if not case.emptyCells.any():
        pytest.skip("no Cell::empty in this case")
    vf = _fields(case, 1)
    phi = H.face_flux(case.Sf)
    got = _plan(case).correction(FIT, 0, vf, phi)
    nI = case.nInternalFaces
    up = np.where(phi[:nI] > 0, case.owner[:nI], case.neighbour)
    assert np.all(got[:nI][case.emptyCells[up]] == 0.0)


def test_scheme_objects_mirror_reference_api():
    case = case_of("hex8")
    vf = _fields(case, 3)
    phi = H.rotating_flux(case)
    s = wx.WENOUpwindFit(case, phi, "2 1")
    assert s.corrected() and s.typeName == "WENOUpwindFit"
    assert np.array_equal(s.weights(vf), (phi >= 0).astype(float))
    want = Oracle("c").correction(case, FIT, 1, vf, phi)
    assert H.normwise_err(s.correction(vf), want) < TOL
    s01 = wx.WENOUpwindFit01(case, phi, "2")
    want = Oracle("c").correction(case, FIT01, 1, vf, phi)
    assert H.normwise_err(s01.correction(vf), want) < TOL
    with pytest.raises(wx.WenoextError):
        wx.WENOUpwindFit(case, phi, "3 1")  # preprocessed for order 2


@pytest.mark.parametrize("nProc", [(2, 1, 1), (2, 2, 1), (2, 2, 2)])
@pytest.mark.parametrize("scheme,bounded", SCHEMES)
def test_multirank_in_process(nProc, scheme, bounded):
    """All ranks of a decomposed case as plans on ONE GPU (exchanges are
    device-to-device copies): halo exchange, interior/boundary split, processor
    patches and the patch limiter against the oracle's MPI emulation."""
    import torch
    n = (8, 8, 6)
    R = int(np.prod(nProc))
    cases = [wx.BoxCase(n, 2, nProc=nProc, myProc=r, useClassCache=False) for r in range(R)]
    wx.exchange_halo_lists(cases)
    rng = np.random.default_rng(3)
    vfs = [_fields(c, 3, seed=10 + r) for r, c in enumerate(cases)]
    fluxes = [H.rotating_flux(c) for c in cases]
    for f in fluxes:
        f[rng.uniform(size=f.shape) < 0.02] = 0.0
    wants = correction_multirank(Oracle("c"), cases, scheme, bounded, vfs, fluxes)
    plans = [wx.Plan(c.desc) for c in cases]
    try:
        wx.attach_local(plans)
        dvs = [torch.from_numpy(v).cuda() for v in vfs]
        dfs = [torch.from_numpy(f).cuda() for f in fluxes]
        outs = wx.correction_local(plans, scheme, bounded, dvs, dfs)
        torch.cuda.synchronize()
        for r in range(R):
            assert plans[r].info.nHaloCells > 0
            assert H.normwise_err(outs[r].cpu().numpy(), wants[r]) < TOL
    finally:
        for p in plans:
            p.destroy()


def test_no_gpu_no_fallback_message():
    # the product path has no CPU fallback: a bad device index must fail loudly
    case = case_of("hex8")
    with pytest.raises(wx.WenoextError):
        wx.Plan(case.desc, device=63)


@pytest.mark.parametrize("sides", [(0,) * 6, (2, 2, 0, 0, 0, 0)])
@pytest.mark.parametrize("scheme,bounded,nComp", [(FIT, 1, 3), (FIT01, 1, 1), (FIT, 0, 2)])
def test_pipelined_host_path_equals_device_path(sides, scheme, bounded, nComp):
    """Host buffers on a serial plan take the slab-pipelined path (H2D, kernel and D2H
    overlapped); it must return exactly what the device-buffer path returns."""
    import torch
    case = wx.BoxCase((48, 48, 40), 2, sides=sides)
    plan = wx.Plan(case.desc)
    vf = _fields(case, nComp, seed=5)
    flux = H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.02)
    host = np.asarray(plan.correction(scheme, bounded, vf, flux)).reshape(case.nFaces, -1)
    dvf = torch.from_numpy(np.ascontiguousarray(vf)).cuda()
    dfl = torch.from_numpy(flux).cuda()
    out = torch.empty((case.nFaces, nComp), dtype=torch.float64, device="cuda")
    plan.correction(scheme, bounded, dvf, dfl, out=out)
    np.testing.assert_array_equal(host, out.cpu().numpy())
    ref = Oracle("c").correction(case, scheme, bounded, vf, flux)
    assert H.normwise_err(host, ref) <= TOL
    plan.destroy()


@pytest.mark.parametrize("env", [
    {"WENOEXT_CELLS": "1"},                       # one cell per warp (meshes with > 8 stencils per cell)
    {"WENOEXT_SLOTS": "1"},                       # ring of one chunk step: no prefetch at all
    {"WENOEXT_SLOTS": "2", "WENOEXT_WARPS": "3"},
    {"WENOEXT_SMEM_KB": "48"},                    # few warps fit
    {"WENOEXT_GRID": "3"},                        # 3 CTAs take all tiles
    {"WENOEXT_CELLS": "1", "WENOEXT_SLOTS": "2", "WENOEXT_WARPS": "5"},
])
@pytest.mark.parametrize("name", ["hex8", "tets_bc", "hex_o3", "cavity"])
def test_launch_geometries(monkeypatch, env, name):
    """The kernel's pipeline protocol must not depend on the launch geometry the plan
    picks: force other cell-per-warp mappings, ring depths, warp and CTA counts."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    case = case_of(name)
    plan = wx.Plan(case.desc)  # reads the overrides at creation
    vf = _fields(case, 3, seed=2)
    flux = H.rotating_flux(case)
    for scheme, bounded in SCHEMES:
        got = plan.correction(scheme, bounded, vf, flux)
        ref = Oracle("c").correction(case, scheme, bounded, vf, flux)
        assert H.normwise_err(got, ref) <= TOL, (env, name, scheme, bounded)
    plan.destroy()


def test_plans_on_two_devices_in_one_process():
    """The opt-in to >48 KB dynamic shared memory is per function AND per device: a second plan on
    another GPU of the same process must launch as well (kernels.cu launchT)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    case = case_of("hex8")
    vf = _fields(case, 3, seed=2)
    flux = H.rotating_flux(case)
    want = Oracle("c").correction(case, FIT, 1, vf, flux)
    for dev in (0, 1, 0):
        plan = wx.Plan(case.desc, device=dev)
        try:
            got = plan.correction(FIT, 1, vf, flux)
            assert H.normwise_err(np.asarray(got).reshape(want.shape), want) < TOL
        finally:
            plan.destroy()


@pytest.mark.parametrize("host", [True, False], ids=["host", "device"])
def test_set_face_flux_once_per_time_step(host):
    """faceFlux_ is the same field for every correction() of a time step (WENOUpwindFit.H:75): uploaded
    once with wenoext_set_face_flux, later calls pass faceFlux = NULL and give the same bits."""
    import torch
    case = wx.BoxCase((40, 40, 36), 2)  # large enough for the slab-pipelined host path
    plan = wx.Plan(case.desc)
    try:
        flux = H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.02)
        fields = [_fields(case, 3, seed=3), np.ascontiguousarray(_fields(case, 1, seed=4)).reshape(-1)]
        conv = (lambda x: x) if host else (lambda x: torch.from_numpy(x).cuda())
        back = (lambda x: np.asarray(x)) if host else (lambda x: x.cpu().numpy())
        want = [back(plan.correction(FIT, 1, conv(v), conv(flux))) for v in fields]
        with pytest.raises(wx.WenoextError):
            plan.correction(FIT, 1, conv(fields[0]), None)
        plan.set_face_flux(conv(flux))
        for v, w in zip(fields, want):
            got = back(plan.correction(FIT, 1, conv(v), None))
            np.testing.assert_array_equal(got, w)
        plan.set_face_flux(None)
        with pytest.raises(wx.WenoextError):
            plan.correction(FIT, 1, conv(fields[0]), None)
    finally:
        plan.destroy()


def test_pinned_host_buffers_from_the_library():
    """wenoext_host_alloc memory through the host-buffer call: same result as pageable numpy arrays."""
    case = wx.BoxCase((40, 40, 36), 2)
    plan = wx.Plan(case.desc)
    try:
        vf = _fields(case, 3, seed=6)
        flux = H.rotating_flux(case)
        want = np.asarray(plan.correction(FIT, 1, vf, flux))
        pv, pf = wx.pinned_empty(vf.shape), wx.pinned_empty(flux.shape)
        po = wx.pinned_empty((case.nFaces, 3))
        pv[...] = vf
        pf[...] = flux
        po[...] = np.nan
        plan.correction(FIT, 1, pv, pf, out=po)
        np.testing.assert_array_equal(np.asarray(po), want)
    finally:
        plan.destroy()


@pytest.mark.parametrize("name", ["hex8", "cyclic", "tets_bc"])
def test_labels64_descriptor_gives_the_same_bits(name):
    """wenoext_plan_create_labels64: the descriptor of a WM_LABEL_SIZE=64 build (every label array int64)
    creates the same plan: bitwise the same correction as the 32-bit descriptor, and within TOL of the oracle."""
    case = case_of(name)
    vf, phi = _fields(case, 3), H.rotating_flux(case)
    d64, keep = wx.labels64(case.desc)
    p64 = wx.Plan(d64)
    for scheme, bounded in SCHEMES:
        got = np.asarray(p64.correction(scheme, bounded, vf, phi))
        assert np.array_equal(got, np.asarray(_plan(case).correction(scheme, bounded, vf, phi)))
        assert H.normwise_err(got, Oracle("c").correction(case, scheme, bounded, vf, phi)) < TOL
    del keep


@pytest.mark.parametrize("slabs", ["3", "6", "8"])
def test_first_host_buffer_call_of_a_fresh_plan_with_many_slabs(monkeypatch, slabs):
    """The slab pipeline of a host-buffer call (copies, packs on their own stream, slab kernels on alternating streams,
    output copies) on the FIRST call of a fresh plan, whose staging buffers hold nothing yet: a read ahead of its
    data cannot hide behind an identical previous result. 6 and more slabs are tapered."""
    import torch
    monkeypatch.setenv("WENOEXT_E2E_SLABS", slabs)
    case = wx.BoxCase((48, 48, 40), 2)
    plan = wx.Plan(case.desc)
    try:
        vf = _fields(case, 3, seed=9)
        flux = H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.02)
        host = np.asarray(plan.correction(FIT, 1, vf, flux)).reshape(case.nFaces, -1)
        dev = plan.correction(FIT, 1, torch.from_numpy(np.ascontiguousarray(vf)).cuda(), torch.from_numpy(flux).cuda())
        np.testing.assert_array_equal(host, dev.cpu().numpy())
    finally:
        plan.destroy()
