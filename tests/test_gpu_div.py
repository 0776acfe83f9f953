"""wenoext_div: fvc::div(faceFlux, vf) = fvc::surfaceIntegrate(faceFlux *
interpolate(vf)) formed on the device, against a numpy restatement of OpenFOAM's
surfaceIntegrate loops fed with (a) the GPU's own face values (bit-exact: same summation
order) and (b) the oracle's correction (1e-12)."""
import numpy as np
import pytest

import wenoext_b200 as wx
from tests import helpers as H
from tests.oracle_binding import Oracle

pytestmark = pytest.mark.gpu
FIT, FIT01 = 0, 1


def surface_integrate(case, flux, corr, vf, bnd, vol):
    """fvc::surfaceIntegrate of flux * face value: internal faces in face order
    (ivf[owner] += ssf; ivf[neighbour] -= ssf), then the patches, then / V. For one cell the
    faces where it is the neighbour have lower indices than the ones it owns."""
    nI = case.nInternalFaces
    vf2, corr2, bnd2 = (np.asarray(x).reshape(len(x), -1) for x in (vf, corr, bnd))
    P, N = case.owner[:nI], case.neighbour
    up = np.where(flux[:nI] > 0, P, N)
    ssf = flux[:nI, None] * (vf2[up] + corr2[:nI])
    div = np.zeros_like(vf2)
    np.subtract.at(div, N, ssf)
    np.add.at(div, P, ssf)
    val = bnd2.copy()
    for (ptype, start, size, nbr) in case.patches:
        if ptype == 2 and size:  # cyclic: upwind value + correction, neighbour cell is local
            f = np.arange(start, start + size)
            own = case.owner[f]
            (qs,) = [ps for i, (pt, ps, pz, pn) in enumerate(case.patches) if i == nbr]
            other = case.owner[qs + (f - start)]
            upv = np.where(flux[f, None] > 0, vf2[own], vf2[other])
            val[f - nI] = upv + corr2[f]
    np.add.at(div, case.owner[nI:], flux[nI:, None] * val)
    return div / vol[:, None]


@pytest.mark.parametrize("name,sides", [("box", (0,) * 6), ("cyclic", (2, 2, 0, 0, 0, 0))])
@pytest.mark.parametrize("scheme,bounded,nComp", [(FIT, 1, 3), (FIT, 0, 1), (FIT01, 1, 1), (FIT, 1, 6)])
def test_div_matches_surface_integrate(name, sides, scheme, bounded, nComp):
    case = wx.BoxCase((10, 8, 6), 2, sides=sides, useClassCache=False)
    plan = wx.Plan(case.desc)
    rng = np.random.default_rng(3)
    v3 = H.vector_field(case.C, seed=6)
    vf = np.ascontiguousarray(np.concatenate([v3, 0.5 * v3 + 0.2], axis=1)[:, :nComp])
    flux = H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.02)
    nB = case.nFaces - case.nInternalFaces
    bnd = rng.uniform(-1, 1, (nB, nComp))
    vol = rng.uniform(0.5, 1.5, case.nCells) * 1e-3
    got = np.asarray(plan.div(scheme, bounded, vf, flux, bnd, vol)).reshape(case.nCells, nComp)
    corr = np.asarray(plan.correction(scheme, bounded, vf, flux)).reshape(case.nFaces, nComp)
    np.testing.assert_array_equal(got, surface_integrate(case, flux, corr, vf, bnd, vol))
    ref = Oracle("c").correction(case, scheme, bounded, vf if nComp > 1 else vf[:, 0], flux)
    want = surface_integrate(case, flux, np.asarray(ref).reshape(case.nFaces, nComp), vf, bnd, vol)
    assert H.normwise_err(got, want) <= 1e-12
    plan.destroy()


def test_div_is_conservative_at_full_size():
    """Telescoping: sum_cells div * V = sum over boundary faces of flux * value (internal faces
    cancel), at the bench workload size, device buffers."""
    import torch
    case = wx.BoxCase((128, 128, 128), 2)
    plan = wx.Plan(case.desc)
    vf = H.vector_field(case.C, seed=8)
    flux = H.face_flux(case.Sf, U=(1.0, -1.0, 0.5))
    nI = case.nInternalFaces
    bnd = np.full((case.nFaces - nI, 3), 0.25)
    vol = np.full(case.nCells, 1.0 / 128 ** 3)
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()  # noqa: E731
    div = plan.div(FIT, 1, t(vf), t(flux), t(bnd), t(vol)).cpu().numpy()
    total = (div * vol[:, None]).sum(axis=0)
    want = (flux[nI:, None] * bnd).sum(axis=0)
    scale = np.abs(flux[nI:, None] * bnd).sum(axis=0)
    assert np.all(np.abs(total - want) <= 1e-9 * scale)
    plan.destroy()


@pytest.mark.parametrize("sides", [(0,) * 6, (2, 2, 0, 0, 0, 0)], ids=["walls", "cyclic_x"])
def test_div_host_buffers_pipelined_equals_device_buffers(sides):
    """Host buffers: the reconstruction runs slab-pipelined and the face field stays on the device; only
    [nCells][nComp] comes back. Same bits as the device-buffer call; the volumes and the flux of an
    earlier call can be reused (cellVolumes = NULL, faceFlux = NULL after wenoext_set_face_flux)."""
    import torch
    case = wx.BoxCase((48, 48, 40), 2, sides=sides)
    plan = wx.Plan(case.desc)
    try:
        rng = np.random.default_rng(4)
        vf = H.vector_field(case.C, seed=9)
        flux = H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.02)
        bnd = rng.uniform(-1, 1, (case.nFaces - case.nInternalFaces, 3))
        vol = rng.uniform(0.5, 1.5, case.nCells) * 1e-3
        t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()  # noqa: E731
        want = plan.div(FIT, 1, t(vf), t(flux), t(bnd), t(vol)).cpu().numpy()
        got = np.asarray(plan.div(FIT, 1, vf, flux, bnd, vol))
        np.testing.assert_array_equal(got, want)
        # second "time step": new field, same mesh volumes, flux uploaded once
        vf2 = H.vector_field(case.C, seed=10)
        want2 = plan.div(FIT, 1, t(vf2), t(flux), t(bnd), t(vol)).cpu().numpy()
        plan.set_face_flux(flux)
        got2 = np.asarray(plan.div(FIT, 1, vf2, None, bnd, None))
        np.testing.assert_array_equal(got2, want2)
        # and the face field of an ordinary call is unaffected by the skipped copies before it
        corr = np.asarray(plan.correction(FIT, 1, vf2, None))
        corr_d = plan.correction(FIT, 1, t(vf2), t(flux)).cpu().numpy()
        np.testing.assert_array_equal(corr, corr_d)
    finally:
        plan.destroy()


@pytest.mark.parametrize("slabs", ["2", "6", "8"])
def test_div_first_host_buffer_call_of_a_fresh_plan_with_many_slabs(monkeypatch, slabs):
    """Host-buffer wenoext_div forms the divergence slab by slab under the kernels (cells whose faces are final
    after slab k). On the FIRST call of a fresh plan nothing stale can stand in for data that is read too early
    (the flux on boundary faces, at the end of the flux field, once was)."""
    import torch
    monkeypatch.setenv("WENOEXT_E2E_SLABS", slabs)
    case = wx.BoxCase((48, 48, 40), 2)
    plan = wx.Plan(case.desc)
    try:
        rng = np.random.default_rng(4)
        vf = H.vector_field(case.C, seed=9)
        flux = H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.02)
        bnd = rng.uniform(-1, 1, (case.nFaces - case.nInternalFaces, 3))
        vol = rng.uniform(0.5, 1.5, case.nCells) * 1e-3
        got = np.asarray(plan.div(FIT, 1, vf, flux, bnd, vol))  # first call
        t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()  # noqa: E731
        want = plan.div(FIT, 1, t(vf), t(flux), t(bnd), t(vol)).cpu().numpy()
        np.testing.assert_array_equal(got, want)
    finally:
        plan.destroy()
