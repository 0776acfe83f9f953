"""Full-size checks at the bench workload (BASELINE.json config 3: 128^3 hex box, order 2,
vector field): size-independent properties of the scheme plus the oracle on the whole mesh.
The setup uses the translation-invariant fast path (tests/test_box_fast.py ties it to the
generic per-cell setup)."""
import os

import numpy as np
import pytest

import wenoext_b200 as wx
from tests import helpers as H
from tests.oracle_binding import Oracle, correction_threaded

pytestmark = pytest.mark.gpu
FIT, FIT01 = 0, 1
N = 128


@pytest.fixture(scope="module")
def big():
    case = wx.BoxCase((N, N, N), 2)
    plan = wx.Plan(case.desc)
    yield case, plan
    plan.destroy()


def _flux(case):
    return H.face_flux(case.Sf, U=(1.0, -1.0, 0.5))


def test_constant_field_gives_exact_zero(big):
    case, plan = big
    vf = np.full((case.nCells, 3), 0.375)
    for scheme, bounded in ((FIT, 0), (FIT, 1), (FIT01, 1)):
        corr = plan.correction(scheme, bounded, vf, _flux(case))
        assert not np.any(corr)


def test_linear_field_is_reproduced(big):
    """Every stencil reproduces a degree-1 polynomial exactly, so whatever the nonlinear
    weights are: correction = face average - cell average = grad . (x_f - x_c)."""
    case, plan = big
    g = np.array([0.7, -1.3, 0.4])
    vf = case.C @ g + 2.0
    flux = _flux(case)
    corr = plan.correction(FIT, 0, vf, flux)
    nI = case.nInternalFaces
    P, Nb = case.owner[:nI], case.neighbour
    fc = 0.5 * (case.C[P] + case.C[Nb])
    up = np.where(flux[:nI] > 0, P, Nb)
    exact = np.einsum("ij,j->i", fc - case.C[up], g)
    live = ~case.emptyCells[up]  # Cell::empty cells give zero correction (WENOCoeff.C:333)
    assert live.sum() > 0.99 * nI
    err = np.abs(corr[:nI] - exact)
    assert err[live].max() <= 1e-10 * np.abs(exact).max(), (err[live].max(), int(np.argmax(err * live)))
    assert not np.any(corr[:nI][~live])


def test_quadratic_field_is_independent_of_the_weights(big):
    """All stencils of an order-2 scheme fit a quadratic field with the same polynomial,
    so the weighted sum does not depend on (p, dm). (The face value itself is not the
    exact face average on quadrilateral faces: the reference subtracts the cell mean
    with the last triangle's area only, geometryWENO.C:853-869.)"""
    case, plan = big
    x, y, z = case.C[:, 0], case.C[:, 1], case.C[:, 2]
    vf = 1 + x - 2 * y + 0.5 * z + x * x - y * y + 0.3 * z * z + 0.7 * x * y
    flux = _flux(case)
    a = plan.correction(FIT, 0, vf, flux)
    plan.set_params(1.0, 1.0, 1e-40)
    try:
        b = plan.correction(FIT, 0, vf, flux)
    finally:
        plan.set_params(4.0, 1000.0, 1e-40)
    assert H.normwise_err(a, b) <= 1e-9


def test_interpolate_is_upwind_value_plus_correction(big):
    case, plan = big
    X = case.C
    vf = H.vector_field(X, seed=11)
    flux = _flux(case)
    nI = case.nInternalFaces
    fv = np.empty((nI, 3))
    corr = plan.correction(FIT, 1, vf, flux, faceValue=fv)
    up = np.where(flux[:nI] > 0, case.owner[:nI], case.neighbour)
    np.testing.assert_array_equal(fv, vf[up] + corr[:nI])


def test_whole_mesh_against_oracle(big):
    """1e-12 relative on all 6.2 M faces, bounded and unbounded."""
    case, plan = big
    vf = H.vector_field(case.C, seed=3)
    vf[:, 1] += H.step_field(case.C)
    flux = H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.01)
    orc = Oracle("c")
    T = max(1, (os.cpu_count() or 2) - 1)
    for bounded in (0, 1):
        ref = correction_threaded(orc, case, bounded, vf, flux, T)
        got = plan.correction(FIT, bounded, vf, flux)
        assert H.normwise_err(got, ref) <= 1e-12


def test_fit01_keeps_a_01_field_in_bounds(big):
    case, plan = big
    vf = H.zalesak_like(case.C)
    flux = H.rotating_flux(case)
    nI = case.nInternalFaces
    corr = plan.correction(FIT01, 1, vf, flux)
    up = np.where(flux[:nI] > 0, case.owner[:nI], case.neighbour)
    face = vf[up] + corr[:nI]
    assert face.min() >= -1e-12 and face.max() <= 1 + 1e-12
