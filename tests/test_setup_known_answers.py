"""Known answers of the reference's own unit tests for the setup helpers.

The reference pins these pieces of the WENOBase setup by known-answer tests
(SURVEY.md section 4); the host restatement in wenoext_b200/csrc/setup/ is checked
against the same numbers through tests/native/libsetup_hooks.so:
  tests/src/geometryWENO-BasicFunc-Test.C:50-145   Jacobian inverse / transformPoint
  tests/src/geometryWENO-BasicFunc-Test.C:149-203  gaussQuad (Matlab value + regression value)
  tests/src/geometryWENO-BasicFunc-Test.C:206-478  initIntegrals / transformIntegral, orders 2-4
  tests/src/mathFunctionsWENO-Test.C:45-131        det, eigenvalues, inverse of 3x3 matrices
"""
import ctypes as C
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def hooks():
    L = C.CDLL(os.path.join(HERE, "native", "libsetup_hooks.so"))
    dp = C.POINTER(C.c_double)
    L.hook_det3.restype = C.c_double
    L.hook_det3.argtypes = [dp]
    L.hook_inv3.argtypes = [dp, dp]
    L.hook_eigen_magnitudes.argtypes = [dp, dp]
    L.hook_transform_point.argtypes = [dp, dp, dp, dp]
    L.hook_gauss_quad.restype = C.c_double
    L.hook_gauss_quad.argtypes = [C.c_int, C.c_int, C.c_int, dp, dp, dp, dp]
    L.hook_cell_integrals.restype = C.c_int
    L.hook_cell_integrals.argtypes = [C.POINTER(C.c_int32), dp, dp, C.c_int, C.c_int, dp, dp, dp, dp, dp]
    return L


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(C.POINTER(C.c_double))


def test_det_of_3x3(hooks):
    # mathFunctionsWENO-Test.C:47-57
    A, p = _d([[2, -5, 3], [0, 7, -2], [-1, 4, 1]])
    assert hooks.hook_det3(p) == 41


def test_inverse_of_3x3(hooks):
    # mathFunctionsWENO-Test.C:112-130
    A, p = _d([[0, -3, -2], [1, -4, -2], [-3, 4, 1]])
    out, po = _d(np.zeros((3, 3)))
    hooks.hook_inv3(p, po)
    np.testing.assert_allclose(out, [[4, -5, -2], [5, -6, -2], [-8, 9, 3]], rtol=0, atol=1e-12)


@pytest.mark.parametrize("A,eig", [
    # mathFunctionsWENO-Test.C:62-74 and :76-89 (the reference returns signed values; the
    # restatement returns the magnitudes its only caller, cond(), uses)
    ([[5, 2, 0], [2, 5, 0], [-3, 4, 6]], [7, 6, 3]),
    ([[1, 1, 0.1], [0, 2, 0], [1, 0, 0.1]], [2, 1.1, 0]),
])
def test_eigenvalues_of_3x3(hooks, A, eig):
    A, p = _d(A)
    out, po = _d(np.zeros(3))
    hooks.hook_eigen_magnitudes(p, po)
    np.testing.assert_allclose(sorted(out, reverse=True), eig, rtol=1e-12, atol=1e-12)


def test_jacobi_inverse_and_transform_point(hooks):
    # geometryWENO-BasicFunc-Test.C:61-143: tet (0,0,0),(1,0,0),(0,2,0),(0,0,3) -> J = diag(1,2,3);
    # transformPoint(JInv, (2,2,3), (1,0,0)) = (1,1,1)
    J, pj = _d(np.diag([1.0, 2.0, 3.0]))
    Jinv, pi = _d(np.zeros((3, 3)))
    hooks.hook_inv3(pj, pi)
    np.testing.assert_allclose(Jinv @ J, np.eye(3), atol=1e-15)
    xp, pxp = _d([2, 2, 3])
    x0, px0 = _d([1, 0, 0])
    out, po = _d(np.zeros(3))
    hooks.hook_transform_point(pi, pxp, px0, po)
    np.testing.assert_allclose(out, [1, 1, 1], rtol=1e-14)


def test_gauss_quad_known_answers(hooks):
    # geometryWENO-BasicFunc-Test.C:155-180 (Matlab value) and :184-200 (regression value);
    # Catch::Approx default tolerance is 1.2e-5 relative (epsilon = 100 * float eps)
    p0, pp0 = _d([0.333, 0.333, 1])
    v0, pv0 = _d([0, 0, 0])
    v1, pv1 = _d([0, 1, 0])
    v2, pv2 = _d([1, 0, 0])
    assert hooks.hook_gauss_quad(1, 1, 1, pp0, pv0, pv1, pv2) == pytest.approx(0.0277777, rel=1.2e-5)
    v0, pv0 = _d([0.05, 0.21, 0.08])
    v1, pv1 = _d([0.03, 1.5, 0.03])
    v2, pv2 = _d([1.87, 0.1, 0.01])
    assert hooks.hook_gauss_quad(1, 2, 4, pp0, pv0, pv1, pv2) == pytest.approx(0.0039776813, rel=1.2e-5)


@pytest.mark.parametrize("order", [2, 3, 4])
def test_volume_integrals_of_a_cuboid_cell(hooks, order):
    # geometryWENO-BasicFunc-Test.C:206-478: cell 33 of the 100 x 100 x 1 case
    # (tests/Cases/2DMesh-cyclic/system/blockMeshDict: [0,1]^2 x [0,0.1]). In the reference
    # frame the cell is a unit cube about its centre: moments 1, 1/12 (one exponent 2),
    # 1/144 (two exponents 2), 1/80 (one exponent 4), everything odd is 0 (+-1e-9);
    # |det JInv| = 1e5 = 1 / cell volume.
    n = (C.c_int32 * 3)(100, 100, 1)
    lo, plo = _d([0, 0, 0])
    hi, phi = _d([1, 1, 0.1])
    s = order + 1
    vol, pv = _d(np.zeros(s ** 3))
    tr, pt = _d(np.zeros(s ** 3))
    J, pJ = _d(np.zeros((3, 3)))
    rp, prp = _d(np.zeros(3))
    det, pdet = _d(np.zeros(1))
    assert hooks.hook_cell_integrals(n, plo, phi, 33, order, pv, pt, pJ, prp, pdet) == 0
    expect = np.zeros((s, s, s))
    for a in range(s):
        for b in range(s):
            for c in range(s):
                if a + b + c > order:
                    continue  # the reference only fills n+m+l <= polOrder
                if a % 2 or b % 2 or c % 2:
                    continue
                v = 1.0
                for e in (a, b, c):
                    v *= {0: 1.0, 2: 1.0 / 12.0, 4: 1.0 / 80.0}[e]
                expect[a, b, c] = v
    for got in (vol.reshape(s, s, s), tr.reshape(s, s, s)):
        for a in range(s):
            for b in range(s):
                for c in range(s):
                    if a + b + c <= order:
                        assert got[a, b, c] == pytest.approx(expect[a, b, c], rel=1.2e-5, abs=1e-9), (a, b, c)
    assert abs(np.linalg.det(J)) == pytest.approx(1e5, rel=1.2e-5)
