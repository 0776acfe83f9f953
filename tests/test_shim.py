"""The OpenFOAM-side shim (shim/openfoam/*.C): compiled, linked and RUN against the OpenFOAM stand-in
oracle/of_stub (`make -C shim check`). The same harness that drives the reference's own classes for
the parity pin (oracle/ref_harness.cpp) here drives the shim's WENOUpwindFit<Type> /
WENOUpwindFit01<Type> through run-time selection, the (mesh, faceFlux, Istream) constructor and
WENODict; correction() stages through pinned buffers and calls libwenoext_b200.so. What a user of
`libs ("libWENOEXT.so")` would get, short of OpenFOAM itself."""
import os
import subprocess

import numpy as np
import pytest

from tests import helpers as H
from tests.golden.make_golden import CASES, SCHEMES, input_sets, make_case
from tests.oracle_binding import SHIM_SO, Oracle, ReferenceBuild, ShimBuild, reference_build_available

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-12


def test_shim_compiles_against_the_openfoam_standin():
    """`make -C shim check` (a no-op without /root/reference, where the prebuilt library is kept)"""
    subprocess.run(["make", "-C", os.path.join(ROOT, "shim"), "check"], check=True, stdout=subprocess.DEVNULL)
    assert os.path.exists(SHIM_SO)


def test_shim_registers_both_schemes_and_fails_loudly_without_a_gpu():
    import wenoext_b200 as wx
    case = make_case("hex6_o2")
    shim = ShimBuild(case)
    assert "shim" in shim.provenance
    vf, flux = input_sets(case)["sine_u111"]
    with pytest.raises(RuntimeError, match="Unknown discretisation scheme"):
        shim.correction("WENOUpwindFit02", 2, 1, vf, flux)
    if reference_build_available():
        # both harness libraries in one process keep their own scheme tables and WENOBase registries
        ref = ReferenceBuild(case)
        want = Oracle("c").correction(case, 0, 1, vf, flux)
        assert H.normwise_err(ref.correction("WENOUpwindFit", 2, 1, vf, flux), want) <= TOL
    if wx.lib().wenoext_device_count() > 0:
        pytest.skip("a GPU is present")
    for name in ("WENOUpwindFit", "WENOUpwindFit01"):
        with pytest.raises(RuntimeError, match="wenoext_b200"):  # FatalError of the shim: no CPU fallback
            shim.correction(name, 2, 1, vf, flux)


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_shim_on_gpu_equals_reference_classes(name):
    """drop-in: same scheme names, same tokens, same fields in and out - the reference's classes on the
    CPU against the shim's classes on the GPU (goldens where the reference build is absent)"""
    case = make_case(name)
    order = CASES[name]["order"]
    shim = ShimBuild(case)
    ref = ReferenceBuild(case) if reference_build_available() else None
    g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    for sname, (vf, flux) in input_sets(case).items():
        for scheme, bounded in SCHEMES:
            got, w = shim.correction(scheme, order, bounded, vf, flux, want_weights=True)
            want = ref.correction(scheme, order, bounded, vf, flux) if ref else g[f"{sname}/corr_s{scheme}b{bounded}"]
            np.testing.assert_array_equal(w, (flux >= 0).astype(float))  # weights() = pos(faceFlux)
            if not np.any(want):
                assert not np.any(got)
            else:
                assert H.normwise_err(got, want) <= TOL


@pytest.mark.gpu
def test_shim_wenodict_and_field_types():
    """WENODict p / dm / epsilon reach the plan (WENOCoeff.C:87-89); all five registered field types run
    (tensor = three component batches on the GPU)"""
    case = make_case("tets3_o2")
    shim = ShimBuild(case)
    orc = Oracle("c")
    rng = np.random.default_rng(5)
    flux = H.rotating_flux(case)
    try:
        shim.set_dict(p=2.7, dm=100.0, epsilon=1e-30)
        vf = H.vector_field(case.C, seed=8)
        for scheme, bounded in SCHEMES:
            got = shim.correction(scheme, 2, bounded, vf, flux)
            want = orc.correction(case, scheme, bounded, vf, flux, p=2.7, dm=100.0, eps=1e-30)
            assert H.normwise_err(got, want) <= TOL
    finally:
        shim.set_dict()
    for typeName, nComp in (("scalar", 1), ("sphericalTensor", 1), ("symmTensor", 6), ("tensor", 9)):
        vf = np.stack([H.sin_field(case.C) * (1 + 0.3 * c) + 0.01 * rng.uniform(-1, 1, case.nCells)
                       for c in range(nComp)], axis=1)
        for scheme, bounded in SCHEMES:
            got = shim.correction(scheme, 2, bounded, vf, flux, typeName=typeName)
            for c in range(nComp):
                want = orc.correction(case, scheme, bounded, vf[:, c].copy(), flux)
                assert H.normwise_err(got[:, c], want) <= TOL
