"""The oracle against the reference's OWN code (the parity pin).

oracle/_ref/libweno_ref_<isa>.so holds the reference's unmodified hot-path translation units
(makeWENOUpwindFit.C, makeWENOUpwindFit01.C -> WENOUpwindFit.C, WENOUpwindFit01.C, WENOCoeff.C,
List3D.H, Blaze 3.8) compiled against the OpenFOAM stand-in oracle/of_stub; it is built in this
container (where /root/reference exists) and travels prebuilt to the GPU box. These tests run the
reference through the interface OpenFOAM itself uses - run-time selection by name, the
(mesh, faceFlux, Istream) constructor, WENODict, weights(), correction() - serial and with one
thread per MPI rank, and compare the C oracle with it on fresh inputs (not the committed goldens)."""
import numpy as np
import pytest

from tests import helpers as H
from tests.golden.make_golden import (CASES, PAR_CASES, SCHEMES, input_sets, make_case,
                                      make_par_cases, par_input_sets)
from tests.oracle_binding import (Oracle, ReferenceBuild, correction_multirank,
                                  reference_build_available)

pytestmark = pytest.mark.skipif(not reference_build_available(),
                                reason="oracle/_ref/libweno_ref_*.so not built (needs /root/reference)")
TOL = 1e-12


def _close(got, ref):
    if not np.any(ref):
        assert not np.any(got)
    else:
        assert H.normwise_err(got, ref) <= TOL


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_equals_reference_build(name):
    case = make_case(name)
    order = CASES[name]["order"]
    ref = ReferenceBuild(case)
    assert "compiled unmodified" in ref.provenance
    for variant in ("c", "blaze"):
        orc = Oracle(variant)
        for sname, (vf, flux) in input_sets(case).items():
            vf = vf + 0.01 * np.cos(7.0 * np.arange(vf.size)).reshape(vf.shape)  # not the golden inputs
            for scheme, bounded in SCHEMES:
                want = ref.correction(scheme, order, bounded, vf, flux)
                _close(orc.correction(case, scheme, bounded, vf, flux), want)


@pytest.mark.parametrize("name", list(PAR_CASES))
def test_oracle_equals_reference_build_decomposed(name):
    """collectData (WENOCoeff.C:224-296) and swapData (WENOUpwindFit.C:326-385) as the reference
    wrote them, over the threaded Pstream stand-in, against the oracle's emulation."""
    cases = make_par_cases(name)
    order = PAR_CASES[name]["order"]
    ref = ReferenceBuild(cases)
    orc = Oracle("c")
    for sname, (vfs, fls) in par_input_sets(cases).items():
        for scheme, bounded in SCHEMES:
            want = ref.correction(scheme, order, bounded, vfs, fls)
            got = correction_multirank(orc, cases, scheme, bounded, vfs, fls)
            for r in range(len(cases)):
                _close(np.asarray(got[r]).reshape(want[r].shape), want[r])


def test_weights_and_selection():
    """weights() = pos(faceFlux) (WENOUpwindFit.H:223-229); unknown scheme names are refused by
    the run-time selection table."""
    case = make_case("hex6_o2")
    ref = ReferenceBuild(case)
    vf, flux = input_sets(case)["vector_zero_flux"]
    _, w = ref.correction("WENOUpwindFit", 2, 1, vf, flux, want_weights=True)
    np.testing.assert_array_equal(w, (flux >= 0).astype(float))
    with pytest.raises(RuntimeError, match="Unknown discretisation scheme"):
        ref.correction("WENOUpwindFit02", 2, 1, vf, flux)


@pytest.mark.parametrize("p,dm,eps", [(2.7, 100.0, 1e-30), (0.0, 1000.0, 1e-40), (6.0, 10.0, 1e-6)])
def test_wenodict_parameters(p, dm, eps):
    """WENODict p / dm / epsilon (WENOCoeff.C:87-89): p is used as an unsigned exponent
    (WENOCoeff.H:149-157), so 2.7 acts as 2 and 0 makes every gamma equal."""
    case = make_case("hex6_o2")
    ref = ReferenceBuild(case)
    ref.set_dict(p=p, dm=dm, epsilon=eps)
    vf, flux = input_sets(case)["vector_step_noise"]
    orc = Oracle("c")
    try:
        for scheme, bounded in SCHEMES:
            want = ref.correction(scheme, 2, bounded, vf, flux)
            got = orc.correction(case, scheme, bounded, vf, flux, p=p, dm=dm, eps=eps)
            _close(got, want)
            base = orc.correction(case, scheme, bounded, vf, flux)
            assert H.normwise_err(base, want) > 1e-9  # the parameters really took effect
    finally:
        ref.set_dict()  # back to the reference's defaults for the tests that follow


@pytest.mark.parametrize("typeName,nComp", [("sphericalTensor", 1), ("symmTensor", 6), ("tensor", 9)])
def test_other_field_types(typeName, nComp):
    """The registration macro instantiates five field types (makeWENOUpwindFit.C:37); components are
    independent, so each equals the scalar run of that component."""
    case = make_case("tets3_o2")
    ref = ReferenceBuild(case)
    X = case.C
    rng = np.random.default_rng(5)
    vf = np.stack([H.sin_field(X) * (1 + 0.3 * c) + 0.01 * rng.uniform(-1, 1, case.nCells)
                   for c in range(nComp)], axis=1)
    flux = H.rotating_flux(case)
    orc = Oracle("c")
    for scheme, bounded in SCHEMES:
        want = ref.correction(scheme, 2, bounded, vf, flux, typeName=typeName)
        for c in range(nComp):
            got = orc.correction(case, scheme, bounded, vf[:, c].copy(), flux)
            _close(got, want[:, c])
