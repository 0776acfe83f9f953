"""The 2-D bulk-copy path of the cell kernel (one strided box per part for the four cells of a
pass, plan.cu: buildTmaMaps) against the per-lane 1-D copies and the oracle. By default the
path is only taken when padding the records to a uniform stride is nearly free (large
meshes); WENOEXT_TMA2D=2 forces it, so that small meshes with many short boundary records mix
fast and general passes."""
import os

import numpy as np
import pytest

import wenoext_b200 as wx
from tests import helpers as H
from tests.oracle_binding import Oracle

pytestmark = pytest.mark.gpu


def _run(case, mode, scheme, bounded, vf, phi):
    old = os.environ.get("WENOEXT_TMA2D")
    os.environ["WENOEXT_TMA2D"] = mode
    try:
        plan = wx.Plan(case.desc)
        out = plan.correction(scheme, bounded, vf, phi)
        plan.destroy()
        return out
    finally:
        if old is None:
            os.environ.pop("WENOEXT_TMA2D", None)
        else:
            os.environ["WENOEXT_TMA2D"] = old


@pytest.mark.parametrize("kw", [dict(n=(13, 12, 11), order=2), dict(n=(20, 20, 20), order=2, sides=(2, 2, 0, 0, 0, 0)),
                                dict(n=(24, 24, 1), order=2, hi=(0.1, 0.1, 0.01), sides=(0, 0, 0, 0, 3, 3))])
@pytest.mark.parametrize("ncomp", [1, 3])
def test_forced_2d_copies_equal_1d_copies_and_oracle(kw, ncomp):
    case = wx.BoxCase(**kw)
    Cn = case.C / case.C.max(axis=0)
    vf = np.ascontiguousarray(H.vector_field(Cn)[:, :ncomp]) if ncomp > 1 else np.ascontiguousarray(H.vector_field(Cn)[:, 0])
    phi = H.rotating_flux(case)
    for scheme, bounded in ((0, 1), (1, 1)):
        a = _run(case, "2", scheme, bounded, vf, phi)
        b = _run(case, "0", scheme, bounded, vf, phi)
        assert np.array_equal(a, b)  # same arithmetic, only the copies differ
        ref = Oracle("c").correction(case, scheme, bool(bounded), vf, phi)
        assert H.normwise_err(a, ref) < 1e-12
