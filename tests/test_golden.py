"""Committed golden vectors (tests/golden/*.npz, made by tests/golden/make_golden.py).

CPU: the plain-C oracle restatement reproduces the goldens (which were produced with
the reference's vendored Blaze arithmetic) to 1e-12; GPU: the CUDA path, through the
C ABI, reproduces them to the same tolerance. Neither reads /root/reference."""
import os

import numpy as np
import pytest

from tests import helpers as H
from tests.golden.make_golden import CASES, SCHEMES, make_case
from tests.oracle_binding import Oracle

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-12
_cases = {}


def _load(name):
    if name not in _cases:
        _cases[name] = (make_case(name), np.load(os.path.join(HERE, "golden", name + ".npz")))
    return _cases[name]


def _sets(g):
    return sorted({k.split("/")[0] for k in g.files if "/" in k})


def _check(got, ref):
    """SURVEY.md 8(c) metric; an all-zero golden must be reproduced as exact zeros."""
    if not np.any(ref):
        assert not np.any(got)
        return
    assert H.normwise_err(got, ref) <= TOL


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_reproduces_golden(name):
    case, g = _load(name)
    assert int(g["nCells"]) == case.nCells and int(g["nFaces"]) == case.nFaces
    orc = Oracle("c")
    for s in _sets(g):
        for scheme, bounded in SCHEMES:
            got = orc.correction(case, scheme, bounded, g[f"{s}/vf"], g[f"{s}/flux"])
            _check(got, g[f"{s}/corr_s{scheme}b{bounded}"])


@pytest.mark.parametrize("name", list(CASES))
def test_golden_properties(name):
    """Domain properties the goldens themselves must satisfy: a constant field gives an
    exactly zero correction; the 01 scheme keeps face values of a 0/1 field in [0, 1]."""
    case, g = _load(name)
    for scheme, bounded in SCHEMES:
        assert not np.any(g[f"constant/corr_s{scheme}b{bounded}"])
    vf, flux = g["zalesak_rot/vf"], g["zalesak_rot/flux"]
    corr = g["zalesak_rot/corr_s1b1"]
    nInt = case.nInternalFaces
    up = np.where(flux[:nInt] > 0, case.owner[:nInt], case.neighbour)
    face = vf[up] + corr[:nInt].reshape(nInt)
    assert face.min() >= -1e-12 and face.max() <= 1 + 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_cuda_reproduces_golden(name):
    import wenoext_b200 as wx
    case, g = _load(name)
    plan = wx.WENOBase.instance(case)
    for s in _sets(g):
        for scheme, bounded in SCHEMES:
            got = plan.correction(scheme, bounded, g[f"{s}/vf"], g[f"{s}/flux"])
            ref = g[f"{s}/corr_s{scheme}b{bounded}"]
            _check(np.asarray(got).reshape(ref.shape), ref)
