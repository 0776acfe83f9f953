"""Committed golden vectors (tests/golden/*.npz, made by tests/golden/make_golden.py).

The goldens were produced by the REFERENCE'S OWN CODE: its unmodified hot-path translation
units compiled against the OpenFOAM stand-in oracle/of_stub (see make_golden.py). CPU: the
plain-C oracle restatement reproduces them to 1e-12, serial and decomposed; GPU: the CUDA
path, through the C ABI, reproduces them to the same tolerance (decomposed cases as in-process
ranks on one GPU). Neither reads /root/reference."""
import os

import numpy as np
import pytest

from tests import helpers as H
from tests.golden.make_golden import CASES, PAR_CASES, SCHEMES, make_case, make_par_cases
from tests.oracle_binding import Oracle, correction_multirank

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-12
_cases = {}


def _load(name):
    if name not in _cases:
        _cases[name] = (make_case(name), np.load(os.path.join(HERE, "golden", name + ".npz")))
    return _cases[name]


def _load_par(name):
    if name not in _cases:
        _cases[name] = (make_par_cases(name), np.load(os.path.join(HERE, "golden", name + ".npz")))
    return _cases[name]


def _sets(g):
    return sorted({k.split("/")[0] for k in g.files if "/" in k})


def test_goldens_come_from_the_reference_build():
    for name in list(CASES) + list(PAR_CASES):
        g = np.load(os.path.join(HERE, "golden", name + ".npz"))
        prov = str(g["provenance"])
        assert "makeWENOUpwindFit.C" in prov and "compiled unmodified" in prov


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


@pytest.mark.parametrize("name", list(PAR_CASES))
def test_oracle_reproduces_decomposed_golden(name):
    """The oracle's emulation of the two exchange steps (collectData, swapData) against the
    reference's own MPI-style run (one thread per rank over the Pstream stand-in)."""
    cases, g = _load_par(name)
    R = int(g["nRanks"])
    assert R == len(cases)
    orc = Oracle("c")
    for s in _sets(g):
        vfs = [g[f"{s}/r{r}/vf"] for r in range(R)]
        fls = [g[f"{s}/r{r}/flux"] for r in range(R)]
        for scheme, bounded in SCHEMES:
            outs = correction_multirank(orc, cases, scheme, bounded, vfs, fls)
            for r in range(R):
                ref = g[f"{s}/r{r}/corr_s{scheme}b{bounded}"]
                _check(np.asarray(outs[r]).reshape(ref.shape), ref)


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


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(PAR_CASES))
def test_cuda_reproduces_decomposed_golden(name):
    """All ranks of the decomposed case as plans on one GPU (halo exchange, processor patches and
    the patch limiter through the library's own phases) against the reference's MPI-style run."""
    import torch

    import wenoext_b200 as wx
    cases, g = _load_par(name)
    R = len(cases)
    plans = [wx.Plan(c.desc) for c in cases]
    try:
        wx.attach_local(plans)
        for s in _sets(g):
            vfs = [torch.from_numpy(np.ascontiguousarray(g[f"{s}/r{r}/vf"])).cuda() for r in range(R)]
            fls = [torch.from_numpy(np.ascontiguousarray(g[f"{s}/r{r}/flux"])).cuda() for r in range(R)]
            for scheme, bounded in SCHEMES:
                outs = wx.correction_local(plans, scheme, bounded, vfs, fls)
                torch.cuda.synchronize()
                for r in range(R):
                    ref = g[f"{s}/r{r}/corr_s{scheme}b{bounded}"]
                    _check(outs[r].cpu().numpy().reshape(ref.shape), ref)
    finally:
        for p in plans:
            p.destroy()
