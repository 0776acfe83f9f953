"""Generate the golden correction() vectors under tests/golden/.

PROVENANCE: the outputs are produced by the REFERENCE'S OWN CODE. oracle/Makefile compiles the
unmodified translation units libWENOEXT/WENOUpwindFit/makeWENOUpwindFit.C and
libWENOEXT/WENOUpwindFit01/makeWENOUpwindFit01.C (hence WENOUpwindFit.C, WENOUpwindFit01.C,
WENOCoeff.C, List3D.H and the vendored Blaze 3.8) where they lie under /root/reference against the
OpenFOAM stand-in oracle/of_stub/, and tests/oracle_binding.ReferenceBuild drives them through
run-time selection ("WENOUpwindFit 2 1"), weights() and correction(vf) - one thread per MPI rank for
the decomposed cases, so collectData and swapData run as written. The reference's own tests pin no
per-face value, so these vectors are what pins the C oracle and the CUDA path.
The stencil data (the WENOBase content) comes from the host setup stand-in, which is what all three
(reference build, oracle, GPU plan) consume in the same flattened form.

    python tests/golden/make_golden.py        (rewrites tests/golden/*.npz; needs /root/reference)

Inputs follow SURVEY.md 8(c)/(d): the reference's analytic test field, a Zalesak-like indicator,
polynomial / constant fields, seeded noise, fluxes with mixed signs and exact zeros.
Each file holds, for one mesh case: cell values and face fluxes per input set (per rank for the
decomposed cases, keys "<set>/r<rank>/..."), and the correction for every (scheme, bounded) pair.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import wenoext_b200 as wx  # noqa: E402
from tests import helpers as H  # noqa: E402

FIT, FIT01 = 0, 1
SCHEMES = [(FIT, 0), (FIT, 1), (FIT01, 1)]

CASES = {
    # name: BoxCase arguments (kept in sync with tests/test_golden.py)
    "hex6_o2": dict(n=(6, 6, 6), order=2),
    "cavity20_o2": dict(n=(20, 20, 1), order=2, hi=(0.1, 0.1, 0.01), sides=(0, 0, 0, 0, 3, 3)),
    "tets3_o2": dict(n=(3, 3, 3), order=2, tetSplit=6),
    "cyclic2d12_o2": dict(n=(12, 12, 1), order=2, sides=(2, 2, 2, 2, 3, 3)),
    "hex6_o3": dict(n=(6, 6, 6), order=3),
}
# decomposed cases: `decomposePar simple` blocks, one MPI rank each
PAR_CASES = {
    "hex8x6x6_o2_np2": dict(n=(8, 6, 6), order=2, nProc=(2, 1, 1)),
    "hex6_o2_np8": dict(n=(6, 6, 6), order=2, nProc=(2, 2, 2)),
    "cyclicy10x6x6_o2_np2": dict(n=(10, 6, 6), order=2, nProc=(2, 1, 1), sides=(0, 0, 2, 2, 0, 0)),
}


def make_case(name):
    kw = dict(CASES[name])
    n, order = kw.pop("n"), kw.pop("order")
    return wx.BoxCase(n, order, useClassCache=False, **kw)


def make_par_cases(name):
    kw = dict(PAR_CASES[name])
    n, order, nProc = kw.pop("n"), kw.pop("order"), kw.pop("nProc")
    R = int(np.prod(nProc))
    cases = [wx.BoxCase(n, order, nProc=nProc, myProc=r, useClassCache=False, **kw) for r in range(R)]
    wx.exchange_halo_lists(cases)
    return cases


def unit_coords(case, lo=None, hi=None):
    lo = case.C.min(axis=0) if lo is None else lo
    hi = case.C.max(axis=0) if hi is None else hi
    return (case.C - lo) / (hi - lo + 1e-300)


def input_sets(case, X=None):
    """name -> (vf [nCells] or [nCells][3], flux [nFaces])"""
    X = unit_coords(case) if X is None else X
    sets = {}
    # (i) reference test field, U = (1,1,1)                (WENOUpwindFit-Test.C:87-94,:199,:220)
    sets["sine_u111"] = (H.sin_field(X), H.face_flux(case.Sf))
    # (ii) Zalesak-like indicator with solid-body rotation (WENOUpwindFit-AdvectionTest.C:90-103,:173)
    sets["zalesak_rot"] = (H.zalesak_like(X), H.rotating_flux(case))
    # (iii) polynomial of degree 2 (reproduced exactly at order >= 2)
    sets["quadratic"] = (1.0 + X[:, 0] - 2 * X[:, 1] + 0.5 * X[:, 0] * X[:, 1] + X[:, 2] ** 2
                         if case.C[:, 2].max() > case.C[:, 2].min() else
                         1.0 + X[:, 0] - 2 * X[:, 1] + 0.5 * X[:, 0] * X[:, 1],
                         H.face_flux(case.Sf, U=(1.0, -1.0, 0.5)))
    # (iv) constant field: exact zero correction
    sets["constant"] = (np.full(case.nCells, 0.75), H.face_flux(case.Sf))
    # (v) smooth + step + seeded noise, vector valued
    v = H.vector_field(X, seed=2)
    v[:, 1] += H.step_field(X)
    sets["vector_step_noise"] = (v, H.face_flux(case.Sf, U=(1.0, -1.0, 0.5)))
    # (vi) mixed-sign fluxes with 5 % exact zeros
    sets["vector_zero_flux"] = (H.vector_field(X, seed=3),
                                H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.05, seed=7))
    return sets


def par_input_sets(cases):
    """name -> ([vf per rank], [flux per rank]); the fields are functions of the global coordinates
    (unit box) and the processor-patch fluxes of the two sides are equal and opposite by
    construction (S_f . U with opposite S_f; the zero pattern is not used here because the two
    sides would draw different random faces)."""
    per_rank = [input_sets(c, X=c.C.copy()) for c in cases]
    names = [k for k in per_rank[0] if k != "vector_zero_flux"]
    return {k: ([s[k][0] for s in per_rank], [s[k][1] for s in per_rank]) for k in names}


def main():
    from tests.oracle_binding import ReferenceBuild
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name in CASES:
        case = make_case(name)
        order = CASES[name]["order"]
        ref = ReferenceBuild(case)
        data = {"nCells": np.int64(case.nCells), "nFaces": np.int64(case.nFaces),
                "nInternalFaces": np.int64(case.nInternalFaces),
                "provenance": np.array(ref.provenance)}
        for sname, (vf, flux) in input_sets(case).items():
            data[f"{sname}/vf"] = vf
            data[f"{sname}/flux"] = flux
            for scheme, bounded in SCHEMES:
                data[f"{sname}/corr_s{scheme}b{bounded}"] = ref.correction(scheme, order, bounded, vf, flux)
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **data)
        print(name, case.nCells, "cells ->", os.path.getsize(path) // 1024, "KiB")
    for name in PAR_CASES:
        cases = make_par_cases(name)
        order = PAR_CASES[name]["order"]
        ref = ReferenceBuild(cases)
        data = {"nRanks": np.int64(len(cases)), "provenance": np.array(ref.provenance)}
        for sname, (vfs, fluxes) in par_input_sets(cases).items():
            for r in range(len(cases)):
                data[f"{sname}/r{r}/vf"] = vfs[r]
                data[f"{sname}/r{r}/flux"] = fluxes[r]
            for scheme, bounded in SCHEMES:
                corr = ref.correction(scheme, order, bounded, vfs, fluxes)
                for r in range(len(cases)):
                    data[f"{sname}/r{r}/corr_s{scheme}b{bounded}"] = corr[r]
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **data)
        print(name, sum(c.nCells for c in cases), "cells ->", os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
