"""Generate the golden correction() vectors under tests/golden/.

The reference itself cannot run in this container (every hot-path translation unit
needs OpenFOAM) and its own tests pin no per-face value, so the
golden outputs are produced by the strongest available stand-in: the oracle
variant that performs the three matrix products with the reference's vendored
Blaze 3.8 (oracle/weno_oracle_blaze.cpp -> oracle/_ref/), on the stencil data of
the host setup restatement. Inputs follow SURVEY.md 8(c)/(d): the reference's
analytic test field, a Zalesak-like indicator, polynomial / constant fields,
seeded noise, and fluxes with mixed signs and exact zeros.

    python tests/golden/make_golden.py        (rewrites tests/golden/*.npz)

Each file holds, for one mesh case: cell values and face fluxes per input set,
and the oracle's correction for every (scheme, bounded) pair.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import wenoext_b200 as wx  # noqa: E402
from tests import helpers as H  # noqa: E402
from tests.oracle_binding import Oracle  # noqa: E402

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


def make_case(name):
    kw = dict(CASES[name])
    n, order = kw.pop("n"), kw.pop("order")
    return wx.BoxCase(n, order, useClassCache=False, **kw)


def unit_coords(case):
    lo, hi = case.C.min(axis=0), case.C.max(axis=0)
    return (case.C - lo) / (hi - lo + 1e-300)


def input_sets(case):
    """name -> (vf [nCells] or [nCells][3], flux [nFaces])"""
    X = unit_coords(case)
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


def main():
    out_dir = os.path.dirname(os.path.abspath(__file__))
    orc = Oracle("blaze")
    for name in CASES:
        case = make_case(name)
        data = {"nCells": np.int64(case.nCells), "nFaces": np.int64(case.nFaces),
                "nInternalFaces": np.int64(case.nInternalFaces)}
        for sname, (vf, flux) in input_sets(case).items():
            data[f"{sname}/vf"] = vf
            data[f"{sname}/flux"] = flux
            for scheme, bounded in SCHEMES:
                corr = orc.correction(case, scheme, bounded, vf, flux)
                data[f"{sname}/corr_s{scheme}b{bounded}"] = corr
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **data)
        print(name, case.nCells, "cells ->", os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
