"""Synthetic fields and comparison metric shared by the tests."""
import numpy as np


def sin_field(C):
    """psi = sum_i sin(pi x_i): the reference's own test field
    (tests/src/WENOUpwindFit-Test.C:87-94 of the reference)."""
    return np.sin(np.pi * C).sum(axis=1)


def vector_field(C, seed=1, noise=1e-3):
    psi = sin_field(C)
    rng = np.random.default_rng(seed)
    v = np.stack([psi, psi * psi, np.cos(np.pi * C).sum(axis=1)], axis=1)
    return v + noise * rng.uniform(-1, 1, v.shape)


def step_field(C, x0=0.5):
    return np.where(C[:, 0] < x0, 1.0, 0.0) + 0.1 * sin_field(C)


def zalesak_like(C):
    """0/1 indicator of a slotted disk (tests/src/WENOUpwindFit-AdvectionTest.C:90-103)."""
    x, y = C[:, 0] - 0.5, C[:, 1] - 0.75
    r = np.sqrt(x * x + y * y)
    inside = (r < 0.15) & ~((np.abs(x) < 0.025) & (C[:, 1] < 0.85))
    return inside.astype(np.float64)


def face_flux(Sf, U=(1.0, 1.0, 1.0), zero_frac=0.0, seed=7):
    """phi_f = S_f . U (WENOUpwindFit-Test.C:199,:220); optionally force a fraction
    of the faces to exactly zero flux to exercise the third branch of
    WENOUpwindFit.C:258-281."""
    phi = Sf @ np.asarray(U, dtype=np.float64)
    if zero_frac > 0:
        rng = np.random.default_rng(seed)
        phi[rng.uniform(size=phi.shape) < zero_frac] = 0.0
    return phi


def rotating_flux(case):
    """solid-body rotation about (0.5,0.5): mixed-sign fluxes."""
    fc = 0.5 * (case.C[case.owner[:case.nInternalFaces]] + case.C[case.neighbour])
    # boundary faces: use owner centre (good enough for a sign pattern)
    fcb = case.C[case.owner[case.nInternalFaces:]]
    fc = np.concatenate([fc, fcb])
    U = np.stack([-(fc[:, 1] - 0.5), fc[:, 0] - 0.5, 0.3 * np.ones(len(fc))], axis=1)
    return np.einsum("ij,ij->i", case.Sf, U)


def normwise_err(a, b):
    """Per component max|a-b| / max|b| (SURVEY.md 8c parity metric)."""
    a = np.asarray(a, dtype=np.float64).reshape(len(a), -1)
    b = np.asarray(b, dtype=np.float64).reshape(len(b), -1)
    scale = np.maximum(np.abs(b).max(axis=0), 1e-300)
    return float((np.abs(a - b).max(axis=0) / scale).max())
