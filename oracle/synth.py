"""Seeded synthetic "surrogate output" fields (TEST INFRASTRUCTURE -- see oracle/__init__.py).

SURVEY.md section 8(d): smooth multi-mode fields plus 1 % noise, so PDE residuals are O(1e-2)
of the field scale like a trained surrogate's; density/pressure-like fields strictly
positive.  NumPy's PCG64 generator is platform independent, so the same seed gives the same
fields in the build container and on the GPU box.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def smooth_field(rng, shape, positive=False, noise=0.01, modes=4):
    """``shape`` = (BS, Nt, Nx[, Ny]).  u = sum_m a_m sin(2pi(m x + phi_m)) cos(2pi(m y + psi_m))
    exp(-lambda_m t) + noise * N(0,1);  positive=True maps it to 1 + 0.1 u."""
    bs, nt = shape[0], shape[1]
    sp = shape[2:]
    t = np.linspace(0.0, 1.0, nt, dtype=F32).reshape((1, nt) + (1,) * len(sp))
    x = np.linspace(0.0, 1.0, sp[0], endpoint=False, dtype=F32).reshape(
        (1, 1, sp[0]) + (1,) * (len(sp) - 1))
    y = None
    if len(sp) == 2:
        y = np.linspace(0.0, 1.0, sp[1], endpoint=False, dtype=F32).reshape(1, 1, 1, sp[1])
    out = np.zeros(shape, dtype=F32)
    bshape = (bs,) + (1,) * (len(shape) - 1)
    for m in range(1, modes + 1):
        a = rng.uniform(0.5, 1.0, bshape).astype(F32)
        phi = rng.uniform(0.0, 1.0, bshape).astype(F32)
        psi = rng.uniform(0.0, 1.0, bshape).astype(F32)
        lam = rng.uniform(0.1, 1.0, bshape).astype(F32)
        term = a * np.sin(F32(2 * np.pi) * (m * x + phi)) * np.exp(-lam * t)
        if y is not None:
            term = term * np.cos(F32(2 * np.pi) * (m * y + psi))
        out += term.astype(F32)
    out += F32(noise) * rng.standard_normal(shape, dtype=F32)
    if positive:
        out = (F32(1.0) + F32(0.1) * out / F32(modes)).astype(F32)
    return out


def fields(seed, n_fields, shape, positive=()):
    """[BS, F, Nt, Nx(, Ny)] stack of independent smooth fields; indices in ``positive`` are
    strictly positive (MHD rho, p)."""
    rng = np.random.default_rng(seed)
    fs = [smooth_field(rng, shape, positive=(j in positive)) for j in range(n_fields)]
    return np.stack(fs, axis=1)


def normal_fields(seed, n_fields, shape):
    """Plain N(0,1) worst case for parity."""
    rng = np.random.default_rng(seed)
    return rng.standard_normal((shape[0], n_fields) + tuple(shape[1:]), dtype=F32)


# Physical constants of the BASELINE configs.
BURGERS = dict(dx=2.0 / 1000 * 5, dt=1.25 / 500 * 10, nu=0.002)
WAVE = dict(c2=0.25)                     # (c*dt/dx)^2
NS = dict(dx=0.01, dy=0.01, dt=0.01, nu=0.001)
MHD = dict(dx=1.0 / 128, dy=1.0 / 128, dt=0.01, gamma=5.0 / 3.0)
ADVECTION = dict(nx=200, nt=100, x_min=0.0, x_max=2.0, t_end=0.5, v=1.0)
