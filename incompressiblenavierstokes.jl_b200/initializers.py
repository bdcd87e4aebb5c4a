"""Initial conditions: src/initializers.jl.  Host NumPy sampling (runs once) followed by the
DEVICE projection (`insb200_project`), so every initial field is discretely divergence free with
respect to the same operators the stepper uses."""
from __future__ import annotations

import numpy as np
import scipy.fft as sfft

from . import api


def _axis(v, b, D):
    sh = [1] * D
    sh[b] = -1
    return v.reshape(sh)


def velocityfield(setup, ufunc, t=0.0, psolver=None, doproject=True):
    """velocityfield(setup, ufunc, t; psolver, doproject)  src/initializers.jl:13-46.
    `ufunc(dim, x...)` is evaluated on broadcastable coordinate arrays (dim is 1-based)."""
    g = setup.grid
    D, T = g.dimension, g.T
    u = np.zeros(g.N + (D,), dtype=T, order="F")
    for a in range(D):
        xin = [_axis(g.xu[a][b][g.Iu[a][b]], b, D) for b in range(D)]
        u[g.Iu[a] + (a,)] = np.broadcast_to(np.asarray(ufunc(a + 1, *xin), dtype=T), g.Np)
    ud = api.to_device(u, setup)
    api.apply_bc_u_(ud, t, setup)
    if doproject:
        api.project_(ud, setup, psolver)
        api.apply_bc_u_(ud, t, setup)
    return ud


def create_spectrum(setup, kp, rng):
    """create_spectrum(; setup, kp, rng)  src/initializers.jl:82-181 on the GLOBAL grid
    (Orlandi-type spectrum, mirrored random phases, random unit vectors projected ⟂ k)."""
    g = setup.grid
    D, T = g.dimension, g.T
    ctype = np.complex64 if T == np.float32 else np.complex128
    Nglob = tuple(n + 2 for n in g.n_global)
    tau = T(2 * np.pi)
    K = tuple(n // 2 for n in g.n_global)
    kk = [_axis(np.arange(K[a]).astype(T), a, D) for a in range(D)]
    k = np.zeros(K, dtype=T)
    for a in range(D):
        k = k + kk[a] ** 2
    k = np.sqrt(k)
    A = T(T(8) * tau / T(3)) / T(kp) ** 5
    amp = np.sqrt(A * k ** 4 * np.exp(-tau * (k / T(kp)) ** 2)).astype(ctype) * T(np.prod(Nglob))
    xi = [rng.random(K).astype(T) for _ in range(D)]
    for al in range(D):
        amp = np.concatenate((amp, np.flip(amp, axis=al)), axis=al)
        xi = [np.concatenate((xi[b], np.flip((-1 if al == b else 1) * xi[b], axis=al)), axis=al) for b in range(D)]
    xis = xi[0]
    for b in range(1, D):
        xis = xis + xi[b]
    amp = (np.exp(1j * (tau * xis).astype(T)).astype(ctype) * amp).astype(ctype)
    KK = tuple(2 * v for v in K)
    kf = [_axis(np.arange(KK[a]).astype(T), a, D) for a in range(D)]
    knorm = np.zeros(KK, dtype=T)
    for a in range(D):
        knorm = knorm + kf[a] ** 2
    knorm = np.sqrt(knorm)
    if D == 2:
        th = rng.random(KK).astype(T)
        e = [np.cos(2 * np.pi * th).astype(T), np.sin(2 * np.pi * th).astype(T)]
    else:
        th = rng.random(KK).astype(T)
        ph = rng.random(KK).astype(T)
        e = [(np.sin(np.pi * th) * np.cos(2 * np.pi * ph)).astype(T),
             (np.sin(np.pi * th) * np.sin(2 * np.pi * ph)).astype(T), np.cos(np.pi * th).astype(T)]
    ke = sum(e[a] * kf[a] for a in range(D))
    with np.errstate(divide="ignore", invalid="ignore"):
        for a in range(D):
            e0 = e[a][(0,) * D]
            e[a] = (e[a] - kf[a] * ke / knorm ** 2).astype(T)
            e[a][(0,) * D] = e0
    enorm = np.sqrt(sum(e[a] ** 2 for a in range(D)))
    return np.stack([amp * (e[a] / enorm) for a in range(D)], axis=-1).astype(ctype)


def random_field(setup, t=0.0, A=1, kp=10, psolver=None, rng=None):
    """random_field(setup, t; A, kp, psolver, rng)  src/initializers.jl:189-219.  The spectrum is
    generated for the GLOBAL grid on the host (every rank draws the same stream from `rng`), the
    local slab is cut out, and the projection runs on the device (distributed when nranks > 1)."""
    g = setup.grid
    D, T = g.dimension, g.T
    rng = np.random.default_rng(0) if rng is None else rng
    uhat = create_spectrum(setup, kp, rng)
    u = sfft.ifftn(uhat, axes=tuple(range(D)))
    u = (T(A) * u.real).astype(T)
    u = np.pad(u, [(1, 1)] * D + [(0, 0)], mode="wrap")  # pad_circular (initializers.jl:209)
    b = setup.backend
    ul = api.local_slab(u, D, b.rank, b.nranks)
    ud = api.to_device(ul, setup)
    api.apply_bc_u_(ud, t, setup)
    api.project_(ud, setup, psolver)
    return api.apply_bc_u_(ud, t, setup)
