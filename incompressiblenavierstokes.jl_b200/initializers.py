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


def create_spectrum_device(setup, kp, rng):
    """create_spectrum on the device (torch ops on setup.device; same formulas and the same host random stream
    as `create_spectrum`, so both agree to rounding): returns a complex tensor shaped (D, 2K_D, .., 2K_1)."""
    import torch

    g = setup.grid
    D, T = g.dimension, g.T
    dev, rt = setup.device, setup.tdtype
    ct = torch.complex64 if T == np.float32 else torch.complex128
    Nglob = tuple(n + 2 for n in g.n_global)
    tau = float(T(2 * np.pi))
    K = tuple(n // 2 for n in g.n_global)

    def ax(v, a):  # Julia axis a (0-based) is torch axis D-1-a of the reversed-shape tensor
        sh = [1] * D
        sh[D - 1 - a] = -1
        return v.reshape(sh)

    def up(x):  # host array in Julia axis order -> device tensor in reversed (C-order) axes
        return torch.as_tensor(np.ascontiguousarray(np.transpose(x)), device=dev)

    k2 = sum(ax(torch.arange(K[a], device=dev, dtype=rt), a) ** 2 for a in range(D))
    k = torch.sqrt(k2.expand(tuple(reversed(K))))
    A = float(T(T(8) * T(tau) / T(3)) / T(kp) ** 5)
    amp = torch.sqrt(A * k ** 4 * torch.exp(-tau * (k / float(kp)) ** 2)) * float(T(np.prod(Nglob)))
    xi = [up(rng.random(K).astype(T)) for _ in range(D)]
    for al in range(D):
        d = D - 1 - al
        amp = torch.cat((amp, torch.flip(amp, dims=(d,))), dim=d)
        xi = [torch.cat((xi[b], torch.flip((-1 if al == b else 1) * xi[b], dims=(d,))), dim=d) for b in range(D)]
    xis = xi[0]
    for b in range(1, D):
        xis = xis + xi[b]
    amp = torch.polar(amp, tau * xis)
    KK = tuple(2 * v for v in K)
    kf = [ax(torch.arange(KK[a], device=dev, dtype=rt), a) for a in range(D)]
    knorm2 = sum(q ** 2 for q in kf).expand(tuple(reversed(KK)))
    th = up(rng.random(KK).astype(T))
    if D == 2:
        e = [torch.cos(2 * np.pi * th), torch.sin(2 * np.pi * th)]
    else:
        ph = up(rng.random(KK).astype(T))
        e = [torch.sin(np.pi * th) * torch.cos(2 * np.pi * ph), torch.sin(np.pi * th) * torch.sin(2 * np.pi * ph),
             torch.cos(np.pi * th)]
    ke = sum(e[a] * kf[a] for a in range(D))
    zero = (0,) * D
    for a in range(D):
        e0 = e[a][zero].clone()
        e[a] = e[a] - kf[a] * ke / knorm2
        e[a][zero] = e0
    enorm = torch.sqrt(sum(q ** 2 for q in e))
    return torch.stack([amp * (e[a] / enorm).to(ct) for a in range(D)], dim=0)


def random_field(setup, t=0.0, A=1, kp=10, psolver=None, rng=None, ondevice=False):
    """random_field(setup, t; A, kp, psolver, rng)  src/initializers.jl:189-219.  The spectrum is
    generated for the GLOBAL grid (every rank draws the same stream from `rng`), the local slab is cut out,
    and the projection runs on the device (distributed when nranks > 1).  `ondevice=True` (single rank) also
    builds the spectrum and takes the inverse FFT on the device (cuFFT through torch.fft, as the reference does
    through CUDA.jl), so that a 512^3 start field costs seconds instead of a minute of host time."""
    g = setup.grid
    D, T = g.dimension, g.T
    rng = np.random.default_rng(0) if rng is None else rng
    b = setup.backend
    if ondevice:
        import torch

        if getattr(b, "nranks", 1) > 1:
            raise NotImplementedError("ondevice=True builds the whole field on one device")
        uhat = create_spectrum_device(setup, kp, rng)
        u = torch.fft.ifftn(uhat, dim=tuple(range(1, D + 1))).real * float(T(A))
        ud = torch.nn.functional.pad(u.unsqueeze(0), (1, 1) * D, mode="circular").squeeze(0).to(setup.tdtype).contiguous()
    else:
        uhat = create_spectrum(setup, kp, rng)
        u = sfft.ifftn(uhat, axes=tuple(range(D)))
        u = (T(A) * u.real).astype(T)
        u = np.pad(u, [(1, 1)] * D + [(0, 0)], mode="wrap")  # pad_circular (initializers.jl:209)
        ul = api.local_slab(u, D, b.rank, b.nranks)
        ud = api.to_device(ul, setup)
    api.apply_bc_u_(ud, t, setup)
    api.project_(ud, setup, psolver)
    return api.apply_bc_u_(ud, t, setup)
