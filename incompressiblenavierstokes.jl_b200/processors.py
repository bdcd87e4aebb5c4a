"""Processors: the observer API of src/processors.jl, headless.

`processor(initialize, finalize)` pairs listen on the `Observable` of `(; u, temp, t, n)` that
`solve_unsteady` updates (src/processors.jl:39-40).  Plotting processors (Makie: realtimeplotter,
animator, fieldplot) need a graphics stack and are unavailable on this box; `energy_history`
returns the points `energy_history_plot` would draw (ext/...MakieExt.jl:281-295) as arrays.
"""
from __future__ import annotations

import time
from types import SimpleNamespace

import numpy as np

from . import api


class Processor:
    def __init__(self, initialize, finalize=None, nupdate=1):
        self.initialize = initialize
        self.finalize = finalize or (lambda initialized, state: initialized)
        self.nupdate = nupdate


def processor(initialize, finalize=None, nupdate=1):
    """processor(initialize, finalize)  src/processors.jl:39-40"""
    return Processor(initialize, finalize, nupdate)


def timelogger(setup, showiter=False, showt=True, showdt=True, showmax=True, showspeed=True, nupdate=1,
               sink=print):
    """timelogger(; ...)  src/processors.jl:45-72; `umax = maximum(abs, u)` is one device reduction."""

    def initialize(state):
        told = [state.get().t]
        oldtime = [time.time()]
        lines = []

        def on(s):
            dt = s.t - told[0]
            told[0] = s.t
            if s.n % nupdate:
                return
            now = time.time()
            itertime = (now - oldtime[0]) / nupdate
            oldtime[0] = now
            msg = []
            if showiter:
                msg.append(f"Iteration {s.n}")
            if showt:
                msg.append(f"t = {float(s.t):g}")
            if showdt:
                msg.append(f"Δt = {float(dt):.2g}")
            if showmax:
                msg.append(f"umax = {api.maxabs(s.u, setup):.2g}")
            if showspeed:
                msg.append(f"itertime = {itertime:.2g}")
            lines.append("\t".join(msg))
            if sink:
                sink(lines[-1])

        state.on(on)
        return lines

    return Processor(initialize, nupdate=nupdate)


def fieldsaver(setup, nupdate=1):
    """fieldsaver(; setup, nupdate)  src/processors.jl:290-300: host copies of (u, t, n)."""

    def initialize(state):
        states = []

        def on(s):
            if s.n % nupdate:
                return
            states.append(SimpleNamespace(u=api.to_host(s.u).copy(order="F"), temp=None, t=s.t, n=s.n))

        state.on(on)
        return states

    return Processor(initialize, nupdate=nupdate)


def energy_history(setup, nupdate=1):
    """Data of energy_history_plot (MakieExt.jl:281-295): (t, E) after every `nupdate` steps,
    E = sum(Ω·kinetic_energy(u)[Ip]) -- computed by the device reduction."""

    def initialize(state):
        pts = []

        def on(s):
            if s.n % nupdate:
                return
            pts.append((float(s.t), api.total_kinetic_energy(s.u, setup)))

        on(state.get())
        state.on(on)
        return pts

    return Processor(initialize, lambda pts, state: np.array(pts), nupdate=nupdate)


def observefield(setup, fieldname, nupdate=1, psolver=None, logtol=None):
    """observefield(state; setup, fieldname, logtol)  src/processors.jl:77-198 for `1 | 2 | 3` (one interpolated
    velocity component), `:velocity`, `:velocitynorm`, `:vorticity`, `:pressure`, `:Qfield`, `:temperature` -- all
    computed on the device by the library -- and `:velocity_raw` (u[Ip, :] as stored).  Returns host arrays
    restricted to Ip."""
    g = setup.grid
    D = g.dimension
    if fieldname not in (1, 2, 3, "velocity", "velocitynorm", "vorticity", "pressure", "velocity_raw", "Qfield", "temperature"):
        raise ValueError("Unknown fieldname")  # processors.jl:128
    logtol = np.finfo(g.T).eps if logtol is None else logtol  # processors.jl:81
    if fieldname in (1, 2, 3) and fieldname > D:
        raise ValueError("Unknown fieldname")

    def initialize(state):
        out = []
        up = api.vectorfield(setup) if fieldname in (1, 2, 3, "velocity", "velocitynorm") else None
        upnorm = api.scalarfield(setup) if fieldname == "velocitynorm" else None
        w = api._wfield(setup) if fieldname == "vorticity" else None
        wp = api._wfield(setup) if fieldname == "vorticity" else None

        def on(s):
            if s.n % nupdate:
                return
            if fieldname == "pressure":
                f = api.to_host(api.pressure(s.u, None, s.t, setup, psolver))[g.Ip]
            elif fieldname == "velocity_raw":
                f = api.to_host(s.u)[g.Ip]
            elif fieldname == "temperature":
                f = api.to_host(s.temp)[g.Ip]
            elif fieldname == "Qfield":  # processors.jl:172-176: Qfield!, then `@. Qin = log(max(logtol, Qin))` (device broadcast)
                Q = api.Qfield(s.u, setup)
                f = api.to_host(Q.clamp_min_(float(logtol)).log_())[g.Ip]
            elif fieldname in (1, 2, 3):
                f = api.to_host(api.interpolate_u_p_(up, s.u, setup))[g.Ip + (fieldname - 1,)]
            elif fieldname == "velocity":
                f = api.to_host(api.interpolate_u_p_(up, s.u, setup))[g.Ip]
            elif fieldname == "velocitynorm":
                api.interpolate_u_p_(up, s.u, setup)
                f = api.to_host(api.velocitynorm_(upnorm, up, setup))[g.Ip]
            else:  # vorticity: processors.jl:152-155
                api.apply_bc_u_(s.u, s.t, setup)
                api.vorticity_(w, s.u, setup)
                f = api.to_host(api.interpolate_w_p_(wp, w, setup))[g.Ip]
            out.append(f.copy())

        state.on(on)
        return out

    return Processor(initialize, nupdate=nupdate)


def spectral_stuff(setup, npoint=100, a=None):
    """spectral_stuff(setup; npoint, a)  src/utils.jl:49-108: wavenumber shells of the energy spectrum.
    `inds` are 0-based linear indices (first axis fastest) into the K1 x .. x KD block of low modes."""
    g = setup.grid
    D, T, Np = g.dimension, g.T, g.Np
    a = T((1 + np.sqrt(5)) / 2) if a is None else T(a)
    K = tuple(n // 2 for n in Np)
    grids = np.meshgrid(*[np.arange(k, dtype=np.int64) for k in K], indexing="ij")
    k = np.sqrt(sum(q.astype(np.float64) ** 2 for q in grids)).astype(T).reshape(-1, order="F")
    kmax = min(K) - 1
    isort = np.argsort(k, kind="stable")
    ksort = k[isort]
    kappa = np.exp(np.linspace(0.0, np.log(float(kmax)), npoint))
    kappa = np.array(sorted({int(v) for v in np.round(kappa)}), dtype=np.int64)
    inds = []
    for ki in kappa:
        if D == 2:  # dyadic binning
            lo, hi = T(ki) / a, T(ki) * a
        else:       # linear binning
            lo, hi = T(ki) - T(0.01), T(ki) + T(1) - T(0.01)
        inds.append(isort[int(np.searchsorted(ksort, T(lo), side="left")):int(np.searchsorted(ksort, T(hi), side="left"))])
    return inds, kappa, K


def observespectrum(setup, npoint=100, a=None, nupdate=1):
    """observespectrum(state; setup, npoint, a)  src/processors.jl:303-332, on the device: per component a
    complex FFT of u[Ip] (cuFFT through torch.fft, as the reference goes through CUDA.jl's CUFFT), |.|^2 of the
    K1 x .. x KD low modes / (2 prod(Np)^2), summed over components and over each shell.  Single rank.
    `finalize` returns (ehat history [nobs, nshell], kappa)."""
    import torch

    g = setup.grid
    D, T, Np = g.dimension, g.T, g.Np
    if getattr(setup.backend, "nranks", 1) > 1:
        raise NotImplementedError("observespectrum needs the whole field on one device")
    inds, kappa, K = spectral_stuff(setup, npoint, a)
    dev = setup.device
    shell = torch.full((int(np.prod(K)),), len(inds), dtype=torch.int64, device=dev)  # modes outside every shell -> dump bin
    multi = np.zeros(int(np.prod(K)), dtype=np.int64)
    for i, ix in enumerate(inds):
        multi[ix] += 1
    if multi.max() > 1:  # overlapping (dyadic, 2-D) shells: fall back to one masked sum per shell
        masks = [torch.as_tensor(ix, device=dev) for ix in inds]
    else:
        masks = None
        for i, ix in enumerate(inds):
            shell[torch.as_tensor(ix, device=dev)] = i
    scale = T(2 * int(np.prod(Np)) ** 2)
    ctype = torch.complex64 if T == np.float32 else torch.complex128
    interior = tuple(slice(1, n + 1) for n in reversed(Np))  # device fields are (comp, N_D, .., N_1)
    low = tuple(slice(0, k) for k in reversed(K))

    def initialize(state):
        out = []

        def on(s):
            if s.n % nupdate:
                return
            e = None
            for al in range(D):
                uhat = torch.fft.fftn(s.u[(al,) + interior].to(ctype))[low]
                ea = (uhat.real ** 2 + uhat.imag ** 2) / scale
                e = ea if e is None else e + ea
            ef = e.reshape(-1)  # C-order flatten of (K_D, .., K_1) == first axis fastest
            if masks is None:
                eh = torch.zeros(len(inds) + 1, dtype=ef.dtype, device=dev).index_add_(0, shell, ef)[:-1]
            else:
                eh = torch.stack([ef[m].sum() for m in masks])
            out.append(eh.cpu().numpy())

        state.on(on)
        return out

    return Processor(initialize, lambda o, state: (np.array(o), kappa), nupdate=nupdate)


# --------------------------------------------------------------------------------------------
# raw slab checkpoints (restart of long runs, solver.jl:35-46 takes `ustart`, `tlims[1]`; the reference itself
# only writes VTK through fieldsaver / vtk_writer).  Format: `<stem>.rank<r>of<P>.json` (D, global n, slab range,
# dtype, t, n) next to `<stem>.rank<r>of<P>.bin`, the rank's field exactly as in memory: reference layout
# (N1, .., ND_local + 2, D) column-major, ghosts included.  Every rank writes and reads only its own slab.
# --------------------------------------------------------------------------------------------
def _ckpt_paths(stem, setup):
    b = setup.backend
    r, p = getattr(b, "rank", 0), getattr(b, "nranks", 1)
    base = f"{stem}.rank{r}of{p}"
    return base + ".json", base + ".bin"


def save_checkpoint(stem, state, setup):
    """Write the state `(; u, t, n)` of this rank; returns the two file names."""
    import json

    g = setup.grid
    jp, bp = _ckpt_paths(stem, setup)
    u = state.u.detach().cpu().contiguous().numpy()
    with open(bp, "wb") as f:
        f.write(u.tobytes())
    meta = {"format": "insb200-slab-1", "D": g.dimension, "n_global": list(g.n_global), "N_local": list(g.N),
            "dtype": np.dtype(g.T).name, "t": float(state.t), "n": int(state.n),
            "rank": getattr(setup.backend, "rank", 0), "nranks": getattr(setup.backend, "nranks", 1),
            "layout": "column-major (N1,..,ND,D), ghosts included", "bytes": int(u.nbytes)}
    with open(jp, "w") as f:
        json.dump(meta, f)
    return jp, bp


def load_checkpoint(stem, setup):
    """Read this rank's slab back: returns the state `(; u, temp, t, n)` with u on the device."""
    import json
    from types import SimpleNamespace

    import torch

    g = setup.grid
    jp, bp = _ckpt_paths(stem, setup)
    with open(jp) as f:
        meta = json.load(f)
    if (meta["format"] != "insb200-slab-1" or meta["D"] != g.dimension or list(meta["n_global"]) != list(g.n_global)
            or list(meta["N_local"]) != list(g.N) or meta["dtype"] != np.dtype(g.T).name):
        raise ValueError(f"checkpoint {jp} does not belong to this setup: {meta}")
    shape = (g.dimension,) + tuple(reversed(g.N))
    u = np.fromfile(bp, dtype=g.T)
    if u.size != int(np.prod(shape)):
        raise ValueError(f"checkpoint {bp} is truncated")
    ud = torch.from_numpy(u.reshape(shape)).to(setup.device)
    return SimpleNamespace(u=ud, temp=None, t=g.T(meta["t"]), n=int(meta["n"]))


def checkpointer(setup, stem, nupdate=1):
    """Processor writing `save_checkpoint(f"{stem}_{n:06d}", state, setup)` every `nupdate` steps (the restart
    counterpart of fieldsaver, processors.jl:290-300).  `finalize` returns the list of written stems."""

    def initialize(state):
        written = []

        def on(s):
            if s.n % nupdate:
                return
            name = f"{stem}_{s.n:06d}"
            save_checkpoint(name, s, setup)
            written.append(name)

        state.on(on)
        return written

    return Processor(initialize, nupdate=nupdate)
