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


def observefield(setup, fieldname, nupdate=1, psolver=None):
    """observefield(state; setup, fieldname)  src/processors.jl:77-198 for the fields that live on
    the hot path: `:pressure` (pressure!) and `:velocity_raw` (u[Ip, :] as stored).  Returns host
    arrays restricted to Ip."""

    def initialize(state):
        out = []
        g = setup.grid

        def on(s):
            if s.n % nupdate:
                return
            if fieldname == "pressure":
                p = api.pressure(s.u, None, s.t, setup, psolver)
                out.append(api.to_host(p)[g.Ip].copy())
            elif fieldname == "velocity_raw":
                out.append(api.to_host(s.u)[g.Ip].copy())
            else:
                raise ValueError("Unknown fieldname")

        state.on(on)
        return out

    return Processor(initialize, nupdate=nupdate)
