#!/usr/bin/env python
"""Generates tests/golden/*.npz: input / output vectors of the hot path.

The reference is Julia and ships no golden vectors, and no Julia runtime exists in the
build image, so these vectors come from the NumPy oracle (oracle/ins_oracle.py) AFTER it has passed the
reference's own property / known-answer tests (tests/test_oracle_properties.py).  They pin the oracle
against silent drift and give the CUDA path fixed inputs; parity with the reference's numbers proper
stays "unpinned by reference vectors" (DESIGN.md).

    python tests/golden/make_golden.py        # rewrites the fixtures (deterministic)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import ins_oracle as o  # noqa: E402

CASES = {
    # name: (n, dtype, Re, force, theta, nsteps, dt)
    "rk44_2d_f64": ((24, 16), np.float64, 1000.0, False, None, 8, 2e-3),
    "rk44_2d_f32_kolmogorov": ((16, 32), np.float32, 2000.0, True, None, 8, 1e-3),
    "rk44_3d_f64": ((16, 12, 8), np.float64, 1000.0, False, None, 5, 2e-3),
    "rk44_3d_f32_smag_force": ((32, 16, 12), np.float32, 4000.0, True, 0.17, 5, 1e-3),
}


def kolmogorov(D):
    if D == 2:
        return lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x
    return lambda dim, x, y, z, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x + 0 * z


def make_setup(n, T, Re, force, theta):
    x = tuple(np.linspace(0.0, 1.0, m + 1).astype(T) for m in n)
    s = o.Setup(x, Re=T(Re), bodyforce=kolmogorov(len(n)) if force else None)
    if theta is not None:
        s.closure_model = o.smagorinsky_closure(s)
    return s


def run_case(name):
    n, T, Re, force, theta, nsteps, dt = CASES[name]
    setup = make_setup(n, T, Re, force, theta)
    u0 = o.random_field(setup, kp=3, rng=np.random.default_rng(len(name)))
    m = o.RK44(T)
    ps = o.psolver_spectral(setup)
    st = o.create_stepper(m, setup, ps, u0.copy(order="F"))
    ke = [o.total_kinetic_energy(u0, setup)]
    p = None
    for _ in range(nsteps):
        st, p = o.timestep(m, st, dt, theta=theta, return_p=True)
        ke.append(o.total_kinetic_energy(st.u, setup))
    out = dict(u0=u0, u=st.u, p=p, ke=np.array(ke, dtype=np.float64), t=np.float64(st.t),
               momentum=o.momentum(u0, None, 0.0, setup), divergence=o.divergence(st.u, setup),
               pressure=o.pressure(u0, None, 0.0, setup, ps), cfl=np.float64(o.get_cfl_timestep(u0, setup)))
    if theta is not None:
        out["closure"] = setup.closure_model(u0, theta)
    return out


if __name__ == "__main__":
    for name in CASES:
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **run_case(name))
        print("wrote", name)
