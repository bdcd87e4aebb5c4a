"""Why bench.py caps dt below half the reference's CFL bound in 3-D: RK44 and the diffusion operator.

get_cfl_timestep! (solver.jl:101-125) bounds the step per axis (Re dx^2 / 2).  When that bound is the active one, half of
it gives nu dt 4/dx^2 = 1 per axis, so the stiffest mode of a 3-D grid has z = -3.0, outside RK44's stability interval on the
negative real axis (2.785): amplification 1.375 per step.  Shown here with the ORACLE (the CPU restatement of the reference,
no GPU code involved) at 16^3, Re = 10: the field decays, then round-off in the (pi, pi, pi) mode takes over.
    python tools/rk44_diffusion_limit.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ins_oracle as o  # noqa: E402

n, Re = 16, 10.0
x = tuple(np.linspace(0, 1, n + 1) for _ in range(3))
so = o.Setup(x, Re=Re)
u = 1e-3 * o.random_field(so, kp=2, rng=np.random.default_rng(1))
m = o.RK44(np.float64)
bound = Re / n ** 2 / 2
print(f"reference CFL bound (diffusive, per axis): {bound:.6f}")
for frac in (0.5, 0.8 * 2.785 / 3.0 * 0.5):
    st = o.create_stepper(m, so, o.psolver_spectral(so), u.copy(order="F"))
    dt = frac * bound
    z = -(1 / Re) * dt * 12 * n ** 2
    R = 1 + z + z ** 2 / 2 + z ** 3 / 6 + z ** 4 / 24
    for i in range(160):
        st = o.timestep(m, st, dt)
        if i % 40 == 39:
            print(f"dt = {frac:.3f} x bound, z = {z:.2f}, |R(z)| = {abs(R):.3f}: step {i + 1:3d} max|u| {np.abs(st.u).max():.3e}")
