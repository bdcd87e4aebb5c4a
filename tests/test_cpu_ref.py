"""oracle/cpu_ref.c (C + OpenMP restatement of the reference's launch structure, the timed CPU baseline)
against the NumPy oracle: same algorithm, two independent restatements."""
import numpy as np
import pytest

from oracle import cpu_ref
from oracle import ins_oracle as o


def kolmogorov(D):
    if D == 2:
        return lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x
    return lambda dim, x, y, z, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x + 0 * z


@pytest.mark.parametrize("n,T,force,smag", [
    ((32, 24), np.float64, False, False),
    ((16, 32), np.float32, True, False),
    ((16, 12, 20), np.float64, True, True),
    ((16, 16, 16), np.float32, False, True),
    ((24, 16), np.float64, False, True),
], ids=lambda v: str(v).replace(" ", ""))
def test_cpu_ref_matches_numpy_oracle(n, T, force, smag):
    D = len(n)
    x = tuple(np.linspace(0.0, 1.0, m + 1).astype(T) for m in n)
    setup = o.Setup(x, Re=T(500), bodyforce=kolmogorov(D) if force else None)
    theta = 0.17 if smag else None
    if smag:
        setup.closure_model = o.smagorinsky_closure(setup)
    u0 = o.random_field(setup, kp=3, rng=np.random.default_rng(5))
    m = o.RK44(T)
    st = o.create_stepper(m, setup, o.psolver_spectral(setup), u0.copy(order="F"))
    nsteps, dt = 5, 2e-3
    for _ in range(nsteps):
        st = o.timestep(m, st, dt, theta=theta)
    u = cpu_ref.rk_steps(u0.copy(order="F"), setup, dt, nsteps, m, theta)
    err = o.rel_l2(u, st.u)
    assert err <= (1e-12 if T == np.float64 else 1e-5), err
    assert cpu_ref.num_threads() >= 1
