"""Unsteady body force -- the `issteadybodyforce = false` branch of applybodyforce! (src/operators.jl:870-877), row a6 of
SURVEY 8a: `Fα .+= bodyforce.(α, xin..., t)` is evaluated at every call, in the RK step once per stage at the stage's
time (momentum!(ku[i], u, temp, t, setup) with `t = tstart + c[i] Δt` set AFTER it, step_explicit_runge_kutta.jl:21,32).
The library asks the caller for the field through `insb200_set_bodyforce_callback`; compared with the oracle."""
import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import axes, compare

pytestmark = pytest.mark.gpu


def force2(dim, x, y, t):
    return (dim == 1) * 5 * np.sin(np.pi * 8 * y) * np.cos(3 * t) + (dim == 2) * 0.5 * np.cos(2 * np.pi * x) * (1 + t) + 0 * x * y


def force3(dim, x, y, z, t):
    return (dim == 1) * 2 * np.sin(2 * np.pi * y) * np.cos(40 * t) + (dim == 3) * np.sin(2 * np.pi * x) * (1 + 10 * t) + 0 * x * y * z


@pytest.mark.parametrize("n,T,smag", [((32, 48), np.float64, False), ((32, 16, 24), np.float32, True), ((32, 32, 16), np.float64, False),
                                      ((20, 12, 10), np.float32, False)], ids=lambda v: str(v).replace(" ", ""))
def test_unsteady_bodyforce_steps_match_oracle(n, T, smag):
    import insb200 as ins

    D = len(n)
    f = force2 if D == 2 else force3
    x = axes(n, T)
    so = o.Setup(x, Re=T(500), bodyforce=f, issteadybodyforce=False)
    sc = ins.Setup(x, Re=T(500), bodyforce=f, issteadybodyforce=False)
    assert not so.issteadybodyforce and not sc.issteadybodyforce
    theta = None
    if smag:
        so.closure_model, sc.closure_model, theta = o.smagorinsky_closure(so), ins.smagorinsky_closure(sc), 0.17
    tol = 1e-12 if T == np.float64 else 1e-5
    u0 = o.random_field(so, kp=3, rng=np.random.default_rng(11))
    # momentum! at a time t != 0
    t1 = T(0.37)
    compare("momentum!(F, u, nothing, t, setup)", ins.to_host(ins.momentum(ins.to_device(u0, sc), None, t1, sc)), o.momentum(u0, None, t1, so),
            1e-13 if T == np.float64 else 2e-6)
    # RK44 steps from t0 = 0.1: every stage sees the force of its own time
    mo, mc = o.RK44(T), ins.RKMethods.RK44(T=T)
    dt, nsteps, t0 = 2e-3, 5, 0.1
    st = o.create_stepper(mo, so, o.psolver_spectral(so), u0.copy(order="F"), t=T(t0))
    for _ in range(nsteps):
        st = o.timestep(mo, st, dt, theta=theta)
    stc = ins.create_stepper(mc, sc, None, ins.to_device(u0, sc), t=t0)
    for _ in range(nsteps):
        stc = ins.timestep_(mc, stc, dt, theta=theta)      # timestep!: one C call per step
    compare("u after 5 timestep! calls", ins.to_host(stc.u), st.u, tol)
    assert float(stc.t) == float(st.t)
    ud = ins.to_device(u0, sc)
    ins.api._rk_steps(mc, sc, ud, t0, dt, nsteps, theta)   # all steps inside one C call
    compare("u after insb200_rk_steps", ins.to_host(ud), st.u, tol)
    # a steady run from the same force frozen at t = 0 must differ (the time dependence is really seen)
    s0 = o.Setup(x, Re=T(500), bodyforce=f)
    if smag:
        s0.closure_model = o.smagorinsky_closure(s0)
    st0 = o.create_stepper(mo, s0, o.psolver_spectral(s0), u0.copy(order="F"), t=T(t0))
    for _ in range(nsteps):
        st0 = o.timestep(mo, st0, dt, theta=theta)
    assert o.rel_l2(st0.u, st.u) > 100 * tol


def test_unsteady_bodyforce_callback_errors_are_reported():
    import insb200 as ins

    T = np.float32
    x = axes((16, 16), T)

    def bad(dim, x, y, t):
        raise RuntimeError("boom")

    sc = ins.Setup(x, Re=T(100), bodyforce=lambda dim, x, y, t: 0 * x * y + 1, issteadybodyforce=False)
    u = ins.vectorfield(sc)
    ins.momentum(u, None, 0.0, sc)
    sc2 = ins.Setup(x, Re=T(100), bodyforce=bad, issteadybodyforce=False)
    with pytest.raises(ins._lib.INSB200Error, match="body-force callback failed"):
        ins.momentum(ins.vectorfield(sc2), None, 0.0, sc2)
    assert isinstance(sc2._force_error, RuntimeError)
