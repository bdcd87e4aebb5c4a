"""Steppers beyond RK44: the explicit tableaux of RKMethods.jl and the low-storage
LMWray3 step.  CPU part: the oracle's low-storage restatement of step_lmwray3.jl equals the explicit stage
combination with the unrolled tableau bit for bit (that is how the library runs it), LMWray3 converges with
third order, and the host mirror's tableaux agree with the oracle's separately typed copies.
GPU part: every method advanced on the CUDA path matches the oracle."""
from types import SimpleNamespace

import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import compare, make_pair

TOL = {np.float64: 1e-12, np.float32: 1e-5}


def oracle_setup(n, T=np.float64, Re=500.0):
    x = tuple(np.linspace(0.0, 1.0, m + 1).astype(T) for m in n)
    return o.Setup(x, Re=T(Re))


def unrolled_wray(T):
    m = o.LMWray3(T)
    a, b = m.a, m.b
    A = np.array([[a[0], 0, 0], [b[0], a[1], 0], [b[0], b[1], a[2]]], dtype=T)
    return SimpleNamespace(A=A, b=A[-1].copy(), c=np.array([m.c[1], m.c[2], 1], dtype=T), r=T(0))


@pytest.mark.parametrize("T", [np.float64, np.float32], ids=["f64", "f32"])
def test_lmwray3_equals_unrolled_tableau_bitwise(T):
    so = oracle_setup((24, 16), T)
    u0 = o.random_field(so, kp=3, rng=np.random.default_rng(3))
    ps = o.psolver_spectral(so)
    st_l = o.create_stepper(o.LMWray3(T), so, ps, u0.copy(order="F"))
    st_e = o.create_stepper(None, so, ps, u0.copy(order="F"))
    for _ in range(3):
        st_l = o.timestep_lmwray3(o.LMWray3(T), st_l, 2e-3)
        st_e = o.timestep(unrolled_wray(T), st_e, 2e-3)
    assert np.array_equal(st_l.u, st_e.u)
    assert st_l.t == st_e.t and st_l.n == st_e.n == 3


def test_lmwray3_is_third_order():
    so = oracle_setup((16, 16), np.float64, Re=100.0)
    u0 = o.random_field(so, kp=2, rng=np.random.default_rng(5))
    ps = o.psolver_spectral(so)

    def run(step, method, dt, nsteps):
        st = o.create_stepper(method, so, ps, u0.copy(order="F"))
        for _ in range(nsteps):
            st = step(method, st, dt)
        return st.u

    ref = run(o.timestep, o.RK44(np.float64), 0.04 / 64, 64)
    e1 = o.rel_l2(run(o.timestep_lmwray3, o.LMWray3(), 0.04 / 4, 4), ref)
    e2 = o.rel_l2(run(o.timestep_lmwray3, o.LMWray3(), 0.04 / 8, 8), ref)
    assert 6.0 < e1 / e2 < 10.5, (e1, e2)


def test_mirror_tableaux_match_oracle_copies():
    import insb200 as ins

    for name in o.EXPLICIT_TABLEAUX:
        for T in (np.float64, np.float32):
            mo, mc = o.explicit_method(name, T), getattr(ins.RKMethods, name)(T=T)
            assert mc.A.dtype == T and np.array_equal(mo.A, mc.A) and np.array_equal(mo.c, mc.c), name
    # consistency conditions of every tableau in the mirror: rows of the unshifted A sum to c, b sums to 1
    for name in ins.RKMethods.names:
        m = getattr(ins.RKMethods, name)()
        assert abs(m.b.sum() - 1) < 1e-14 and np.allclose(m.A[-1], m.b), name
        assert np.allclose(np.triu(m.A, 1), 0), name
    w = ins.LMWray3(np.float32)
    assert np.array_equal(w.A, unrolled_wray(np.float32).A) and np.array_equal(w.c, unrolled_wray(np.float32).c)
    with pytest.raises(AttributeError):
        ins.RKMethods.GL2  # implicit: not on the explicit stepper


GPU_CASES = [("LMWray3", (32, 16, 16), np.float64, True), ("LMWray3", (32, 32), np.float32, False),
             ("FE11", (32, 16, 16), np.float32, False), ("SSP33", (32, 16, 16), np.float64, False),
             ("Wray3", (24, 20, 16), np.float64, True), ("RK56", (32, 16, 16), np.float64, True),
             ("HEM5", (32, 32), np.float64, False), ("RK44P2", (32, 16, 32), np.float32, True),
             ("NSSP53", (32, 16, 16), np.float64, False), ("SSP43", (32, 16, 16), np.float32, True),
             ("RK33C2", (16, 16), np.float64, False)]


@pytest.mark.gpu
@pytest.mark.parametrize("name,n,T,smag", GPU_CASES, ids=lambda v: str(v).replace(" ", ""))
def test_methods_match_oracle_on_gpu(name, n, T, smag):
    import insb200 as ins
    from insb200.api import _rk_steps

    so, sc = make_pair(n, T, Re=800.0, force=False, smag=smag)
    theta = 0.17 if smag else None
    u0 = o.random_field(so, kp=3, rng=np.random.default_rng(17))
    nsteps, dt = 5, 1.5e-3
    if name == "LMWray3":
        mo, mc, step = o.LMWray3(T), ins.LMWray3(T), o.timestep_lmwray3
    else:
        mo, mc, step = o.explicit_method(name, T), getattr(ins.RKMethods, name)(T=T), o.timestep
    st = o.create_stepper(mo, so, o.psolver_spectral(so), u0.copy(order="F"))
    for _ in range(nsteps):
        st = step(mo, st, dt, theta=theta)
    # one timestep! through the reference-layout entry, the rest inside one insb200_rk_steps call
    ud = ins.to_device(u0, sc)
    stc = ins.timestep_(mc, ins.create_stepper(mc, sc, ins.default_psolver(sc), ud), dt, theta=theta)
    t, _ = _rk_steps(mc, sc, stc.u, stc.t, dt, nsteps - 1, theta)
    compare(f"{name} u after {nsteps} steps", ins.to_host(stc.u), st.u, TOL[T])
    assert abs(float(t) - float(st.t)) <= 1e-6 * abs(float(st.t))
    div = ins.to_host(ins.divergence(stc.u, sc))[so.grid.Ip]
    assert np.abs(div).max() <= (1e-10 if T == np.float64 else 2e-3)
