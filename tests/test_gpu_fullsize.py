"""Parity at BASELINE.json's FULL sizes (configs 3 and 4) against the oracle.

The oracle cannot step 512^3 in seconds, so the full-size runs use a size-independent property of the
periodic problem: a field of period 1 on the box [0, R)^D, sampled with the same spacing h, evolves
exactly like the field on [0, 1)^D -- every stencil sees the same neighbours, and the Poisson problem has
the same solution (only the modes that are multiples of R are occupied).  The small problem is stepped by
the oracle, the R^D-fold tiled copy by the CUDA path at the full grid size of the BASELINE config:

* stencil operators (momentum!, Smagorinsky closure) must agree BIT FOR BIT with the tiled oracle result
  (h is a power of two; no FFT involved) -- this exercises every tile / chunk boundary of the full grid;
* RK44 steps must agree to north_star's Float32 tolerance (rel-L2 <= 1e-5; the FFTs differ in length);
* kinetic energy scales with the volume, the projected field is divergence free to round-off.
"""
import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import axes

pytestmark = pytest.mark.gpu
T = np.float32
TOL = 1e-5


def tile_periodic(a, reps, D):
    """ghosted small field (n+2, .., [D]) -> ghosted field of the `reps`-fold tiled box."""
    inner = a[(slice(1, -1),) * D]
    big = np.tile(inner, (reps,) * D + (1,) * (a.ndim - D))
    pad = [(1, 1)] * D + [(0, 0)] * (a.ndim - D)
    return np.asfortranarray(np.pad(big, pad, mode="wrap"))


def rel_l2(got, want):
    """relative L2 error of a vector field, accumulated in Float64 component by component (the arrays are GBs)."""
    num = den = 0.0
    for c in range(got.shape[-1]):
        w = want[..., c].astype(np.float64)
        d = got[..., c].astype(np.float64) - w
        num += float((d * d).sum())
        den += float((w * w).sum())
    return np.sqrt(num / max(den, 1e-300))


def kolmogorov(dim, x, y, t):
    """examples/Kolmogorov2D.jl:29-36"""
    return (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x


def build(nsmall, reps, Re, force=False, smag=False):
    import insb200 as ins

    D = len(nsmall)
    xs = axes(nsmall, T, (0.0, 1.0))
    xb = axes(tuple(m * reps for m in nsmall), T, (0.0, float(reps)))
    so = o.Setup(xs, Re=T(Re), bodyforce=kolmogorov if force else None)
    sc = ins.Setup(xb, Re=T(Re))
    assert [float(v) for v in sc.grid.h] == [1.0 / m for m in nsmall]  # same spacing, a power of two
    if force:  # the steady force field of the small box, tiled (setup.jl:25-32 materialises it as a field)
        sc.bodyforce = ins.to_device(tile_periodic(so.bodyforce, reps, D), sc)
        sc.issteadybodyforce = True
        sc.ctx.call("insb200_set_bodyforce", ins.api._ptr(sc.bodyforce))
    if smag:
        so.closure_model = o.smagorinsky_closure(so)
        sc.closure_model = ins.smagorinsky_closure(sc)
    return ins, so, sc


def step_both(ins, so, sc, u0, reps, nsteps, dt, theta):
    D = so.grid.dimension
    mo = o.RK44(T)
    st = o.create_stepper(mo, so, o.psolver_spectral(so), u0.copy(order="F"))
    ke_o = [o.total_kinetic_energy(u0, so)]
    for _ in range(nsteps):
        st = o.timestep(mo, st, dt, theta=theta)
        ke_o.append(o.total_kinetic_energy(st.u, so))
    mc = ins.RKMethods.RK44(T=T)
    ud = ins.to_device(tile_periodic(u0, reps, D), sc)
    stc = ins.create_stepper(mc, sc, ins.default_psolver(sc), ud)
    ke_c = [ins.total_kinetic_energy(ud, sc)]
    for _ in range(nsteps):
        stc = ins.timestep_(mc, stc, dt, theta=theta)
        ke_c.append(ins.total_kinetic_energy(stc.u, sc))
    return st, stc, np.array(ke_o, dtype=np.float64), np.array(ke_c, dtype=np.float64)


def check_run(ins, so, sc, st, stc, ke_o, ke_c, reps):
    D = so.grid.dimension
    want = tile_periodic(st.u, reps, D)
    got = ins.to_host(stc.u)
    assert got.shape == want.shape
    err = rel_l2(got, want)
    assert np.isfinite(err) and err <= TOL, f"u after the steps: rel_l2 = {err:.3e}"
    assert np.allclose(ke_c, ke_o * reps ** D, rtol=TOL, atol=0), (ke_c, ke_o * reps ** D)
    div = ins.to_host(ins.divergence(stc.u, sc))[(slice(1, -1),) * D]
    n = max(sc.grid.N) - 2
    assert np.abs(div).max() / n < 100 * np.finfo(T).eps * float(np.abs(st.u).max())


def test_config4_512cube_smagorinsky_matches_tiled_oracle():
    """BASELINE config 4: DecayingTurbulence3D 512^3 Float32 + Smagorinsky (theta = 0.17), Re = 4e3."""
    reps, ns = 8, (64, 64, 64)
    ins, so, sc = build(ns, reps, 4e3, smag=True)
    assert tuple(sc.grid.N) == (514, 514, 514)
    u0 = o.random_field(so, A=1.0, kp=10, rng=np.random.default_rng(4))
    # stencil operators at full size: bit for bit
    ud = ins.to_device(tile_periodic(u0, reps, 3), sc)
    inner = (slice(1, -1),) * 3
    F = ins.to_host(ins.momentum(ud, None, 0.0, sc))
    assert np.array_equal(F[inner], tile_periodic(o.momentum(u0, None, 0.0, so), reps, 3)[inner]), "momentum! at 512^3"
    del F
    s = ins.to_host(sc.closure_model(ud, T(0.17)))
    assert np.array_equal(s[inner], tile_periodic(so.closure_model(u0, T(0.17)), reps, 3)[inner]), "Smagorinsky at 512^3"
    del s, ud
    # three RK44 steps at the BASELINE time step
    st, stc, ke_o, ke_c = step_both(ins, so, sc, u0, reps, 3, 2.5e-4, T(0.17))
    check_run(ins, so, sc, st, stc, ke_o, ke_c, reps)
    assert ke_c[-1] < ke_c[0]  # decaying turbulence


def test_config3_4096sq_kolmogorov_matches_tiled_oracle():
    """BASELINE config 3: Kolmogorov2D 4096^2 Float32, Re = 2e3, steady force 5 sinpi(8y) stored as a field."""
    reps, ns = 32, (128, 128)
    ins, so, sc = build(ns, reps, 2e3, force=True)
    assert tuple(sc.grid.N) == (4098, 4098)
    u0 = o.random_field(so, A=1e-2, kp=10, rng=np.random.default_rng(3))
    ud = ins.to_device(tile_periodic(u0, reps, 2), sc)
    inner = (slice(1, -1),) * 2
    F = ins.to_host(ins.momentum(ud, None, 0.0, sc))
    assert np.array_equal(F[inner], tile_periodic(o.momentum(u0, None, 0.0, so), reps, 2)[inner]), "momentum! + force at 4096^2"
    del F, ud
    st, stc, ke_o, ke_c = step_both(ins, so, sc, u0, reps, 5, 2.5e-5, None)
    check_run(ins, so, sc, st, stc, ke_o, ke_c, reps)
