"""Temperature equation on the periodic path, right_hand_side!, Qfield!, dissipation_from_strain! and
get_scale_numbers.  The reference holds no value tests for these (test/chainrules.jl
only differentiates them), so the CPU part pins the oracle restatements with known answers and conservation
identities; the GPU part compares the library (C ABI) with the oracle -- bit for bit for the stencil operators
(true divisions, no FMA contraction), within the north-star tolerance after RK steps."""
import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import compare, rand_p, rand_u

TOL = {np.float32: 1e-5, np.float64: 1e-12}


def osetup(n, T=np.float64, Re=100.0, lims=(0.0, 1.0), gdir=2, dodissipation=True, nondim_type=1):
    x = tuple(np.linspace(lims[0], lims[1], m + 1).astype(T) for m in n)
    tq = o.temperature_equation(T(0.71), T(1e6), T(0.1), dodissipation=dodissipation, gdir=gdir, nondim_type=nondim_type)
    return o.Setup(x, Re=T(Re), temperature=tq)


def pair(n, T, lims=(0.0, 1.0), smag=False, **kw):
    import insb200 as ins

    so = osetup(n, T, lims=lims, **kw)
    x = tuple(np.linspace(lims[0], lims[1], m + 1).astype(T) for m in n)
    tq = ins.temperature_equation(T(0.71), T(1e6), T(0.1), dodissipation=kw.get("dodissipation", True),
                                  gdir=kw.get("gdir", 2), nondim_type=kw.get("nondim_type", 1))
    sc = ins.Setup(x, Re=T(kw.get("Re", 100.0)), temperature=tq)
    if smag:
        so.closure_model = o.smagorinsky_closure(so)
        sc.closure_model = ins.smagorinsky_closure(sc)
    return so, sc


# ------------------------------------------------------------------------------------------------
# CPU: the oracle restatements
# ------------------------------------------------------------------------------------------------
def test_oracle_temperature_coefficients():
    for T in (np.float32, np.float64):
        q = o.temperature_equation(T(0.71), T(1e6), T(0.1))
        assert q.alpha1.dtype == T and q.gdir == 2
        assert np.isclose(q.alpha1, np.sqrt(0.71 / 1e6)) and q.alpha2 == 1
        assert np.isclose(q.alpha4, 1 / np.sqrt(0.71e6)) and np.isclose(q.gamma, 1 / 0.1)
        q2 = o.temperature_equation(T(0.71), T(1e6), T(0.1), nondim_type=2)
        assert np.isclose(q2.alpha2, 0.71e6) and q2.alpha4 == 1
        q3 = o.temperature_equation(T(0.71), T(1e6), T(0.1), nondim_type=3)
        assert np.isclose(q3.alpha1, q3.alpha3) and np.isclose(q3.gamma, 1)


def test_oracle_default_Re_follows_temperature():
    """setup.jl:12: without an explicit Re, `Re = 1 / temperature.α1`."""
    for T in (np.float32, np.float64):
        x = tuple(np.linspace(0.0, 1.0, 9).astype(T) for _ in range(2))
        q = o.temperature_equation(T(0.71), T(1e6), T(0.1))
        s = o.Setup(x, temperature=q)
        assert s.Re.dtype == T and s.Re == T(1) / q.alpha1
        assert o.Setup(x).Re == T(1000)


@pytest.mark.gpu
def test_mirror_default_Re_follows_temperature():
    import insb200 as ins

    T = np.float32
    x = tuple(np.linspace(0.0, 1.0, 17).astype(T) for _ in range(2))
    q = ins.temperature_equation(T(0.71), T(1e6), T(0.1))
    sc = ins.Setup(x, temperature=q)
    so = o.Setup(x, temperature=o.temperature_equation(T(0.71), T(1e6), T(0.1)))
    assert sc.Re == so.Re == T(1) / T(q.alpha1)


def test_mirror_temperature_equation_matches_oracle():
    """The host mirror forms the coefficients of setup.jl:56-86 exactly like the oracle (pure host code)."""
    import insb200 as ins

    for T in (np.float32, np.float64):
        for nd in (1, 2, 3):
            a = ins.temperature_equation(T(0.71), T(1e6), T(0.1), nondim_type=nd, gdir=1, dodissipation=False)
            b = o.temperature_equation(T(0.71), T(1e6), T(0.1), nondim_type=nd, gdir=1, dodissipation=False)
            for k in ("alpha1", "alpha2", "alpha3", "alpha4", "gamma"):
                assert getattr(a, k) == getattr(b, k) and getattr(a, k).dtype == T, (k, nd)
            assert (a.gdir, a.dodissipation) == (b.gdir, b.dodissipation)
    with pytest.raises(ValueError):
        ins.temperature_equation(1.0, 1.0, 1.0, nondim_type=4)


@pytest.mark.parametrize("n", [(16, 24), (8, 12, 16)], ids=["2d", "3d"])
def test_oracle_temperature_transport_identities(n):
    so = osetup(n)
    g = so.grid
    ps = o.psolver_spectral(so)
    u = o.apply_bc_u(o.project(rand_u(so, 0, np.float64), so, ps), 0.0, so)
    # constant temperature + divergence-free velocity: no transport, no conduction
    c = o.convection_diffusion_temp(o.scalarfield(so), u, np.full(g.N, 3.0, order="F"), so)
    assert np.abs(c[g.Ip]).max() < 1e-10
    # flux form: the term sums to zero over the periodic box for any (u, temp), and it ADDS to its first argument
    temp = rand_p(so, 1, np.float64)
    c = o.convection_diffusion_temp(o.scalarfield(so), rand_u(so, 2, np.float64), temp, so)
    assert abs(c[g.Ip].sum()) < 1e-9 * np.abs(c[g.Ip]).sum()
    base = rand_p(so, 3, np.float64)
    c2 = o.convection_diffusion_temp(base, rand_u(so, 2, np.float64), temp, so)
    assert np.allclose(c2[g.Ip] - base[g.Ip], c[g.Ip], rtol=0, atol=1e-9) and np.array_equal(c2[0], base[0])
    # pure conduction of a single mode: -alpha4 * lambda_h * temp
    k = 2
    shape = [1] * len(n)
    shape[0] = -1
    tm = o.apply_bc_temp(np.asfortranarray(np.broadcast_to(np.cos(2 * np.pi * k * g.xp[0]).reshape(shape), g.N)), 0.0, so)
    c = o.convection_diffusion_temp(o.scalarfield(so), o.vectorfield(so), tm, so)
    h = g.Delta[0][1]
    lam = (2 - 2 * np.cos(2 * np.pi * k * h)) / h ** 2
    assert np.allclose(c[g.Ip], -so.temperature.alpha4 * lam * tm[g.Ip], rtol=1e-9, atol=1e-12)


def test_oracle_gravity_and_dissipation():
    so = osetup((16, 16), gdir=2)
    g = so.grid
    F = o.gravity(o.vectorfield(so), np.full(g.N, 2.5, order="F"), so)
    assert np.array_equal(F[g.Ip + (1,)], np.full(g.Np, 2.5 * so.temperature.alpha2)) and not F[..., 0].any()
    # linear temperature in y: the average sits at the u2 point (upper face)
    temp = np.asfortranarray(np.broadcast_to(g.xp[1][None, :], g.N))
    F = o.gravity(o.vectorfield(so), temp, so)
    assert np.allclose(F[g.Ip + (1,)], so.temperature.alpha2 * g.x[1][2:-1][None, :])
    # dissipation: sum_b coef (u diff)|_I averaged -- for a uniform velocity diffusion vanishes
    u = o.vectorfield(so)
    u[...] = 1.0
    assert not o.dissipation(o.scalarfield(so), u, so)[g.Ip].any()
    # interior cells away from the first row/column: equals coef * mean of the two face values of u*diff
    u = rand_u(so, 5, np.float64)
    d = o.dissipation(o.scalarfield(so), u, so)
    diff = o.diffusion(u, so)
    q = so.temperature
    coef = so.Re * q.alpha1 / q.gamma
    I = (slice(2, -1), slice(2, -1))
    want = sum(coef * (np.roll(u[..., b] * diff[..., b], 1, axis=b) + u[..., b] * diff[..., b]) / 2 for b in range(2))
    assert np.allclose(d[I], want[I], rtol=1e-12)
    # reference quirk kept: the ghosts of diff are 0, so the first interior row only sees its own face
    ud = u * diff
    first = coef * ud[1, :, 0] / 2 + coef * (np.roll(ud[1, :, 1], 1) + ud[1, :, 1]) / 2
    assert np.allclose(d[1, 2:-1], first[2:-1], rtol=1e-12) and not diff[0].any()


def test_oracle_rk_with_temperature_conserves_heat_without_dissipation():
    so = osetup((16, 16), dodissipation=False)
    g = so.grid
    ps = o.psolver_spectral(so)
    u = o.random_field(so, kp=3, rng=np.random.default_rng(0))
    temp = rand_p(so, 1, np.float64)
    m = o.RK44(np.float64)
    st = o.create_stepper(m, so, ps, u.copy(order="F"), temp=temp.copy(order="F"))
    for _ in range(5):
        st = o.timestep(m, st, 1e-3)
    assert abs(st.temp[g.Ip].sum() - temp[g.Ip].sum()) < 1e-10 * np.abs(temp[g.Ip]).sum()
    assert np.abs(o.divergence(st.u, so)[g.Ip]).max() < 1e-10
    # buoyancy acts: the velocity differs from the run without temperature
    s0 = o.create_stepper(m, so, ps, u.copy(order="F"))
    for _ in range(5):
        s0 = o.timestep(m, s0, 1e-3)
    assert o.rel_l2(st.u, s0.u) > 1e-6


def test_oracle_qfield_strain_dissipation_and_scale_numbers():
    so = osetup((32, 32, 32), Re=50.0, lims=(0.0, 2 * np.pi))
    g = so.grid
    # one Fourier mode u = (cos y, 0, 0): |du/dy| = |sin y|; Q = 0; eps = nu sin^2 y (2 nu * 2 * (S12)^2)
    u = o.vectorfield(so)
    u[..., 0] = np.cos(g.xp[1])[None, :, None]
    u = o.apply_bc_u(u, 0.0, so)
    assert np.abs(o.Qfield(u, so)[g.Ip]).max() < 1e-12
    eps = o.dissipation_from_strain(u, so)
    assert np.allclose(eps[g.Ip], (1 / 50.0) * np.sin(g.xp[1][1:-1])[None, :, None] ** 2, atol=4e-4)
    sn = o.get_scale_numbers(u, so)
    # <u_i u_i> = 1/2; the reference's uavg sums the component loop D times over ALL components (operators.jl:1577-1585)
    assert np.isclose(sn.uavg, np.sqrt(3 * 0.5), rtol=1e-3)
    hh = 2 * np.pi / 32  # the 4-point mean of one-sided differences is the centred difference: factor sin(h)/h
    assert np.isclose(sn.eps, 0.5 / 50.0 * (np.sin(hh) / hh) ** 2, rtol=1e-9)
    # single mode |k| = 1: E(1) = |uhat|^2 / (2 N^2) = 1/8 => L = 3 pi / (2 uavg^2) / 8
    assert np.isclose(sn.L, 3 * np.pi / 2 / sn.uavg ** 2 / 8, rtol=1e-9)
    assert np.isclose(sn.tau, sn.L / sn.uavg) and np.isclose(sn.Re_int, sn.L * sn.uavg * 50.0)
    assert np.isclose(sn.eta, ((1 / 50.0) ** 3 / sn.eps) ** 0.25) and np.isclose(sn.lam, np.sqrt(5 / 50.0 / sn.eps) * sn.uavg)
    # Q of a solid-body-like rotation u = (-sin y.., ) : Q = -du1/dy du2/dx > 0 inside the vortex of a TGV cell
    u = o.vectorfield(so)
    u[..., 0] = (np.sin(g.xu[0][0])[:, None] * np.cos(g.xu[0][1])[None, :])[:, :, None]
    u[..., 1] = (-np.cos(g.xu[1][0])[:, None] * np.sin(g.xu[1][1])[None, :])[:, :, None]
    u = o.apply_bc_u(u, 0.0, so)
    Q = o.Qfield(u, so)[g.Ip]
    x, y = g.xp[0][1:-1, None, None], g.xp[1][None, 1:-1, None]
    # analytic: Q = -(1/2)(u_x^2 + v_y^2 + 2 u_y v_x) = sin^2 x sin^2 y - cos^2 x cos^2 y
    assert np.abs(Q - (np.sin(x) ** 2 * np.sin(y) ** 2 - np.cos(x) ** 2 * np.cos(y) ** 2)).max() < 0.03


def test_oracle_right_hand_side_is_divergence_free():
    so = osetup((16, 16, 16))
    ps = o.psolver_spectral(so)
    u = rand_u(so, 0, np.float64)
    f = o.apply_bc_u(o.right_hand_side(u, 0.0, so, ps), 0.0, so)  # project! leaves the ghosts stale (pressure.jl:69-82)
    assert np.abs(o.divergence(f, so)[so.grid.Ip]).max() < 1e-9 * np.abs(f).max() * 16


# ------------------------------------------------------------------------------------------------
# GPU: the library against the oracle
# ------------------------------------------------------------------------------------------------
GRIDS = [((32, 64), (0.0, 1.0)), ((16, 8, 32), (0.0, 1.0)), ((20, 12), (0.0, 3.0)), ((12, 10, 14), (0.0, 0.7))]
IDS = ["2d_pow2", "3d_pow2", "2d_any_h", "3d_any_h"]


@pytest.mark.gpu
@pytest.mark.parametrize("T", [np.float32, np.float64], ids=["f32", "f64"])
@pytest.mark.parametrize("n,lims", GRIDS, ids=IDS)
def test_temperature_operators_bit_exact(n, lims, T):
    import insb200 as ins

    gd = 1 + (len(n) - 1) % 2  # exercise both a middle and the first axis
    so, sc = pair(n, T, lims=lims, gdir=gd)
    g = so.grid
    u, temp, base = rand_u(so, 0, T), rand_p(so, 1, T), rand_p(so, 2, T)
    du, dt_, db = ins.to_device(u, sc), ins.to_device(temp, sc), ins.to_device(base, sc)
    # bit-exact when h is a power of two; otherwise the oracle's Delta vectors differ from cell to cell in the last
    # bit (rounding of the coordinate vector), while the library takes ONE h per axis (ctx.cuh): round-off tolerance
    pow2 = lims == (0.0, 1.0) and all((m & (m - 1)) == 0 for m in n)
    rt = 50 * np.finfo(T).eps
    got = ins.to_host(ins.convection_diffusion_temp_(db.clone(), du, dt_, sc))
    compare("convection_diffusion_temp!", got, o.convection_diffusion_temp(base, u, temp, so), rt, exact=pow2)
    F0 = rand_u(so, 3, T)
    got = ins.to_host(ins.gravity_(ins.to_device(F0, sc), dt_, sc))
    compare("gravity!", got, o.gravity(F0, temp, so), rt, exact=pow2)
    got = ins.to_host(ins.dissipation_(db.clone(), ins.vectorfield(sc), du, sc))
    compare("dissipation!", got[g.Ip], o.dissipation(base, u, so)[g.Ip], rt, exact=pow2)
    got = ins.to_host(ins.momentum(du, dt_, 0.0, sc))
    compare("momentum!(temp)", got[g.Ip], o.momentum(u, temp, 0.0, so)[g.Ip], rt, exact=pow2)
    raw = np.asfortranarray(np.random.default_rng(4).standard_normal(g.N).astype(T))
    got = ins.to_host(ins.apply_bc_temp(ins.to_device(raw, sc), 0.0, sc))
    compare("apply_bc_temp!", got, o.apply_bc_temp(raw, 0.0, so), 0, exact=True)
    compare("Qfield!", ins.to_host(ins.Qfield(du, sc))[g.Ip], o.Qfield(u, so)[g.Ip], rt, exact=pow2)
    compare("dissipation_from_strain!", ins.to_host(ins.dissipation_from_strain(du, sc))[g.Ip],
            o.dissipation_from_strain(u, so)[g.Ip], rt, exact=pow2)


@pytest.mark.gpu
@pytest.mark.parametrize("T", [np.float32, np.float64], ids=["f32", "f64"])
@pytest.mark.parametrize("n,smag,diss,method", [((32, 32), False, True, "RK44"), ((16, 16, 32), False, True, "RK44"),
                                                ((16, 32, 16), True, False, "RK44"), ((24, 20), False, True, "Wray3"),
                                                ((32, 32, 32), False, True, "RK44P2")],
                         ids=["2d", "3d", "3d_smag_nodiss", "2d_wray3_any_n", "3d_rk44p2"])
def test_rk_steps_with_temperature(n, smag, diss, method, T):
    import insb200 as ins

    so, sc = pair(n, T, smag=smag, dodissipation=diss)
    g = so.grid
    ps = o.psolver_spectral(so)
    u = o.random_field(so, kp=3, rng=np.random.default_rng(0))
    temp = o.apply_bc_temp(np.asfortranarray(np.random.default_rng(1).random(g.N).astype(T)), 0.0, so)
    mo, mc = o.explicit_method(method, T), getattr(ins.RKMethods, method)(T=T)
    theta = 0.17 if smag else None
    st = o.create_stepper(mo, so, ps, u.copy(order="F"), temp=temp.copy(order="F"))
    stc = ins.create_stepper(mc, sc, None, ins.to_device(u, sc), temp=ins.to_device(temp, sc))
    nstep = 5
    p = None
    for i in range(nstep):
        if i == nstep - 1:
            st, p = o.timestep(mo, st, 1e-3, theta=theta, return_p=True)
        else:
            st = o.timestep(mo, st, 1e-3, theta=theta)
    pc = ins.scalarfield(sc)
    for i in range(nstep):
        stc = ins.timestep_(mc, stc, 1e-3, theta=theta, p=pc if i == nstep - 1 else None)
    tol = TOL[T]
    compare("u after 5 steps", ins.to_host(stc.u), st.u, tol)
    compare("temp after 5 steps", ins.to_host(stc.temp), st.temp, tol)
    compare("p of the last stage", ins.to_host(pc)[g.Ip], p[g.Ip], tol)  # measured <= 7e-15 / 3.3e-6 (profiles/r2_tolerances.txt)
    assert abs(float(stc.t) - float(st.t)) <= 4 * np.finfo(T).eps and stc.n == nstep
    div = o.divergence(ins.to_host(stc.u), so)[g.Ip]
    assert np.abs(div).max() < 2e3 * np.finfo(T).eps * max(1.0, np.abs(st.u).max()) * max(n)
    # the same through solve_unsteady (fixed dt: one insb200_rk_steps_temp call)
    state, _ = ins.solve_unsteady(sc, (0.0, nstep * 1e-3), ins.to_device(u, sc), tempstart=ins.to_device(temp, sc), method=mc,
                                  dt=1e-3, theta=theta)
    compare("solve_unsteady u", ins.to_host(state.u), st.u, tol)
    compare("solve_unsteady temp", ins.to_host(state.temp), st.temp, tol)


@pytest.mark.gpu
def test_temperature_errors_are_loud():
    import insb200 as ins

    x = tuple(np.linspace(0.0, 1.0, 17).astype(np.float32) for _ in range(2))
    sc = ins.Setup(x, Re=np.float32(100))
    with pytest.raises(ValueError):
        ins.gravity(ins.scalarfield(sc), sc)
    with pytest.raises(ins._lib.INSB200Error, match="insb200_set_temperature"):
        sc.ctx.call("insb200_gravity", ins.api._ptr(ins.vectorfield(sc)), ins.api._ptr(ins.scalarfield(sc)))
    with pytest.raises(AssertionError):
        ins.Setup(x, temperature=ins.temperature_equation(np.float32(1), np.float32(1e3), np.float32(1),
                                                          boundary_conditions=(("periodic", "periodic"), ("dirichlet", "dirichlet"))))


@pytest.mark.gpu
@pytest.mark.parametrize("T", [np.float32, np.float64], ids=["f32", "f64"])
@pytest.mark.parametrize("n", [(64, 32), (32, 16, 32), (24, 40, 20)], ids=["2d", "3d", "3d_any_n"])
def test_right_hand_side_and_scale_numbers(n, T):
    import insb200 as ins
    from tests.util import make_pair

    so, sc = make_pair(n, T, Re=200.0, force=len(n) == 2)
    g = so.grid
    ps = o.psolver_spectral(so)
    u = o.random_field(so, kp=4, rng=np.random.default_rng(3))
    du = ins.to_device(u, sc)
    ucopy = du.clone()
    f = ins.create_right_hand_side(sc, None)(du, None, 0.0)
    assert bool((du == ucopy).all())  # sciml.jl:41: u is not touched
    want = o.right_hand_side(u, 0.0, so, ps)
    compare("right_hand_side!", ins.to_host(f)[g.Ip], want[g.Ip], 50 * TOL[T])
    sn, so_n = ins.get_scale_numbers(du, sc), o.get_scale_numbers(u, so)
    for k in ("uavg", "eps", "eta", "lam", "Relam", "L", "tau", "Re_int"):
        a, b = float(getattr(sn, k)), float(getattr(so_n, k))
        assert abs(a - b) <= 20 * TOL[T] * abs(b), (k, a, b)
    # observefield(:Qfield) and (:temperature)
    st = ins.Observable(ins.get_state(ins.create_stepper(None, sc, None, du)))
    out = ins.observefield(sc, "Qfield").initialize(st)
    st.set(st.get())
    want = np.log(np.maximum(np.finfo(T).eps, o.Qfield(u, so)[g.Ip]))
    compare("observefield(:Qfield)", out[-1], want, 1e-5 if T == np.float32 else 1e-12)
