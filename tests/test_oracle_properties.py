"""Pin the oracle against the reference's own property / known-answer tests.

Each test cites the reference test it restates (paths relative to /root/reference).  The
reference fixtures use non-uniform Dirichlet grids; the hot path's admissible domain is
periodic + uniform (src/utils.jl:1-13), so the same properties are checked there (they hold
for any BC) -- plus once on a NON-uniform periodic grid to exercise the metric arrays.
"""
import numpy as np
import pytest

from oracle import ins_oracle as o


def make_setup(D, n, T=np.float64, lims=(0.0, 1.0), Re=1000.0, nonuniform=False, **kw):
    xs = []
    for a in range(D):
        x = np.linspace(lims[0], lims[1], n + 1).astype(T)
        if nonuniform:
            s = np.linspace(0, 1, n + 1)
            x = (lims[0] + (lims[1] - lims[0]) * (s + 0.1 * (a + 1) / 3 * np.sin(2 * np.pi * s) / (2 * np.pi))).astype(T)
        xs.append(x)
    return o.Setup(tuple(xs), Re=T(Re), **kw)


def weights_u(setup, a):
    """Omega_u for component a: Du along a, D along the others (test/operators.jl:71-85)."""
    g = setup.grid
    D = g.dimension
    w = 1.0
    for b in range(D):
        v = g.Delta_u[b] if a == b else g.Delta[b]
        sh = [1] * D
        sh[b] = -1
        w = w * v.reshape(sh)
    return w


@pytest.mark.parametrize("D", [2, 3])
@pytest.mark.parametrize("nonuniform", [False, True])
def test_divergence_is_minus_gradient_transpose(D, nonuniform):
    """test/operators.jl:58-88  `pDv ≈ -vGp`."""
    setup = make_setup(D, 16, nonuniform=nonuniform)
    rng = np.random.default_rng(0)
    g = setup.grid
    v = o.apply_bc_u(np.asfortranarray(rng.standard_normal(g.N + (D,))), 0.0, setup)
    p = o.apply_bc_p(np.asfortranarray(rng.standard_normal(g.N)), 0.0, setup)
    Dv = o.scalewithvolume(o.divergence(v, setup), setup, inplace_form=False)
    Gp = o.pressuregradient(p, setup)
    pDv = (p * Dv)[g.Ip].sum()
    vGp = sum((v[..., a] * weights_u(setup, a) * Gp[..., a])[g.Iu[a]].sum() for a in range(D))
    assert np.isclose(pDv, -vGp, rtol=1e-10)


@pytest.mark.parametrize("D", [2, 3])
def test_convection_skew_symmetric(D):
    """test/operators.jl:107-128 `uCu ≈ 0 atol=1e-12` (u divergence free, BC applied)."""
    setup = make_setup(D, 16)
    g = setup.grid
    u = o.random_field(setup, rng=np.random.default_rng(3))
    c = o.convection(u, setup)
    uCu = sum((u[..., a] * weights_u(setup, a) * c[..., a])[g.Iu[a]].sum() for a in range(D))
    scale = sum((np.abs(u[..., a] * weights_u(setup, a) * c[..., a]))[g.Iu[a]].sum() for a in range(D))
    assert abs(uCu) < 1e-12 * max(scale, 1.0)


@pytest.mark.parametrize("D", [2, 3])
def test_diffusion_dissipative(D):
    """test/operators.jl:130-151 `uDu ≤ 0`."""
    setup = make_setup(D, 16)
    g = setup.grid
    u = o.random_field(setup, rng=np.random.default_rng(4))
    d = o.diffusion(u, setup)
    uDu = sum((u[..., a] * weights_u(setup, a) * d[..., a])[g.Iu[a]].sum() for a in range(D))
    assert uDu <= 0


@pytest.mark.parametrize("D", [2, 3])
@pytest.mark.parametrize("nonuniform", [False, True])
def test_fused_convection_diffusion(D, nonuniform):
    """test/operators.jl:153-160 `cd ≈ c + d`."""
    setup = make_setup(D, 16, nonuniform=nonuniform)
    rng = np.random.default_rng(5)
    u = o.apply_bc_u(np.asfortranarray(rng.standard_normal(setup.grid.N + (D,))), 0.0, setup)
    cd = o.convectiondiffusion(u, setup)
    assert np.allclose(cd, o.convection(u, setup) + o.diffusion(u, setup), rtol=1e-12, atol=1e-10)


@pytest.mark.parametrize("D", [2, 3])
def test_laplacian_negative_and_equals_div_grad(D):
    """test/operators.jl:90-105: negativity; and L = Omega * D * G (the identity the
    spectral solver diagonalises, src/pressure.jl:305-311)."""
    setup = make_setup(D, 16)
    g = setup.grid
    rng = np.random.default_rng(6)
    p = o.apply_bc_p(np.asfortranarray(rng.standard_normal(g.N)), 0.0, setup)
    Lp = o.laplacian(p, setup)
    assert (p * o.scalewithvolume(Lp, setup))[g.Ip].sum() <= 0
    G = o.apply_bc_u(o.pressuregradient(p, setup), 0.0, setup)
    WDG = o.scalewithvolume(o.divergence(G, setup), setup)
    assert np.allclose(Lp[g.Ip], WDG[g.Ip], rtol=1e-10, atol=1e-9)


def test_spectral_poisson_known_answer():
    """test/psolvers.jl:1-32: p_exact = (cos 2x + cos 2y)/4 on [0,2pi]^2, n = 32."""
    n = 32
    x = np.linspace(0, 2 * np.pi, n + 1)
    setup = o.Setup((x, x), Re=1e3)
    g = setup.grid
    X, Y = np.meshgrid(g.xp[0], g.xp[1], indexing="ij")
    p_exact = np.asfortranarray(0.25 * (np.cos(2 * X) + np.cos(2 * Y)))
    p_exact = o.apply_bc_p(p_exact, 0.0, setup)
    lap = o.laplacian(p_exact, setup)
    psolver = o.psolver_spectral(setup)
    p = o.apply_bc_p(o.poisson(psolver, lap), 0.0, setup)
    assert np.allclose(p[g.Ip], p_exact[g.Ip], rtol=np.sqrt(np.finfo(float).eps), atol=1e-12)


@pytest.mark.parametrize("D,T", [(2, np.float64), (3, np.float64), (2, np.float32), (3, np.float32)])
def test_poisson_identity_and_projection(D, T):
    """3-D / Float32 are unpinned by the reference: pin by identities.
    (i) laplacian(psolve(f)) == f - mean(f);  (ii) post-projection divergence ~ eps."""
    setup = make_setup(D, 16, T=T)
    g = setup.grid
    rng = np.random.default_rng(7)
    psolver = o.psolver_spectral(setup)
    f = np.asfortranarray(rng.standard_normal(g.N).astype(T))
    f[g.Ip] -= f[g.Ip].mean()
    p = o.apply_bc_p(psolver(f), 0.0, setup)
    tol = 2e-4 if T == np.float32 else 1e-11
    assert o.rel_l2(o.laplacian(p, setup)[g.Ip], f[g.Ip]) < tol
    u = o.apply_bc_u(np.asfortranarray(rng.standard_normal(g.N + (D,)).astype(T)), 0.0, setup)
    u = o.apply_bc_u(o.project(u, setup, psolver), 0.0, setup)
    assert u.dtype == T
    div = o.divergence(u, setup)[g.Ip]
    eps = np.finfo(T).eps
    assert np.abs(div).max() < 2000 * eps * np.abs(u).max() * 16


def test_taylor_green_2d_second_order():
    """examples/TaylorGreenVortex2D.jl:29-73: analytic solution, error ∝ n^-2 (Re=2000,
    Δt=0.01; shortened horizon t∈[0,0.5] keeps the CPU suite fast, the spatial error
    dominates either way)."""
    Re = 2.0e3

    def sol(t):
        return lambda dim, x, y: (-np.sin(x) * np.cos(y) if dim == 1 else np.cos(x) * np.sin(y)) * np.exp(-2 * t / Re) + 0 * x * y

    errs = []
    nlist = [8, 16, 32, 64]
    for n in nlist:
        x = np.linspace(0.0, 2 * np.pi, n + 1)
        setup = o.Setup((x, x), Re=Re)
        psolver = o.psolver_spectral(setup)
        ustart = o.velocityfield(setup, sol(0.0), 0.0, psolver=psolver)
        ut = o.velocityfield(setup, sol(0.5), 0.5, psolver=psolver, doproject=False)
        state, _ = o.solve_unsteady(setup, (0.0, 0.5), ustart, dt=0.01, psolver=psolver)
        Ip = setup.grid.Ip
        errs.append(np.sqrt(((state.u[Ip] - ut[Ip]) ** 2).sum()) / np.sqrt((ut[Ip] ** 2).sum()))
    rates = [np.log2(errs[i] / errs[i + 1]) for i in range(len(errs) - 1)]
    assert all(1.8 < r < 2.2 for r in rates), (errs, rates)


def test_rk44_tableau_shift():
    """methods.jl:231-236 + RKMethods.jl:515-521."""
    m = o.RK44()
    assert np.allclose(m.A, [[.5, 0, 0, 0], [0, .5, 0, 0], [0, 0, 1, 0], [1 / 6, 1 / 3, 1 / 3, 1 / 6]])
    assert np.allclose(m.c, [.5, .5, 1, 1])


def test_inviscid_energy_conservation_3d():
    """Skew-symmetric convection + exact projection: with Re → ∞ the RK44 energy error is
    O(Δt^4) -- the 'energy-conservation check' of BASELINE config 2 (TaylorGreenVortex3D)."""
    n = 16
    x = np.linspace(0.0, 1.0, n + 1)
    setup = o.Setup((x, x, x), Re=1e12)

    def U(dim, x, y, z):
        if dim == 1:
            return np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y) * np.sin(2 * np.pi * z) / 2
        if dim == 2:
            return -np.cos(2 * np.pi * x) * np.sin(2 * np.pi * y) * np.sin(2 * np.pi * z) / 2
        return 0 * x * y * z

    u = o.velocityfield(setup, U)
    E0 = o.total_kinetic_energy(u, setup)
    state, _ = o.solve_unsteady(setup, (0.0, 0.05), u, dt=0.005)
    E1 = o.total_kinetic_energy(state.u, setup)
    assert abs(E1 - E0) / E0 < 1e-8


def test_smagorinsky_dissipative_and_shape():
    """test/operators.jl:207 (`smagorinsky_closure(setup)(u, 0.1) isa Array`) + the physical
    property that the eddy-viscosity term removes energy."""
    for D in (2, 3):
        setup = make_setup(D, 16)
        g = setup.grid
        u = o.random_field(setup, rng=np.random.default_rng(8))
        s = o.smagorinsky_closure(setup)(u, 0.17)
        assert s.shape == u.shape and np.isfinite(s).all()
        usu = sum((u[..., a] * weights_u(setup, a) * s[..., a])[g.Iu[a]].sum() for a in range(D))
        assert usu < 0


def test_cfl_and_bodyforce():
    """solver.jl:101-125; setup.jl:25-32 (Kolmogorov force 5 sinpi(8y), Kolmogorov2D.jl:29-36)."""
    n = 32
    x = np.linspace(0.0, 1.0, n + 1)
    setup = o.Setup((x, x), Re=2e3, bodyforce=lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x)
    g = setup.grid
    assert setup.issteadybodyforce and setup.bodyforce.shape == g.N + (2,)
    assert np.allclose(setup.bodyforce[3, :, 0], 5 * np.sin(8 * np.pi * g.xu[0][1]))
    assert np.all(setup.bodyforce[..., 1] == 0)
    u = o.random_field(setup, A=1e-2, rng=np.random.default_rng(9))
    dt = o.get_cfl_timestep(u, setup)
    h = 1.0 / n
    expect = min(2e3 * h * h / 2, (h / np.abs(u[g.Ip])).min())
    assert np.isclose(dt, expect)
