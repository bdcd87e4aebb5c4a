"""GPU parity, operator by operator: the CUDA kernels (called through the C ABI via the host
mirror) against the oracle on identical inputs.

Bit-exact cases: unit box with n = 2^m per axis, so 1/h is exact and -- with -fmad=false -- the
stencil kernels perform exactly the reference's IEEE operations.  Other grids and anything behind
an FFT (cuFFT vs pocketfft) are compared with the tolerances of BASELINE.json's north_star:
rel-L2 <= 1e-12 (Float64), <= 1e-5 (Float32).
"""
import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import compare, make_pair, rand_p, rand_u

pytestmark = pytest.mark.gpu

TOL = {np.float64: 1e-12, np.float32: 1e-5}
GRIDS = [((16, 32), True), ((24, 20), False), ((16, 8, 32), True), ((12, 20, 10), False)]
CASES = [(n, ex, T) for (n, ex) in GRIDS for T in (np.float64, np.float32)]


def ids(c):
    return f"{'x'.join(map(str, c[0]))}-{c[2].__name__}"


@pytest.fixture(params=CASES, ids=ids)
def case(request):
    import insb200 as ins

    n, exact, T = request.param
    so, sc = make_pair(n, T, force=True)
    return ins, n, exact, T, so, sc


def test_apply_bc(case):
    ins, n, exact, T, so, sc = case
    rng = np.random.default_rng(1)
    u = np.asfortranarray(rng.standard_normal(so.grid.N + (len(n),)).astype(T))
    p = np.asfortranarray(rng.standard_normal(so.grid.N).astype(T))
    compare("apply_bc_u", ins.to_host(ins.apply_bc_u_(ins.to_device(u, sc), 0.0, sc)), o.apply_bc_u(u, 0.0, so), 0, exact=True)
    compare("apply_bc_p", ins.to_host(ins.apply_bc_p_(ins.to_device(p, sc), 0.0, sc)), o.apply_bc_p(p, 0.0, so), 0, exact=True)


def test_momentum_and_terms(case):
    ins, n, exact, T, so, sc = case
    u = rand_u(so, 2, T)
    ud = ins.to_device(u, sc)
    tol = 20 * np.finfo(T).eps
    compare("momentum", ins.to_host(ins.momentum(ud, None, 0.0, sc)), o.momentum(u, None, 0.0, so), tol, exact)
    compare("convection", ins.to_host(ins.convection(ud, sc)), o.convection(u, so), tol, exact)
    compare("diffusion", ins.to_host(ins.diffusion(ud, sc)), o.diffusion(u, so), tol, exact)
    compare("convectiondiffusion", ins.to_host(ins.convectiondiffusion(ud, sc)), o.convectiondiffusion(u, so), tol, exact)


def test_divergence_gradient(case):
    ins, n, exact, T, so, sc = case
    u, p = rand_u(so, 3, T), rand_p(so, 4, T)
    ud, pd = ins.to_device(u, sc), ins.to_device(p, sc)
    tol = 20 * np.finfo(T).eps
    Ip = so.grid.Ip
    compare("divergence", ins.to_host(ins.divergence(ud, sc))[Ip], o.divergence(u, so)[Ip], tol, exact)
    compare("pressuregradient", ins.to_host(ins.pressuregradient(pd, sc)), o.pressuregradient(p, so), tol, exact)
    compare("applypressure", ins.to_host(ins.applypressure(ud, pd, sc)), o.applypressure(u, p, so), tol, exact)
    compare("scalewithvolume", ins.to_host(ins.scalewithvolume_(pd.clone(), sc)), o.scalewithvolume(p, so), tol, exact)


def test_smagorinsky(case):
    ins, n, exact, T, so, sc = case
    u = rand_u(so, 5, T)
    want = o.smagorinsky_closure(so)(u, 0.17)
    got = ins.to_host(ins.smagorinsky_closure(sc)(ins.to_device(u, sc), 0.17))
    compare("smagorinsky", got, want, 50 * np.finfo(T).eps, exact)


def test_reductions(case):
    ins, n, exact, T, so, sc = case
    u = rand_u(so, 6, T)
    ud = ins.to_device(u, sc)
    E, Eo = ins.total_kinetic_energy(ud, sc), float(o.total_kinetic_energy(u.astype(np.float64), so))
    assert abs(E - Eo) <= (1e-12 if T == np.float64 else 2e-6) * abs(Eo), (E, Eo)
    assert ins.maxabs(ud, sc) == float(np.abs(u).max())
    dt, dto = ins.get_cfl_timestep_(None, ud, sc), o.get_cfl_timestep(u, so)
    assert dt == dto if exact else np.isclose(dt, dto, rtol=10 * np.finfo(T).eps), (dt, dto)


def test_psolve_project_pressure(case):
    """Everything behind the FFT: tolerance of north_star, plus the identities that pin the
    3-D / Float32 solves the reference's tests leave unpinned."""
    ins, n, exact, T, so, sc = case
    tol = TOL[T]  # north_star's: measured 2e-7 ... 1.3e-6 (Float32), 3e-16 ... 3.5e-15 (Float64), profiles/r2_tolerances.txt
    Ip = so.grid.Ip
    psol_o = o.psolver_spectral(so)
    f = rand_p(so, 7, T)
    f[Ip] -= f[Ip].mean()
    fd = ins.to_device(f, sc)
    got = ins.to_host(ins.psolver_spectral(sc)(fd))
    compare("psolve (random right-hand side)", got[Ip], psol_o(f)[Ip], tol)
    u = rand_u(so, 8, T)
    uo, po = o.project(u, so, psol_o, return_p=True)
    ud, pd = ins.to_device(u, sc), ins.scalarfield(sc)
    ins.project_(ud, sc, p=pd)
    compare("project.u (random non-solenoidal u)", ins.to_host(ud)[Ip], uo[Ip], tol)
    compare("project.p (random non-solenoidal u)", ins.to_host(pd), po, tol)
    ins.apply_bc_u_(ud, 0.0, sc)
    div = ins.to_host(ins.divergence(ud, sc))[Ip]
    hmin = 1.0 / max(n)
    assert np.abs(div).max() * hmin < 200 * np.finfo(T).eps * np.abs(u).max(), np.abs(div).max()
    compare("pressure", ins.to_host(ins.pressure(ins.to_device(u, sc), None, 0.0, sc)), o.pressure(u, None, 0.0, so, psol_o), tol)


def test_poisson_known_answer_gpu():
    """test/psolvers.jl:1-32 on the CUDA path."""
    import insb200 as ins

    n = 32
    x = np.linspace(0, 2 * np.pi, n + 1)
    so = o.Setup((x, x), Re=1e3)
    sc = ins.Setup((x, x), Re=1e3)
    X, Y = np.meshgrid(so.grid.xp[0], so.grid.xp[1], indexing="ij")
    p_exact = o.apply_bc_p(np.asfortranarray(0.25 * (np.cos(2 * X) + np.cos(2 * Y))), 0.0, so)
    lap = o.laplacian(p_exact, so)
    p = ins.to_host(ins.apply_bc_p_(ins.poisson(ins.psolver_spectral(sc), ins.to_device(lap, sc)), 0.0, sc))
    assert np.allclose(p[so.grid.Ip], p_exact[so.grid.Ip], rtol=np.sqrt(np.finfo(float).eps), atol=1e-12)


def test_operator_properties_gpu():
    """test/operators.jl:58-88 (D = -G') and :107-128 (skew-symmetry) evaluated with the CUDA
    operators themselves at 64^3 Float64."""
    import insb200 as ins

    n = (64, 64, 64)
    so, sc = make_pair(n, np.float64)
    h = 1.0 / 64
    Ip = so.grid.Ip
    v, p = rand_u(so, 9, np.float64), rand_p(so, 10, np.float64)
    Dv = ins.to_host(ins.scalewithvolume_(ins.divergence(ins.to_device(v, sc), sc), sc))
    Gp = ins.to_host(ins.pressuregradient(ins.to_device(p, sc), sc))
    pDv, vGp = (p * Dv)[Ip].sum(), (v * Gp)[Ip].sum() * h ** 3
    assert np.isclose(pDv, -vGp, rtol=1e-10)
    u = ins.random_field(sc, rng=np.random.default_rng(11))
    c = ins.to_host(ins.convection(u, sc))
    uh = ins.to_host(u)
    assert abs((uh * c)[Ip].sum()) < 1e-12 * np.abs(uh * c)[Ip].sum()
    d = ins.to_host(ins.diffusion(u, sc))
    assert (uh * d)[Ip].sum() <= 0


@pytest.mark.parametrize("n,T", [((32, 48), np.float64), ((16, 32, 16), np.float32)], ids=["2d_f64", "3d_f32"])
def test_random_field_on_device_matches_host_generation(n, T):
    """random_field (initializers.jl:189-219) with the spectrum and the inverse FFT on the device against the
    host generation (same random stream) and against the oracle's."""
    import insb200 as ins

    so, sc = make_pair(n, T)
    ud = ins.to_host(ins.random_field(sc, kp=4, rng=np.random.default_rng(5), ondevice=True))
    uh = ins.to_host(ins.random_field(sc, kp=4, rng=np.random.default_rng(5)))
    uo = o.random_field(so, kp=4, rng=np.random.default_rng(5))
    tol = TOL[T]
    compare("random_field: device vs host generation", ud, uh, tol)
    compare("random_field: device generation vs oracle", ud, uo, tol)
    div = ins.to_host(ins.divergence(ins.to_device(ud, sc), sc))[so.grid.Ip]
    assert np.abs(div).max() < (1e-3 if T == np.float32 else 1e-10)


def test_error_behaviour():
    """assert_uniform_periodic (src/utils.jl:1-13) and argument validation of the C ABI."""
    import insb200 as ins

    with pytest.raises(AssertionError):
        ins.Setup((np.linspace(0, 1, 16), np.linspace(0, 1, 17)))  # odd n
    x = np.linspace(0, 1, 17) ** 2
    with pytest.raises(AssertionError):
        ins.Setup((x, x))  # non-uniform
    ax = np.linspace(0, 1, 17)
    s = ins.Setup((ax, ax))
    with pytest.raises(ValueError):
        ins.momentum(ins.scalarfield(s), None, 0.0, s)  # wrong shape
    with pytest.raises(ins._lib.INSB200Error):
        s.ctx.call("insb200_convectiondiffusion", None, None, 3)


def ins_mod():
    import insb200 as ins

    return ins


@pytest.mark.parametrize("T", [np.float64, np.float32], ids=["f64", "f32"])
@pytest.mark.parametrize("n", [(16, 16, 8), (32, 16, 16), (16, 12, 32), (32, 32, 64), (8, 8, 128), (8, 16, 256),
                               (16, 8, 512), (8, 8, 1024), (8, 8, 2048)], ids=str)
def test_zsolve_matches_cufft_and_oracle(n, T):
    """psolver_spectral with the fused last-axis kernel (forward z FFT + symbol + inverse z FFT, radix 8/4/2)
    against the all-cuFFT path and the oracle."""
    so, sc = make_pair(n, T)
    f = rand_p(so, 5, T)
    f[so.grid.Ip] -= f[so.grid.Ip].mean(dtype=np.float64).astype(T)
    want = o.psolver_spectral(so)(f)[so.grid.Ip]
    got = {}
    for z in (1, 0):
        sc.ctx.call("insb200_set_param", b"zsolve", float(z))
        got[z] = ins_mod().to_host(ins_mod().psolver_spectral(sc)(ins_mod().to_device(f, sc)))[so.grid.Ip]
    tol = 1e-12 if T == np.float64 else 1e-5  # north_star's; measured 4e-16 / 3e-7 (profiles/r2_tolerances.txt)
    compare("psolver_spectral: k_zsolve path vs oracle", got[1], want, tol)
    compare("psolver_spectral: cuFFT path vs oracle", got[0], want, tol)
    compare("psolver_spectral: k_zsolve vs cuFFT path", got[1], got[0], tol)
