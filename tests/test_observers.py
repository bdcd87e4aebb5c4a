"""Observer fields either side of the stepper: vorticity!, interpolate_u_p!, interpolate_w_p!,
the velocity norm of observefield and observespectrum.  CPU part pins the oracle restatements with known
answers; GPU part compares the library (C ABI) and the device-side processors with the oracle."""
import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import compare, make_pair


def osetup(n, T=np.float64):
    return o.Setup(tuple(np.linspace(0.0, 1.0, m + 1).astype(T) for m in n), Re=T(1000))


def test_oracle_vorticity_of_a_shear_layer_and_interpolations():
    so = osetup((64, 32))
    g = so.grid
    u = o.vectorfield(so)
    # u1 = sin(2 pi y) sampled at the u1 points (x faces, y centres): vorticity = -du1/dy = -2 pi cos(2 pi y) at the corners
    u[..., 0] = np.sin(2 * np.pi * g.xp[1])[None, :]
    u = o.apply_bc_u(u, 0.0, so)
    w = o.vorticity(u, so)
    ycorner = g.x[1][1:]  # corner I sits at the upper right of cell I
    exact = -2 * np.pi * np.cos(2 * np.pi * ycorner[: g.N[1] - 1])
    assert np.abs(w[1:-1, : g.N[1] - 1] - exact[None, :]).max() < 2 * np.pi * (2 * np.pi / 32) ** 2
    wp = o.interpolate_w_p(w, so)
    assert np.abs(wp[g.Ip] + 2 * np.pi * np.cos(2 * np.pi * g.xp[1][1:-1])[None, :]).max() < 0.1
    up = o.interpolate_u_p(u, so)
    assert np.array_equal(up[g.Ip + (0,)], u[g.Ip + (0,)])  # u1 does not depend on x
    assert np.allclose(o.velocitynorm(up, so)[g.Ip], np.abs(up[g.Ip + (0,)]))


@pytest.mark.parametrize("n", [(32, 32), (16, 16, 32)], ids=["2d", "3d"])
def test_oracle_spectrum_of_a_single_mode(n):
    so = osetup(n)
    g = so.grid
    D = len(n)
    k0 = 3
    u = o.vectorfield(so)
    shape = [1] * D
    shape[0] = -1
    u[..., 1] = np.cos(2 * np.pi * k0 * g.xp[0]).reshape(shape)  # u2 = cos(2 pi k0 x)
    ehat, kappa = o.observespectrum(u, so)
    i0 = int(np.where(kappa == k0)[0][0])
    # |uhat(k0)|^2 = (N/2)^2, only the +k0 mode lies in the low-mode block: e = 1/8
    assert abs(ehat[i0] - 0.125) < 1e-12
    if D == 3:  # linear shells are disjoint: nothing anywhere else
        assert np.abs(np.delete(ehat, i0)).max() < 1e-20


def test_mirror_spectral_stuff_matches_oracle():
    import insb200 as ins

    for n, T in (((32, 48), np.float32), ((16, 32, 16), np.float64)):
        so = osetup(n, T)
        st = o.spectral_stuff(so, npoint=40)
        sc = type("S", (), {"grid": so.grid})()
        inds, kappa, K = ins.spectral_stuff(sc, npoint=40)
        assert K == st.K and np.array_equal(kappa, st.kappa)
        assert all(np.array_equal(np.sort(a), np.sort(b)) for a, b in zip(inds, st.inds))


@pytest.mark.gpu
@pytest.mark.parametrize("n,T", [((32, 16), np.float64), ((64, 32), np.float32), ((16, 32, 16), np.float64), ((32, 16, 16), np.float32)],
                         ids=lambda v: str(v).replace(" ", ""))
def test_observer_operators_match_oracle_bitwise(n, T):
    import insb200 as ins

    so, sc = make_pair(n, T)
    u = o.apply_bc_u(o.random_field(so, kp=3, rng=np.random.default_rng(2)), 0.0, so)
    ud = ins.to_device(u, sc)
    g = so.grid
    lo = tuple(slice(0, m - 1) for m in g.N)
    w_o = o.vorticity(u, so)
    w_c = ins.to_host(ins.vorticity(ud, sc))
    compare("vorticity", w_c[lo], w_o[lo], 0, exact=True)      # power-of-two spacing: bit for bit
    wp_c = ins.to_host(ins.interpolate_w_p(ins.to_device(w_o, sc), sc))
    compare("interpolate_w_p", wp_c[g.Ip], o.interpolate_w_p(w_o, so)[g.Ip], 0, exact=True)
    up_o = o.interpolate_u_p(u, so)
    upd = ins.interpolate_u_p(ud, sc)
    compare("interpolate_u_p", ins.to_host(upd)[g.Ip], up_o[g.Ip], 0, exact=True)
    nrm = ins.to_host(ins.velocitynorm_(ins.scalarfield(sc), upd, sc))
    compare("velocitynorm", nrm[g.Ip], o.velocitynorm(up_o, so)[g.Ip], 1e-6 if T == np.float32 else 1e-15)


@pytest.mark.gpu
@pytest.mark.parametrize("n,T", [((32, 32), np.float64), ((32, 16, 16), np.float32)], ids=["2d_f64", "3d_f32"])
def test_device_processors_match_oracle(n, T):
    import insb200 as ins

    so, sc = make_pair(n, T)
    u0 = o.random_field(so, kp=3, rng=np.random.default_rng(8))
    procs = {"w": ins.observefield(sc, "vorticity", nupdate=2), "un": ins.observefield(sc, "velocitynorm", nupdate=2),
             "u2": ins.observefield(sc, 2, nupdate=2), "spec": ins.observespectrum(sc, npoint=20, nupdate=2)}
    state, out = ins.solve_unsteady(sc, (0.0, 0.004), ins.to_device(u0, sc), dt=1e-3, processors=procs)
    # the same four steps on the oracle
    m = o.RK44(T)
    st = o.create_stepper(m, so, o.psolver_spectral(so), u0.copy(order="F"))
    for _ in range(4):
        st = o.timestep(m, st, 1e-3)
    tol = 1e-5 if T == np.float32 else 1e-12
    assert len(out["w"]) == 2 and len(out["un"]) == 2
    compare("observed vorticity", out["w"][-1], o.observe_vorticity(st.u, st.t, so), tol)  # measured 2e-16 / 1.4e-7 (profiles/r2_tolerances.txt)
    up = o.interpolate_u_p(st.u, so)
    compare("observed velocity norm", out["un"][-1], o.velocitynorm(up, so)[so.grid.Ip], tol)
    compare("observed u2", out["u2"][-1], up[so.grid.Ip + (1,)], tol)
    ehat, kappa = out["spec"]
    ehat_o, kappa_o = o.observespectrum(st.u, so, npoint=20)
    assert np.array_equal(kappa, kappa_o) and ehat.shape == (2, len(kappa))
    assert np.allclose(ehat[-1], ehat_o, rtol=1e-4 if T == np.float32 else 1e-10, atol=0)
