"""BASELINE configs 1 and 2 at their STATED sizes against the oracle ("what's missing" 3).

Config 1: DecayingTurbulence2D periodic 256^2 Float64, Re = 4e3, random_field(kp = 10), tlims (0, 1), adaptive dt with
cfl = 0.9 (examples/DecayingTurbulence2D.jl:22-29, solver.jl:52-69): 805 RK44 steps.  The oracle needs 100 s for that run,
so its result is a committed fixture (tests/golden/config1_256sq_f64.npz, written by tests/golden/make_config1_fixture.py),
which also records how much the flow itself amplifies a one-ulp perturbation of the start field over the run
(rel-L2 1.8e-14): the bar below is north_star's 1e-12.
Config 2: TaylorGreenVortex3D periodic 128^3 Float64, Re = 1e3, dt = 1e-3 (examples/TaylorGreenVortex3D.jl:26-49): the
oracle is stepped here, 3 steps, u and the kinetic-energy history to 1e-12."""
import os

import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import compare

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "config1_256sq_f64.npz")


def test_config1_decaying_turbulence_256sq_f64_adaptive_full_run():
    import insb200 as ins

    g = np.load(GOLD)
    T = np.float64
    x = tuple(np.linspace(0.0, 1.0, 257) for _ in range(2))
    so = o.Setup(x, Re=T(4e3))
    sc = ins.Setup(x, Re=T(4e3))
    u0 = np.asfortranarray(g["u0"])
    # the fixture's start field is what the oracle generates today (same image on the GPU box)
    assert np.array_equal(u0, o.random_field(so, kp=10, A=1, rng=np.random.default_rng(5)))
    nsteps = int(g["nsteps"])
    assert float(g["roundoff_growth_u"]) < 1e-13
    # (a) the whole run inside ONE C call on a host buffer (insb200_solve_unsteady_host: adaptive loop on the device side)
    uh = u0.copy(order="F")
    ns, t, ke = ins.solve_unsteady_host(sc, (0.0, 1.0), uh, cfl=0.9, ke_hist=nsteps + 1)
    assert ns == nsteps and float(t) == float(g["t_end"])
    err_a = compare("config 1, host entry: u(t = 1)", uh, g["u_end"], 1e-12)
    assert np.allclose(ke, g["ke"], rtol=1e-12, atol=0)
    # (b) the reference's own loop: solve_unsteady of the host mirror, one timestep! + get_cfl_timestep! per step
    state, out = ins.solve_unsteady(sc, (0.0, 1.0), ins.to_device(u0, sc), cfl=0.9, processors={"ke": ins.energy_history(sc)})
    assert state.n == nsteps
    err_b = compare("config 1, solve_unsteady: u(t = 1)", ins.to_host(state.u), g["u_end"], 1e-12)
    assert np.allclose(out["ke"][:, 1], g["ke"], rtol=1e-12, atol=0)
    div = ins.to_host(ins.divergence(state.u, sc))[so.grid.Ip]
    assert np.abs(div).max() / 256 < 100 * np.finfo(T).eps * np.abs(g["u_end"]).max()
    print(f"\n  config 1 (805 adaptive steps): rel-L2(u) = {err_a:.2e} (one C call), {err_b:.2e} (per-step loop); bar 1e-12, "
          f"round-off amplification of the flow {float(g['roundoff_growth_u']):.1e}")


def tgv3(dim, x, y, z):
    """examples/TaylorGreenVortex3D.jl:34-41"""
    if dim == 1:
        return np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y) * np.sin(2 * np.pi * z) / 2
    if dim == 2:
        return -np.cos(2 * np.pi * x) * np.sin(2 * np.pi * y) * np.sin(2 * np.pi * z) / 2
    return 0 * x * y * z


@pytest.mark.parametrize("exact", [0, 1], ids=["fast", "exact"])
def test_config2_taylor_green_128cube_f64(exact):
    import insb200 as ins

    T = np.float64
    x = tuple(np.linspace(0.0, 1.0, 129) for _ in range(3))
    so = o.Setup(x, Re=T(1e3))
    sc = ins.Setup(x, Re=T(1e3))
    sc.ctx.call("insb200_set_param", b"exact", float(exact))
    u0 = o.velocityfield(so, tgv3)
    compare("velocityfield at 128^3", ins.to_host(ins.velocityfield(sc, tgv3)), u0, 1e-12)
    ke_fn = lambda s: float(o.total_kinetic_energy(s.u, so))  # noqa: E731
    state_o, out_o = o.solve_unsteady(so, (0.0, 0.003), u0, dt=1e-3, processors={"ke": ke_fn})
    state_c, out_c = ins.solve_unsteady(sc, (0.0, 0.003), ins.to_device(u0, sc), dt=1e-3, processors={"ke": ins.energy_history(sc)})
    assert state_c.n == state_o.n == 3
    err = compare("config 2: u after 3 steps at 128^3", ins.to_host(state_c.u), state_o.u, 1e-12)
    assert np.allclose(out_c["ke"][:, 1], out_o["ke"], rtol=1e-12, atol=0)
    p_c = ins.to_host(ins.pressure(state_c.u, None, state_c.t, sc))[so.grid.Ip]
    p_o = o.pressure(state_o.u, None, state_o.t, so, o.psolver_spectral(so))[so.grid.Ip]
    errp = compare("config 2: pressure!(u) at 128^3", p_c, p_o, 1e-12)
    print(f"\n  config 2 ({'exact' if exact else 'fast'} kernels): rel-L2 u {err:.2e}, p {errp:.2e} (bar 1e-12)")
