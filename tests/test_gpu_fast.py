"""GPU parity of the FAST kernels (csrc/fast.cu: k_stage3d_fast, the fused closure + stage kernel that is the default
of the 3-D step) against (a) the bit-exact kernels of the same library (`set_param("exact", 1)`) and (b) the oracle.

The fast kernel evaluates the reference's operators in flux form with FMA contraction (tests/test_fast_formulas.py
restates the arithmetic): it is not bit-identical, the bar is north_star's -- relative L2 of u and p after N steps
<= 1e-5 (Float32) / <= 1e-12 (Float64), divergence at machine precision, kinetic-energy history to the same tolerance.
"""
import ctypes as C

import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import compare, make_pair

pytestmark = pytest.mark.gpu


def set_param(sc, k, v):
    sc.ctx.call("insb200_set_param", k.encode(), float(v))


@pytest.mark.parametrize("n,T,force,smag,zc", [
    ((64, 32, 40), np.float32, True, True, 0),
    ((64, 32, 40), np.float32, True, True, 7),
    ((32, 16, 12), np.float32, False, True, 2),
    ((64, 32, 40), np.float32, False, False, 0),
    ((32, 48, 10), np.float32, True, False, 3),
    ((64, 32, 24), np.float64, True, False, 5),
], ids=lambda v: str(v).replace(" ", ""))
def test_fast_stage_matches_exact_and_oracle(n, T, force, smag, zc):
    import insb200 as ins
    from insb200.api import _rk_steps

    so, sc = make_pair(n, T, force=force, smag=smag)
    theta = 0.17 if smag else None
    u0 = o.random_field(so, kp=4, rng=np.random.default_rng(5))
    m = ins.RKMethods.RK44(T=T)
    nsteps, dt = 6, 1e-3
    out = {}
    for mode in (1, 0):
        set_param(sc, "exact", mode)
        set_param(sc, "zchunk", zc)
        u = ins.to_device(u0, sc)
        l0 = sc.ctx.lib.insb200_launch_count(sc.ctx.h)
        _, ke = _rk_steps(m, sc, u, 0.0, dt, nsteps, theta, ke_hist=True)
        out[mode] = (ins.to_host(u), ke, sc.ctx.lib.insb200_launch_count(sc.ctx.h) - l0,
                     ins.to_host(ins.divergence(u, sc))[so.grid.Ip])
    if smag:  # the closure pass is gone: one launch less per stage
        assert out[0][2] == out[1][2] - 4 * nsteps, (out[0][2], out[1][2])
    mo = o.RK44(T)
    st = o.create_stepper(mo, so, o.psolver_spectral(so), u0.copy(order="F"))
    ke_o = [o.total_kinetic_energy(u0, so)]
    for _ in range(nsteps):
        st = o.timestep(mo, st, T(dt), theta=theta)
        ke_o.append(o.total_kinetic_energy(st.u, so))
    tol = 1e-5 if T == np.float32 else 1e-12
    r_exact = compare("fast vs exact kernels", out[0][0], out[1][0], tol)
    r_oracle = compare("fast vs oracle", out[0][0], st.u, tol)
    print(f"\n  n={n} {np.dtype(T).name} smag={smag}: rel-L2 fast/exact {r_exact:.2e}, fast/oracle {r_oracle:.2e} (bar {tol:g})")
    assert np.allclose(out[0][1], ke_o, rtol=tol, atol=0)
    assert np.abs(out[0][3]).max() / max(n) < 100 * np.finfo(T).eps * np.abs(st.u).max()


def test_fast_stage_other_tableaux_and_long_rows():
    """Rows with more than 3 earlier k_j take the exact kernels inside a fast run; results stay within tolerance."""
    import insb200 as ins
    from insb200.api import _rk_steps

    T, n = np.float32, (32, 32, 16)
    so, sc = make_pair(n, T, smag=True)
    u0 = o.random_field(so, kp=4, rng=np.random.default_rng(9))
    for name in ("SSP33", "RK56", "RK33C2"):
        mc, mo = getattr(ins.RKMethods, name)(T=T), o.explicit_method(name, T)
        u = ins.to_device(u0, sc)
        _rk_steps(mc, sc, u, 0.0, 1e-3, 3, 0.17)
        st = o.create_stepper(mo, so, o.psolver_spectral(so), u0.copy(order="F"))
        for _ in range(3):
            st = o.timestep(mo, st, T(1e-3), theta=0.17)
        compare(name, ins.to_host(u), st.u, 1e-5)
