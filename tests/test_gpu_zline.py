"""GPU parity of the transpose-free last-axis pressure solve (csrc/zline.cuh, the default of a slab-decomposed grid) on ONE
GPU, where `set_param("zline", 1)` swaps it in for k_zsolve: the sub-chunks of the slab then play the ranks' part (same
kernels, same exchange areas, same chain; only the peer pointers are local).  Against the oracle's psolver_spectral
(src/pressure.jl:323-344) and against the library's own k_zsolve path.  Multi-GPU: tests/mgpu_parity.py."""
import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import compare, make_pair, rand_p

pytestmark = pytest.mark.gpu


def set_param(sc, k, v):
    sc.ctx.call("insb200_set_param", k.encode(), float(v))


@pytest.mark.parametrize("n,T,chunks", [
    ((32, 16, 64), np.float64, 1),
    ((32, 16, 64), np.float64, 4),
    ((24, 20, 48), np.float64, 3),      # not powers of two (k_zsolve does not take these; the line solve does)
    ((64, 32, 128), np.float32, 0),
    ((64, 32, 128), np.float32, 8),
    ((128, 128, 256), np.float32, 0),
], ids=lambda v: str(v).replace(" ", ""))
def test_line_solve_matches_oracle_poisson(n, T, chunks):
    import insb200 as ins

    so, sc = make_pair(n, T)
    f = rand_p(so, 3, T)
    f[so.grid.Ip] -= f[so.grid.Ip].mean()
    want = o.poisson(o.psolver_spectral(so), f)
    ref = ins.to_host(ins.poisson(ins.psolver_spectral(sc), ins.to_device(f, sc)))
    set_param(sc, "zline", 1)
    set_param(sc, "zline_chunks", chunks)
    got = ins.to_host(ins.poisson(ins.psolver_spectral(sc), ins.to_device(f, sc)))
    again = ins.to_host(ins.poisson(ins.psolver_spectral(sc), ins.to_device(f, sc)))  # the second exchange area
    tol = 1e-12 if T == np.float64 else 1e-5
    Ip = so.grid.Ip
    compare("line solve vs oracle", got[Ip], want[Ip], tol)
    compare("line solve vs k_zsolve path", got[Ip], ref[Ip], tol)
    assert np.array_equal(got[Ip], again[Ip])


@pytest.mark.parametrize("T", [np.float32, np.float64], ids=["f32", "f64"])
def test_line_solve_in_rk_steps(T):
    """RK44 steps with the line solve in the projection: same bar as every other path (north_star: 1e-5 / 1e-12)."""
    import insb200 as ins
    from insb200.api import _rk_steps

    n = (64, 32, 64)
    so, sc = make_pair(n, T, force=True)
    set_param(sc, "zline", 1)
    set_param(sc, "zline_chunks", 2)
    u0 = o.random_field(so, kp=4, rng=np.random.default_rng(5))
    m = ins.RKMethods.RK44(T=T)
    nsteps, dt = 5, 1e-3
    u = ins.to_device(u0, sc)
    _rk_steps(m, sc, u, 0.0, dt, nsteps, None)
    mo = o.RK44(T)
    st = o.create_stepper(mo, so, o.psolver_spectral(so), u0.copy(order="F"))
    for _ in range(nsteps):
        st = o.timestep(mo, st, T(dt))
    compare("u after RK44 steps with the line solve", ins.to_host(u), st.u, 1e-12 if T == np.float64 else 1e-5)
    div = ins.to_host(ins.divergence(u, sc))[so.grid.Ip]
    assert np.abs(div).max() / max(n) < 100 * np.finfo(T).eps * np.abs(st.u).max()
