"""Host-side pieces of bench.py that need no GPU: the weak-scaling grids, the time step of the throughput runs (inside
RK44's stability interval for the stiffest diffusion mode at every N -- the 8-GPU TaylorGreenVortex grid was not, with
half the reference's per-axis CFL bound alone; profiles/r2_bench_19_8gpu_note.txt) and the word model of the kernels."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


@pytest.mark.parametrize("name", sorted(bench.WORKLOADS))
@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_time_step_is_inside_the_rk44_stability_interval(name, world):
    wl = bench.WORKLOADS[name]
    n = bench.weak_grid(wl["D"], wl["n"], world)
    assert int(np.prod([float(m) for m in n])) == wl["n"] ** wl["D"] * world
    # the reference's per-axis diffusive bound (solver.jl:113), halved as bench.py does
    dt_cfl = 0.5 * min(wl["Re"] / float(m) ** 2 / 2 for m in n)
    dt = bench.stable_dt(wl, n, wl["n"], dt_cfl)
    z = dt * sum(4.0 * float(m) ** 2 for m in n) / wl["Re"]
    R = 1 - z + z ** 2 / 2 - z ** 3 / 6 + z ** 4 / 24
    assert z <= 0.8 * 2.785 * (1 + 1e-12) and abs(R) < 1.0, (n, dt, z, R)
    assert dt <= wl["dt"] and dt <= dt_cfl


def test_half_the_per_axis_bound_alone_is_unstable_in_3d():
    """what the cap is for: nu dt 4/dx^2 = 1 per axis -> z = -3 in 3-D, |R(z)| = 1.375"""
    z = 3.0
    assert abs(1 - z + z ** 2 / 2 - z ** 3 / 6 + z ** 4 / 24) > 1.3


def test_kernel_word_model_has_the_line_solve_passes():
    w = bench.kernel_words(3, False, False, tma=True, zsolve=True)
    assert w["poisson_forward"] == 2.0 and w["poisson_scale"] == 2.0 and w["cufft_forward"] == 4.0


def test_a_non_finite_run_is_measured_again_once_with_the_other_solver():
    import types

    calls = []

    def run_factory(fail_times):
        def run():
            calls.append(list(args.param))
            if len(calls) <= fail_times:
                raise bench.BlowUp("tgv3d: non-finite u")
            return {"ok": True}
        return run

    args = types.SimpleNamespace(param=["fft_pipe=4"])
    rem = {}
    assert bench.measure_with_fallback(run_factory(1), args, 8, 1, rem, "tgv3d") == {"ok": True}
    assert calls == [["fft_pipe=4"], ["fft_pipe=4", "zline=0"]] and rem["tgv3d"]["with"].startswith("zline=0")
    calls.clear()
    with pytest.raises(SystemExit):  # fails with both solvers
        bench.measure_with_fallback(run_factory(2), types.SimpleNamespace(param=[]), 8, 1, {}, "tgv3d")
    calls.clear()
    with pytest.raises(SystemExit):  # one GPU: nothing to fall back to
        bench.measure_with_fallback(run_factory(1), types.SimpleNamespace(param=[]), 1, 0, {}, "tgv3d")
    calls.clear()
    with pytest.raises(SystemExit):  # the caller chose the solver
        bench.measure_with_fallback(run_factory(1), types.SimpleNamespace(param=["zline = 0"]), 8, 0, {}, "tgv3d")
