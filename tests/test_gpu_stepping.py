"""GPU parity of the time stepper: `timestep!` / `solve_unsteady` on the CUDA path against the
oracle, identical initial arrays (never "same seed").

Tolerances are north_star's: relative L2 of u and p after N steps <= 1e-12 (Float64) / <= 1e-5
(Float32), post-projection divergence at machine precision, kinetic-energy history to the same
tolerance.  Float32 runs are compared with the Float32 oracle.
"""
import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import compare, make_pair

pytestmark = pytest.mark.gpu
TOL = {np.float64: 1e-12, np.float32: 1e-5}


def tgv3(dim, x, y, z):
    """examples/TaylorGreenVortex3D.jl:34-41"""
    if dim == 1:
        return np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y) * np.sin(2 * np.pi * z) / 2
    if dim == 2:
        return -np.cos(2 * np.pi * x) * np.sin(2 * np.pi * y) * np.sin(2 * np.pi * z) / 2
    return 0 * x * y * z


def run_pair(n, T, nsteps, dt, Re=1000.0, force=False, smag=False, A=1.0, theta=None, seed=0):
    import insb200 as ins

    so, sc = make_pair(n, T, Re=Re, force=force, smag=smag)
    u0 = o.random_field(so, A=A, kp=4 if min(n) < 32 else 10, rng=np.random.default_rng(seed))
    # oracle
    method_o = o.RK44(T)
    st = o.create_stepper(method_o, so, o.psolver_spectral(so), u0.copy(order="F"))
    ke_o = [o.total_kinetic_energy(u0, so)]
    p_o = None
    for _ in range(nsteps):
        st, p_o = o.timestep(method_o, st, dt, theta=theta, return_p=True)
        ke_o.append(o.total_kinetic_energy(st.u, so))
    # CUDA, step by step through timestep! (exercises import/export every step)
    method_c = ins.RKMethods.RK44(T=T)
    ud = ins.to_device(u0, sc)
    pd = ins.scalarfield(sc)
    stc = ins.create_stepper(method_c, sc, ins.default_psolver(sc), ud)
    ke_c = [ins.total_kinetic_energy(ud, sc)]
    for _ in range(nsteps):
        stc = ins.timestep_(method_c, stc, dt, theta=theta, p=pd)
        ke_c.append(ins.total_kinetic_energy(stc.u, sc))
    return ins, so, sc, u0, st, p_o, np.array(ke_o, dtype=np.float64), stc, pd, np.array(ke_c)


@pytest.mark.parametrize("n,T,force,smag", [
    ((32, 32), np.float64, False, False),
    ((32, 32), np.float64, True, False),
    ((32, 16), np.float32, True, False),
    ((16, 16, 16), np.float64, False, False),
    ((16, 32, 16), np.float32, False, False),
    ((16, 16, 16), np.float64, False, True),
    ((32, 32), np.float64, False, True),
    ((16, 16, 32), np.float32, True, True),
], ids=lambda v: str(v).replace(" ", ""))
def test_rk44_steps_match_oracle(n, T, force, smag):
    nsteps, dt = 10, 2e-3
    theta = 0.17 if smag else None
    ins, so, sc, u0, st, p_o, ke_o, stc, pd, ke_c = run_pair(n, T, nsteps, dt, force=force, smag=smag, theta=theta)
    Ip = so.grid.Ip
    tol = TOL[T]
    compare("u after N steps", ins.to_host(stc.u), st.u, tol)       # ghosts included: valid on return
    compare("p (last stage multiplier)", ins.to_host(pd)[Ip], p_o[Ip], tol)
    assert np.allclose(ke_c, ke_o, rtol=tol, atol=0), (ke_c, ke_o)
    assert float(stc.t) == float(st.t) and stc.n == nsteps
    div = ins.to_host(ins.divergence(stc.u, sc))[Ip]
    assert np.abs(div).max() / max(n) < 100 * np.finfo(T).eps * max(np.abs(st.u).max(), 1e-30)


def test_rk_steps_equals_repeated_timestep():
    """insb200_rk_steps (device-resident loop) == repeated timestep! (reference layout every
    step): bit-identical, and the kinetic-energy history it returns matches the reduction op."""
    import insb200 as ins
    from insb200.api import _rk_steps

    so, sc = make_pair((16, 16, 16), np.float64)
    u0 = o.random_field(so, kp=4, rng=np.random.default_rng(3))
    m = ins.RKMethods.RK44()
    a = ins.to_device(u0, sc)
    st = ins.create_stepper(m, sc, None, a)
    ke = [ins.total_kinetic_energy(a, sc)]
    for _ in range(5):
        st = ins.timestep_(m, st, 1e-3)
        ke.append(ins.total_kinetic_energy(st.u, sc))
    b = ins.to_device(u0, sc)
    t, hist = _rk_steps(m, sc, b, 0.0, 1e-3, 5, None, ke_hist=True)
    compare("rk_steps vs timestep!", ins.to_host(b), ins.to_host(a), 0, exact=True)
    assert np.allclose(hist, ke, rtol=1e-14) and float(t) == float(st.t)


def test_other_explicit_tableaux():
    """Any explicit tableau is an argument: RK33P2 has a negative and an
    off-diagonal entry, FE11 a single stage."""
    import insb200 as ins

    so, sc = make_pair((32, 32), np.float64)
    u0 = o.random_field(so, kp=4, rng=np.random.default_rng(4))
    for name, A, b, c in [("RK33P2", [[0, 0, 0], [1 / 3, 0, 0], [-1, 2, 0]], [0, 3 / 4, 1 / 4], [0, 1 / 3, 1]),
                          ("FE11", [[0]], [1], [0])]:
        mo = o.runge_kutta_method(A, b, c, 0)
        st = o.create_stepper(mo, so, o.psolver_spectral(so), u0.copy(order="F"))
        for _ in range(3):
            st = o.timestep(mo, st, 2e-3)
        mc = getattr(ins.RKMethods, name)()
        stc = ins.create_stepper(mc, sc, None, ins.to_device(u0, sc))
        for _ in range(3):
            stc = ins.timestep_(mc, stc, 2e-3)
        compare(name, ins.to_host(stc.u), st.u, 1e-12)


def test_solve_unsteady_adaptive_config1_small():
    """BASELINE config 1 (DecayingTurbulence2D, adaptive dt, cfl 0.9) at 64^2 Float64:
    solve_unsteady through the host mirror AND through the host-buffer C entry point."""
    import insb200 as ins

    n = (64, 64)
    so, sc = make_pair(n, np.float64, Re=4e3)
    u0 = o.random_field(so, rng=np.random.default_rng(5))
    state_o, out_o = o.solve_unsteady(so, (0.0, 0.02), u0, processors={"ke": lambda s: float(o.total_kinetic_energy(s.u, so))})
    state_c, out_c = ins.solve_unsteady(sc, (0.0, 0.02), ins.to_device(u0, sc), processors={"ke": ins.energy_history(sc)})
    assert state_c.n == state_o.n and np.isclose(float(state_c.t), float(state_o.t), rtol=1e-14)
    compare("adaptive solve u", ins.to_host(state_c.u), state_o.u, 1e-12)
    assert np.allclose(out_c["ke"][:, 1], out_o["ke"], rtol=1e-12)
    uh = u0.copy(order="F")
    ns, t, ke = ins.solve_unsteady_host(sc, (0.0, 0.02), uh, ke_hist=256)
    assert ns == state_o.n
    compare("host entry u", uh, state_o.u, 1e-12)
    assert np.allclose(ke, out_o["ke"], rtol=1e-12)


def test_taylor_green_3d_energy_config2():
    """BASELINE config 2: TaylorGreenVortex3D periodic 128^3 Float64, Re=1e3, dt=1e-3.  Energy
    check vs reference: (i) CUDA energy history == oracle history at 32^3 (same tolerance as u),
    (ii) at the full 128^3 the inviscid limit conserves energy to O(dt^4) and the viscous run
    dissipates monotonically with dE/dt ~ -2 nu * enstrophy (size-independent properties)."""
    import insb200 as ins

    T = np.float64
    # (i) small, against the oracle
    so, sc = make_pair((32, 32, 32), T, Re=1e3)
    u0 = o.velocityfield(so, tgv3)
    uc = ins.velocityfield(sc, tgv3)
    compare("velocityfield", ins.to_host(uc), u0, 1e-12)
    state_o, out_o = o.solve_unsteady(so, (0.0, 0.01), u0, dt=1e-3, processors={"ke": lambda s: float(o.total_kinetic_energy(s.u, so))})
    state_c, out_c = ins.solve_unsteady(sc, (0.0, 0.01), uc, dt=1e-3, processors={"ke": ins.energy_history(sc)})
    compare("TGV u", ins.to_host(state_c.u), state_o.u, 1e-12)
    assert np.allclose(out_c["ke"][:, 1], out_o["ke"], rtol=1e-12)
    # (ii) full size, properties
    n = 128
    x = tuple(np.linspace(0.0, 1.0, n + 1) for _ in range(3))
    for Re, check in ((1e12, "conserve"), (1e3, "dissipate")):
        s = ins.Setup(x, Re=Re)
        u = ins.velocityfield(s, tgv3)
        state, out = ins.solve_unsteady(s, (0.0, 0.02), u, dt=1e-3, processors={"ke": ins.energy_history(s)})
        E = out["ke"][:, 1]
        if check == "conserve":
            assert abs(E[-1] - E[0]) / E[0] < 1e-10, E
        else:
            assert np.all(np.diff(E) < 0) and 1e-4 < (E[0] - E[-1]) / E[0] < 1e-1, E
        div = ins.to_host(ins.divergence(state.u, s))[s.grid.Ip]
        assert np.abs(div).max() / n < 100 * np.finfo(T).eps


def test_processors_and_observers():
    """processors API (src/processors.jl): timelogger, fieldsaver, observefield(:pressure)."""
    import insb200 as ins

    so, sc = make_pair((32, 32), np.float64)
    u0 = o.random_field(so, kp=4, rng=np.random.default_rng(6))
    lines = []
    procs = {"log": ins.timelogger(sc, nupdate=2, sink=lines.append), "save": ins.fieldsaver(sc, nupdate=4),
             "p": ins.observefield(sc, "pressure", nupdate=4)}
    state, out = ins.solve_unsteady(sc, (0.0, 0.008), ins.to_device(u0, sc), dt=1e-3, processors=procs)
    assert state.n == 8 and len(lines) == 4 and "umax" in lines[0]
    assert [s.n for s in out["save"]] == [4, 8]
    po = o.pressure(out["save"][-1].u, None, 0.0, so, o.psolver_spectral(so))
    compare("observed pressure", out["p"][-1], po[so.grid.Ip], 1e-12)


def set_param(sc, name, value):
    sc.ctx.call("insb200_set_param", name.encode(), float(value))


@pytest.mark.parametrize("n,T,force,smag", [
    ((32, 32, 24), np.float64, False, False),
    ((64, 32, 20), np.float32, True, False),
    ((32, 16, 16), np.float64, False, True),
    ((32, 32, 16), np.float32, True, True),
], ids=lambda v: str(v).replace(" ", ""))
def test_tiled_stage_kernel_matches_oracle(n, T, force, smag):
    """Tile-aligned 3-D grids take the shared-memory stage kernel (fused divergence, projection of
    stage i applied by the loader of stage i+1): same tolerances against the oracle."""
    nsteps, dt = 6, 2e-3
    theta = 0.17 if smag else None
    ins, so, sc, u0, st, p_o, ke_o, stc, pd, ke_c = run_pair(n, T, nsteps, dt, force=force, smag=smag, theta=theta)
    tol = TOL[T]
    compare("u after N steps", ins.to_host(stc.u), st.u, tol)
    compare("p (last stage multiplier)", ins.to_host(pd)[so.grid.Ip], p_o[so.grid.Ip], tol)
    assert np.allclose(ke_c, ke_o, rtol=tol, atol=0), (ke_c, ke_o)


@pytest.mark.parametrize("T", [np.float64, np.float32], ids=["f64", "f32"])
@pytest.mark.parametrize("smag", [False, True], ids=["nosmag", "smag"])
def test_kernel_variants_bit_identical(T, smag):
    """The tiled kernels run the same device arithmetic as the one-thread-per-cell kernels: every
    tuning (tile height, z-chunk, fused / separate gradient) gives bit-identical u after 4 steps."""
    import insb200 as ins
    from insb200.api import _rk_steps

    n = (64, 32, 40)
    so, sc = make_pair(n, T, force=True, smag=smag)
    theta = 0.17 if smag else None
    u0 = o.random_field(so, kp=4, rng=np.random.default_rng(12))
    m = ins.RKMethods.RK44(T=T)
    results = {}
    set_param(sc, "exact", 1)  # the bit-exact kernels; the fast kernels (fast.cu) are compared in tests/test_gpu_fast.py
    for name, params in [("naive", dict(variant=0)),
                         ("tiled16", dict(variant=1, tile_y=16, zchunk=64, fuse_gradient=1)),
                         ("tiled8_zc16", dict(variant=1, tile_y=8, zchunk=16, fuse_gradient=1)),
                         ("tiled16_zc7_nofuse", dict(variant=1, tile_y=16, zchunk=7, fuse_gradient=0)),
                         ("tma_zc64", dict(variant=2, zchunk=64)),
                         ("tma_zc7", dict(variant=2, zchunk=7)),
                         ("tma_zc2", dict(variant=2, zchunk=2)),
                         ("tma_smag_v1", dict(variant=2, zchunk=64, smag_v=1)),
                         ("tma_smag_v2_zc128", dict(variant=2, smag_v=2, smag_zchunk=128)),
                         ("tma_smag_v2_zc7", dict(variant=2, smag_v=2, smag_zchunk=7)),
                         ("tma_smag_v2_zc2", dict(variant=2, smag_v=2, smag_zchunk=2)),
                         ("tma_smag_v2_ty8", dict(variant=2, smag_v=2, smag_zchunk=16, smag_ty=8))]:
        for k, v in params.items():
            set_param(sc, k, v)
        u = ins.to_device(u0, sc)
        _, ke = _rk_steps(m, sc, u, 0.0, 1e-3, 4, theta, ke_hist=True)
        results[name] = (ins.to_host(u), ke)
    ref_u, ref_ke = results["naive"]
    for name, (u, ke) in results.items():
        if "fused" in name and T == np.float32:
            # FFMA2-contracted products (f32x2.cuh): same operations with fewer roundings, not bit-identical
            compare(name, u, ref_u, 2e-6)
            assert np.allclose(ke, ref_ke, rtol=1e-6, atol=0), name
            continue
        compare(name, u, ref_u, 0, exact=True)
        assert np.array_equal(ke, ref_ke), name


@pytest.mark.parametrize("T", [np.float64, np.float32], ids=["f64", "f32"])
@pytest.mark.parametrize("n,smag", [((64, 48), False), ((70, 38), True), ((32, 6), False), ((58, 30), False), ((128, 64), False)],
                         ids=["64x48", "70x38_smag", "32x6", "58x30", "128x64"])
def test_2d_row_marching_kernels_bit_identical(T, n, smag):
    """2-D: the kernels that march several rows per thread (divergence, gradient) perform the arithmetic of
    the one-cell-per-thread kernels: bit-identical u, p and energies, also when the row count is not a multiple of
    the rows per thread."""
    import insb200 as ins
    from insb200.api import _rk_steps

    so, sc = make_pair(n, T, force=True, smag=smag)
    theta = 0.17 if smag else None
    u0 = o.random_field(so, kp=4, rng=np.random.default_rng(5))
    m = ins.RKMethods.RK44(T=T)
    res = {}
    # (rows2d, fuse2d, rows_chunk): unfused one-cell kernels = reference; row-marching divergence / gradient; the fused
    # warp-marching stage kernel (stage2d.cuh; falls back to the unfused path with a closure model) with automatic,
    # tiny and odd row chunks
    configs = [(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 5), (1, 1, 2), (1, 1, 17)]
    for cfg in configs:
        for k, v in zip(("rows2d", "fuse2d", "rows_chunk"), cfg):
            set_param(sc, k, v)
        u = ins.to_device(u0, sc)
        _, ke = _rk_steps(m, sc, u, 0.0, 1e-3, 3, theta, ke_hist=True)
        p = ins.pressure(u, None, 0.0, sc)
        G = ins.pressuregradient(p, sc)
        d = ins.divergence(ins.apply_bc_u(G, 0.0, sc), sc)
        res[cfg] = [ins.to_host(x) for x in (u, p, G, d)] + [ke]
    for cfg in configs[1:]:
        for a, b, name in zip(res[cfg], res[configs[0]], ("u", "p", "grad p", "div grad p", "ke")):
            compare(f"{name} {cfg}", a, b, 0, exact=True)


@pytest.mark.parametrize("n,T,smag", [((32, 32, 16), np.float32, True), ((24, 20, 16), np.float64, False), ((64, 48), np.float64, False),
                                      ((70, 38), np.float32, True)], ids=["3d_tma_smag", "3d_generic", "2d_fused", "2d_smag"])
def test_cuda_graph_replay_is_bit_identical(n, T, smag):
    """Whole RK steps replayed as CUDA graphs (one rank, fixed dt, non-default stream) give the results of the eager
    launches: same kernels, same arguments, buffer rotation restored from the captured state."""
    import torch

    import insb200 as ins
    from insb200.api import _rk_steps

    so, sc = make_pair(n, T, force=True, smag=smag)
    theta = 0.17 if smag else None
    u0 = o.random_field(so, kp=4, rng=np.random.default_rng(9))
    m = ins.RKMethods.RK44(T=T)
    lib, h = sc.ctx.lib, sc.ctx.h
    out = {}
    side = torch.cuda.Stream()
    for graphs in (0, 1):
        set_param(sc, "graphs", graphs)
        with torch.cuda.stream(side):
            u = ins.to_device(u0, sc)
            r0 = lib.insb200_graph_replays(h)
            _, ke = _rk_steps(m, sc, u, 0.0, 1e-3, 9, theta, ke_hist=True)
            st = ins.create_stepper(m, sc, None, u)
            for _ in range(4):  # one step per call: the rotation state carries over between calls
                st = ins.timestep_(m, st, 1e-3, theta=theta)
            side.synchronize()
            out[graphs] = (ins.to_host(st.u), ke, lib.insb200_graph_replays(h) - r0)
    assert out[0][2] == 0 and out[1][2] >= 6, (out[0][2], out[1][2])
    compare("u eager vs graphs", out[1][0], out[0][0], 0, exact=True)
    assert np.array_equal(out[1][1], out[0][1])
    # the legacy default stream cannot be captured: the library runs eagerly there
    r0 = lib.insb200_graph_replays(h)
    u = ins.to_device(u0, sc)
    _rk_steps(m, sc, u, 0.0, 1e-3, 5, theta)
    assert lib.insb200_graph_replays(h) == r0


def test_checkpoint_restart_is_bit_identical(tmp_path):
    """A run interrupted by save_checkpoint / load_checkpoint continues exactly as the uninterrupted one."""
    import insb200 as ins

    so, sc = make_pair((32, 16, 16), np.float32, smag=True)
    u0 = o.random_field(so, kp=3, rng=np.random.default_rng(4))
    ref, _ = ins.solve_unsteady(sc, (0.0, 0.008), ins.to_device(u0, sc), dt=1e-3, theta=0.17)
    stem = str(tmp_path / "run")
    half, out = ins.solve_unsteady(sc, (0.0, 0.004), ins.to_device(u0, sc), dt=1e-3, theta=0.17,
                                   processors={"ck": ins.checkpointer(sc, stem, nupdate=4)})
    assert out["ck"] == [stem + "_000004"]
    st = ins.load_checkpoint(out["ck"][-1], sc)
    assert st.n == 4 and abs(float(st.t) - 0.004) < 1e-9
    rest, _ = ins.solve_unsteady(sc, (float(st.t), 0.008), st.u, dt=1e-3, theta=0.17)
    assert np.array_equal(ins.to_host(rest.u), ins.to_host(ref.u))
    with pytest.raises(ValueError):
        ins.load_checkpoint(out["ck"][-1], make_pair((32, 16, 32), np.float32)[1])
