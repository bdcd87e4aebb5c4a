"""Shared helpers of the GPU parity tests: oracle/CUDA setup pairs and array comparison."""
import numpy as np

from oracle import ins_oracle as o


def axes(n, T, lims=(0.0, 1.0)):
    return tuple(np.linspace(lims[0], lims[1], m + 1).astype(T) for m in n)


def make_pair(n, T, Re=1000.0, lims=(0.0, 1.0), force=False, smag=False):
    """(oracle setup, CUDA setup) on identical grids."""
    import insb200 as ins

    x = axes(n, T, lims)
    D = len(n)
    bf = None
    if force:  # Kolmogorov force of examples/Kolmogorov2D.jl:29-36
        bf = (lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x) if D == 2 else (
            lambda dim, x, y, z, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x + 0 * z)
    so = o.Setup(x, Re=T(Re), bodyforce=bf)
    sc = ins.Setup(x, Re=T(Re), bodyforce=bf)
    if smag:
        so.closure_model = o.smagorinsky_closure(so)
        sc.closure_model = ins.smagorinsky_closure(sc)
    return so, sc


def rand_u(so, seed, T, scale=1.0):
    rng = np.random.default_rng(seed)
    u = np.asfortranarray((scale * rng.standard_normal(so.grid.N + (so.grid.dimension,))).astype(T))
    return o.apply_bc_u(u, 0.0, so)


def rand_p(so, seed, T):
    rng = np.random.default_rng(seed)
    p = np.asfortranarray(rng.standard_normal(so.grid.N).astype(T))
    return o.apply_bc_p(p, 0.0, so)


def compare(name, got, want, rtol, exact=False, target=None):
    """relative L2 (and max) error with a diagnostic that locates the worst entry.
    `target`: north_star's tolerance (1e-12 Float64 / 1e-5 Float32) where `rtol` is wider than it; the measured error is
    then printed as a `[tolerance]` line (pytest -s) so that DESIGN.md's table of widened tolerances can quote it."""
    got = np.asarray(got)
    want = np.asarray(want)
    assert got.shape == want.shape, f"{name}: shape {got.shape} vs {want.shape}"
    g64, w64 = got.astype(np.float64), want.astype(np.float64)
    diff = np.abs(g64 - w64)
    nrm = max(np.sqrt((w64 ** 2).sum()), 1e-300)
    rel = np.sqrt((diff ** 2).sum()) / nrm
    idx = np.unravel_index(np.argmax(diff), diff.shape) if diff.size else ()
    msg = (f"{name}: rel_l2={rel:.3e} max_abs={diff.max() if diff.size else 0:.3e} at {idx} "
           f"got={got[idx] if diff.size else None} want={want[idx] if diff.size else None} "
           f"nbad={(diff > 0).sum()}/{diff.size}")
    if target is not None:
        print(f"\n[tolerance] {name} ({got.dtype}, shape {got.shape}): measured rel-L2 {rel:.3e}, north_star {target:g}, allowed {rtol:g}"
              + ("" if rel <= target else "  <-- ABOVE north_star"))
    if exact:
        assert np.array_equal(got, want), "NOT BIT-EXACT " + msg
    else:
        assert np.isfinite(g64).all() and rel <= rtol, msg
    return rel
