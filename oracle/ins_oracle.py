"""CPU oracle for the periodic RK44 hot path of IncompressibleNavierStokes.jl.

TEST INFRASTRUCTURE ONLY.  This module is a NumPy restatement of the reference
algorithm (Julia, /root/reference, v2.2.0).  It may be imported from `tests/`,
`__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of
`bench.py`, and from nowhere else.  The product (libinsb200.so and the host
mirror in `incompressiblenavierstokes.jl_b200/`) never routes through it.

Parity status: the reference is Julia and there is no Julia runtime in the
build container, and the reference ships no golden vectors.  The oracle is
therefore pinned against every *property / known-answer* test the reference
holds for this path (tests/test_oracle_properties.py):
  test/operators.jl:58-88   D = -G'            test/operators.jl:107-128  convection skew-symmetry
  test/operators.jl:130-151 diffusion <= 0     test/operators.jl:153-160  fused == conv + diff
  test/psolvers.jl:1-32     spectral Poisson known answer
  examples/TaylorGreenVortex2D.jl:29-73  analytic 2-D TGV, error ~ n^-2
3-D / Float32 spectral results are "parity unpinned" by the reference's own
tests; they are pinned here by identities (post-projection
divergence at machine epsilon, laplacian(psolve(f)) == f - mean(f)).

Layout (reference `src/grid.jl:100-276`, `src/initializers.jl:1-6`):
  scalar field  shape (N1,..,ND)      N = n + 2 (one ghost volume each side)
  vector field  shape (N1,..,ND, D)   component index last (slowest in Julia's
                                      column-major memory)
Arrays are created in Fortran order so that `arr.ravel(order="K")` is the exact
byte image of the Julia array.  Index I (Julia, 1-based) == index I-1 here.

Arithmetic is kept in the field dtype (float64 or float32) with the reference's
operation order; NumPy ufuncs do not contract a*b+c into an FMA, like Julia.
"""
from __future__ import annotations

import math
from types import SimpleNamespace

import numpy as np
import scipy.fft as sfft

__all__ = [
    "Setup", "Grid", "scalarfield", "vectorfield", "apply_bc_u", "apply_bc_p",
    "divergence", "scalewithvolume", "pressuregradient", "applypressure",
    "laplacian", "convection", "diffusion", "convectiondiffusion", "applybodyforce",
    "momentum", "smagtensor", "divoftensor", "smagorinsky_closure", "psolver_spectral",
    "poisson", "project", "pressure", "kinetic_energy", "total_kinetic_energy",
    "get_cfl_timestep", "RK44", "runge_kutta_method", "create_stepper", "timestep",
    "solve_unsteady", "create_spectrum", "random_field", "velocityfield",
    "to_julia_bytes", "interior", "rel_l2",
]


# --------------------------------------------------------------------------
# Grid and Setup  (src/grid.jl:100-276, src/setup.jl:2-46)
# --------------------------------------------------------------------------

def _padghost_periodic(x):
    """padghost!(::PeriodicBC, ...) src/boundary_conditions.jl:42-49 (left then right)."""
    dxb = x[-1] - x[-2]
    x = np.concatenate(([x[0] - dxb], x))
    # right pad uses x[2]-x[1] of the *already left-padded* vector, exactly like the
    # reference, which pads left first (grid.jl:115-119)
    dxa = x[1] - x[0]
    x = np.concatenate((x, [x[-1] + dxa]))
    return x


def Grid(x):
    """Periodic-BC restatement of `Grid` (src/grid.jl:100-276).

    `x` : tuple of 1-D coordinate vectors (n+1 volume faces per axis).
    """
    x = tuple(np.array(xa) for xa in x)
    T = x[0].dtype.type
    D = len(x)
    xlims = tuple((xa.min(), xa.max()) for xa in x)
    x = tuple(_padghost_periodic(xa).astype(T) for xa in x)
    N = tuple(len(xa) - 1 for xa in x)
    # offset_u = offset_p = 1 for PeriodicBC (boundary_conditions.jl:79-80)
    Np = tuple(n - 2 for n in N)
    Nu = tuple(tuple(N[b] - 2 for b in range(D)) for a in range(D))
    # index ranges as python slices (0-based, half open): Julia 2:N-1
    Ip = tuple(slice(1, n - 1) for n in N)
    Iu = tuple(Ip for _ in range(D))
    xu = tuple(
        tuple(x[b][1:] if a == b else ((x[b][:-1] + x[b][1:]) / T(2)).astype(T) for b in range(D))
        for a in range(D)
    )
    xp = tuple(((xa[:-1] + xa[1:]) / T(2)).astype(T) for xa in x)
    eps = np.finfo(T).eps
    Delta = tuple(np.maximum(np.diff(xa), eps).astype(T) for xa in x)
    Delta_u = tuple(
        np.maximum(np.concatenate((np.diff(xp[d]), [Delta[d][-1] / T(2)])), eps).astype(T)
        for d in range(D)
    )
    # interpolation weights grid.jl:227-248 ; A[a][b] = (A1, A2)
    A = []
    for a in range(D):
        row = []
        for b in range(D):
            if a == b:
                A1 = np.full(N[a], T(0.5), dtype=T)
                A1[0] = 1
                A2 = np.full(N[a], T(0.5), dtype=T)
                A2[-1] = 1
            else:
                A2 = np.array(
                    [(x[b][i] - xp[b][i - 1]) / Delta_u[b][i - 1] for i in range(1, N[b])], dtype=T
                )
                A1 = (T(1) - A2).astype(T)
                A1 = np.concatenate(([T(1)], A1)).astype(T)
                A2 = np.concatenate((A2, [T(1)])).astype(T)
            row.append((A1, A2))
        A.append(tuple(row))
    return SimpleNamespace(
        xlims=xlims, dimension=D, N=N, Nu=Nu, Np=Np, Iu=Iu, Ip=Ip, x=x, xu=xu, xp=xp,
        Delta=Delta, Delta_u=Delta_u, A=tuple(A), T=T,
    )


def Setup(x, Re=None, bodyforce=None, issteadybodyforce=True, closure_model=None, temperature=None):
    """`Setup(; x, Re, bodyforce, issteadybodyforce, closure_model)` src/setup.jl:2-46.

    Only periodic boundary conditions (the hot path's admissible domain,
    src/utils.jl:1-13).  `bodyforce(dim, x..., t)` uses 1-based `dim`.
    """
    grid = Grid(x)
    T = grid.T
    if Re is None:  # setup.jl:12: Re = isnothing(temperature) ? 1000 : 1 / temperature.α1
        Re = T(1000) if temperature is None else T(1) / T(temperature.alpha1)
    Re = T(Re)
    setup = SimpleNamespace(
        grid=grid, Re=Re, bodyforce=bodyforce, issteadybodyforce=False,
        closure_model=closure_model, temperature=temperature,
    )
    if bodyforce is not None and issteadybodyforce:
        # setup.jl:25-32: materialise the steady force once as a vector field
        F = vectorfield(setup)
        u = vectorfield(setup)
        setup.bodyforce = applybodyforce(F, u, T(0), setup)
        setup.issteadybodyforce = True
    return setup


def scalarfield(setup):
    """src/initializers.jl:1-2"""
    return np.zeros(setup.grid.N, dtype=setup.grid.T, order="F")


def vectorfield(setup):
    """src/initializers.jl:4-6"""
    g = setup.grid
    return np.zeros(g.N + (g.dimension,), dtype=g.T, order="F")


def to_julia_bytes(a):
    """The exact memory image of the Julia array (column-major)."""
    return np.asfortranarray(a).ravel(order="K")


def interior(a, setup):
    """`a[Ip]` / `a[Ip, :]`."""
    return a[setup.grid.Ip]


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.sqrt(np.sum((a - b) ** 2)) / max(np.sqrt(np.sum(b ** 2)), 1e-300))


# slicing helpers -----------------------------------------------------------

def _int(N, shift=None):
    """Slices addressing I + shift for I in the interior 2:N-1 (Julia)."""
    D = len(N)
    shift = (0,) * D if shift is None else shift
    return tuple(slice(1 + s, n - 1 + s) for n, s in zip(N, shift))


def _e(D, b, s=1):
    t = [0] * D
    t[b] += s
    return tuple(t)


def _add(a, b):
    return tuple(i + j for i, j in zip(a, b))


def _ax(v, b, D):
    """reshape a 1-D metric vector along axis b for broadcasting."""
    sh = [1] * D
    sh[b] = -1
    return v.reshape(sh)


def _m1(N, b, v, s=0):
    """metric vector v[I[b]+s] over the interior, broadcastable."""
    D = len(N)
    return _ax(v[1 + s: N[b] - 1 + s], b, D)


# --------------------------------------------------------------------------
# Boundary conditions  (src/boundary_conditions.jl:276-288, 306-318)
# --------------------------------------------------------------------------

def apply_bc_u(u, t, setup, copy=True):
    """apply_bc_u!(u, t, setup): periodic ghost fill, axes in order 1..D, both
    components, over the FULL extent of the other axes (boundary_conditions.jl:97-103)."""
    u = u.copy(order="F") if copy else u
    D = setup.grid.dimension
    for b in range(D):
        ia = [slice(None)] * (D + 1)
        ib = [slice(None)] * (D + 1)
        ja = [slice(None)] * (D + 1)
        jb = [slice(None)] * (D + 1)
        Nb = setup.grid.N[b]
        ia[b], ib[b], ja[b], jb[b] = 0, Nb - 1, 1, Nb - 2
        u[tuple(ia)] = u[tuple(jb)]
        u[tuple(ib)] = u[tuple(ja)]
    return u


def apply_bc_p(p, t, setup, copy=True):
    """apply_bc_p!(p, t, setup) boundary_conditions.jl:306-318 (also used for the
    tensor field sigma, whose trailing axes ride along)."""
    p = p.copy(order="F") if copy else p
    D = setup.grid.dimension
    for b in range(D):
        ia = [slice(None)] * p.ndim
        ib = [slice(None)] * p.ndim
        ja = [slice(None)] * p.ndim
        jb = [slice(None)] * p.ndim
        Nb = setup.grid.N[b]
        ia[b], ib[b], ja[b], jb[b] = 0, Nb - 1, 1, Nb - 2
        p[tuple(ia)] = p[tuple(jb)]
        p[tuple(ib)] = p[tuple(ja)]
    return p


# --------------------------------------------------------------------------
# Operators  (src/operators.jl)
# --------------------------------------------------------------------------

def divergence(u, setup):
    """divergence! src/operators.jl:106-125: div[I] = sum_a (u[I,a]-u[I-e_a,a]) / D[a][I[a]]."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    div = scalarfield(setup)
    d = np.zeros(g.Np, dtype=T)
    for a in range(D):
        d = d + (u[_int(N) + (a,)] - u[_int(N, _e(D, a, -1)) + (a,)]) / _m1(N, a, g.Delta[a])
    div[g.Ip] = d
    return div


def scalewithvolume(p, setup, inplace_form=True):
    """scalewithvolume! src/operators.jl:81-95 (`p *= dx*dy*dz`, volume formed first)
    or the allocating twin :64-78 (`((p*dx)*dy)*dz`) when inplace_form=False."""
    g = setup.grid
    D = g.dimension
    d = [_ax(g.Delta[a], a, D) for a in range(D)]
    if inplace_form:
        vol = d[0] * d[1]
        if D == 3:
            vol = vol * d[2]
        return p * vol
    out = p * d[0]
    out = out * d[1]
    if D == 3:
        out = out * d[2]
    return out


def pressuregradient(p, setup):
    """pressuregradient! src/operators.jl:159-178."""
    g = setup.grid
    D, N = g.dimension, g.N
    G = vectorfield(setup)
    for a in range(D):
        G[_int(N) + (a,)] = (p[_int(N, _e(D, a))] - p[_int(N)]) / _m1(N, a, g.Delta_u[a])
    return G


def applypressure(u, p, setup):
    """applypressure! src/operators.jl:214-233: u[I,a] -= (p[I+e_a]-p[I]) / Du[a][I[a]]."""
    g = setup.grid
    D, N = g.dimension, g.N
    u = u.copy(order="F")
    for a in range(D):
        u[_int(N) + (a,)] -= (p[_int(N, _e(D, a))] - p[_int(N)]) / _m1(N, a, g.Delta_u[a])
    return u


def laplacian(p, setup):
    """laplacian! src/operators.jl:297-364 (periodic branch = the generic `else`)."""
    g = setup.grid
    D, N = g.dimension, g.N
    L = scalarfield(setup)
    # OmegaI = prod(getindex.(Delta, I))  ((d1*d2)*d3)
    om = _m1(N, 0, g.Delta[0])
    for a in range(1, D):
        om = om * _m1(N, a, g.Delta[a])
    acc = np.zeros(g.Np, dtype=g.T)
    for a in range(D):
        acc = acc + om / _m1(N, a, g.Delta[a]) * (
            (p[_int(N, _e(D, a))] - p[_int(N)]) / _m1(N, a, g.Delta_u[a])
            - (p[_int(N)] - p[_int(N, _e(D, a, -1))]) / _m1(N, a, g.Delta_u[a], -1)
        )
    L[g.Ip] = acc
    return L


def _conv_diff_terms(u, setup, do_conv, do_diff, visc):
    """Shared body of convection_kernel! (:392-415), diffusion_kernel! (:551-573) and
    convection_diffusion_kernel! (:647-690).  Operation order follows the fused kernel;
    the separate kernels use `f -= conv/D` resp. `f += visc*(db-da)/D`."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    F = vectorfield(setup)
    two = T(2)
    for a in range(D):
        f = np.zeros(g.Np, dtype=T)
        uI = u[_int(N) + (a,)]
        for b in range(D):
            eb = _e(D, b)
            ea = _e(D, a)
            Dab = _m1(N, b, g.Delta_u[b] if a == b else g.Delta[b])
            um = u[_int(N, _e(D, b, -1)) + (a,)]
            up = u[_int(N, eb) + (a,)]
            if do_conv:
                ua1 = (um + uI) / two
                ua2 = (uI + up) / two
                # A[b][a][2][I[a]-(a==b)] * u[I-e_b, b] + A[b][a][1][I[a]+(a!=b)] * u[I-e_b+e_a, b]
                A1, A2 = g.A[b][a]
                w1 = _m1(N, a, A2, -1 if a == b else 0)
                w2 = _m1(N, a, A1, 0 if a == b else 1)
                ub1 = w1 * u[_int(N, _e(D, b, -1)) + (b,)] + w2 * u[_int(N, _add(_e(D, b, -1), ea)) + (b,)]
                ub2 = _m1(N, a, A2) * u[_int(N) + (b,)] + _m1(N, a, A1, 1) * u[_int(N, ea) + (b,)]
                uu1 = ua1 * ub1
                uu2 = ua2 * ub2
            if do_diff:
                Da = _m1(N, b, g.Delta[b]) if b == a else _m1(N, b, g.Delta_u[b], -1)
                Db = _m1(N, b, g.Delta[b], 1) if b == a else _m1(N, b, g.Delta_u[b])
                d1 = (uI - um) / Da
                d2 = (up - uI) / Db
                eps2 = two * np.finfo(T).eps
                d1 = (Da > eps2).astype(T) * d1
                d2 = (Db > eps2).astype(T) * d2
            if do_conv and do_diff:
                f = f + (visc * (d2 - d1) - (uu2 - uu1)) / Dab
            elif do_conv:
                f = f - (uu2 - uu1) / Dab
            else:
                f = f + visc * (d2 - d1) / Dab
        F[_int(N) + (a,)] = f
    return F


def convection(u, setup):
    """convection! src/operators.jl:378-415 (result on zeros)."""
    return _conv_diff_terms(u, setup, True, False, None)


def diffusion(u, setup, use_viscosity=True):
    """diffusion! src/operators.jl:537-573."""
    T = setup.grid.T
    visc = T(1) / setup.Re if use_viscosity else T(1)
    return _conv_diff_terms(u, setup, False, True, visc)


def convectiondiffusion(u, setup):
    """convectiondiffusion! src/operators.jl:634-690 (result on zeros)."""
    T = setup.grid.T
    return _conv_diff_terms(u, setup, True, True, T(1) / setup.Re)


def applybodyforce(F, u, t, setup):
    """applybodyforce! src/operators.jl:861-879."""
    g = setup.grid
    D = g.dimension
    if setup.issteadybodyforce:
        return F + setup.bodyforce
    F = F.copy(order="F")
    for a in range(D):
        xin = [_ax(g.xu[a][b], b, D) for b in range(D)]
        val = np.vectorize(lambda *xs: setup.bodyforce(a + 1, *xs, t), otypes=[g.T])(*np.broadcast_arrays(*xin))
        F[..., a] = F[..., a] + val.astype(g.T)
    return F


def momentum(u, temp, t, setup):
    """momentum! src/operators.jl:967-976."""
    F = convectiondiffusion(u, setup)
    if setup.bodyforce is not None:
        F = applybodyforce(F, u, t, setup)
    if temp is not None:
        F = gravity(F, temp, setup)
    return F


# Temperature equation ------------------------------------------------------

def temperature_equation(Pr, Ra, Ge, dodissipation=True, gdir=2, nondim_type=1):
    """temperature_equation src/setup.jl:56-86 (periodic temperature boundary conditions only).
    `gdir` is 1-based like the reference; the coefficients carry the dtype of `Pr`."""
    T = np.asarray(Pr).dtype.type
    Pr, Ra, Ge = T(Pr), T(Ra), T(Ge)
    if nondim_type == 1:
        a1 = np.sqrt(Pr / Ra)
        a2 = T(1)
        a3 = Ge * np.sqrt(Pr / Ra)
        a4 = T(1) / np.sqrt(Pr * Ra)
    elif nondim_type == 2:
        a1 = Pr
        a2 = Pr * Ra
        a3 = Ge / Ra
        a4 = T(1)
    elif nondim_type == 3:
        a1 = np.sqrt(Pr * Ge / Ra)
        a2 = Ge
        a3 = np.sqrt(Pr * Ge / Ra)
        a4 = np.sqrt(Ge / (Pr * Ra))
    else:
        raise ValueError("nondim_type must be 1, 2 or 3")
    a1, a2, a3, a4 = T(a1), T(a2), T(a3), T(a4)
    return SimpleNamespace(alpha1=a1, alpha2=a2, alpha3=a3, alpha4=a4, gamma=T(a1 / a3),
                           dodissipation=bool(dodissipation), gdir=int(gdir))


def apply_bc_temp(temp, t, setup, copy=True):
    """apply_bc_temp! src/boundary_conditions.jl:236-246; PeriodicBC -> apply_bc_p! (:338-339)."""
    return apply_bc_p(temp, t, setup, copy=copy)


def _avg(phi, setup, b, shift=None):
    """avg(phi, Delta, I + shift, b) src/operators.jl:59-62 for I in Ip:
    (D[b][I[b]+1]*phi[I] + D[b][I[b]]*phi[I+e_b]) / (D[b][I[b]] + D[b][I[b]+1])."""
    g = setup.grid
    D, N = g.dimension, g.N
    shift = (0,) * D if shift is None else shift
    d0 = _m1(N, b, g.Delta[b], shift[b])
    d1 = _m1(N, b, g.Delta[b], shift[b] + 1)
    return (d1 * phi[_int(N, shift)] + d0 * phi[_int(N, _add(shift, _e(D, b)))]) / (d0 + d1)


def convection_diffusion_temp(c, u, temp, setup):
    """convection_diffusion_temp! src/operators.jl:711-735 (adds to `c`, returns the new array)."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    a4 = setup.temperature.alpha4
    cI = np.zeros(g.Np, dtype=T)
    for b in range(D):
        eb, mb = _e(D, b), _e(D, b, -1)
        dTdx1 = (temp[_int(N)] - temp[_int(N, mb)]) / _m1(N, b, g.Delta_u[b], -1)
        dTdx2 = (temp[_int(N, eb)] - temp[_int(N)]) / _m1(N, b, g.Delta_u[b])
        uT1 = u[_int(N, mb) + (b,)] * _avg(temp, setup, b, mb)
        uT2 = u[_int(N) + (b,)] * _avg(temp, setup, b)
        cI = cI + (-(uT2 - uT1) + a4 * (dTdx2 - dTdx1)) / _m1(N, b, g.Delta[b])
    c = c.copy(order="F")
    c[g.Ip] = c[g.Ip] + cI
    return c


def dissipation(diss, u, setup):
    """dissipation! src/operators.jl:787-808: diff = diffusion!(zeros, u) (its ghosts stay 0), then
    diss[I] += sum_b Re*a1/gamma * (u[I-e_b,b]*diff[I-e_b,b] + u[I,b]*diff[I,b]) / 2."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    tq = setup.temperature
    diff = diffusion(u, setup)
    d = np.zeros(g.Np, dtype=T)
    coef = setup.Re * tq.alpha1 / tq.gamma
    for b in range(D):
        mb = _e(D, b, -1)
        d = d + coef * (u[_int(N, mb) + (b,)] * diff[_int(N, mb) + (b,)] + u[_int(N) + (b,)] * diff[_int(N) + (b,)]) / T(2)
    diss = diss.copy(order="F")
    diss[g.Ip] = diss[g.Ip] + d
    return diss


def gravity(F, temp, setup):
    """gravity! src/operators.jl:916-931: F[I, gdir] += alpha2 * avg(temp, Delta, I, gdir) on Iu[gdir]."""
    g = setup.grid
    N = g.N
    tq = setup.temperature
    gd = tq.gdir - 1
    F = F.copy(order="F")
    F[_int(N) + (gd,)] = F[_int(N) + (gd,)] + tq.alpha2 * _avg(temp, setup, gd)
    return F


# Smagorinsky ---------------------------------------------------------------

def _dx(u, setup, a, b):
    """`dx(u, I, a, b, D[b], Du[b])` src/operators.jl:1023-1033 on all p-cells I in Ip."""
    g = setup.grid
    D, N = g.dimension, g.N
    T = g.T
    ea, eb = _e(D, a), _e(D, b)
    ma, mb = _e(D, a, -1), _e(D, b, -1)
    if a == b:
        return (u[_int(N) + (a,)] - u[_int(N, mb) + (a,)]) / _m1(N, b, g.Delta[b])
    Du0 = _m1(N, b, g.Delta_u[b])
    Dum = _m1(N, b, g.Delta_u[b], -1)
    return (
        (u[_int(N, eb) + (a,)] - u[_int(N) + (a,)]) / Du0
        + (u[_int(N, _add(ma, eb)) + (a,)] - u[_int(N, ma) + (a,)]) / Du0
        + (u[_int(N) + (a,)] - u[_int(N, mb) + (a,)]) / Dum
        + (u[_int(N, ma) + (a,)] - u[_int(N, _add(ma, mb)) + (a,)]) / Dum
    ) / T(4)


def smagtensor(u, theta, setup):
    """smagtensor! src/operators.jl:1135-1151.  Returns sigma of shape (N..., D, D)
    (the reference stores SMatrix{D,D} per cell); only Ip is written."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    theta = T(theta)
    G = [[_dx(u, setup, a, b) for b in range(D)] for a in range(D)]  # G[a][b] = d_b u_a
    S = [[(G[a][b] + G[b][a]) / T(2) for b in range(D)] for a in range(D)]
    # sum(S .* S): column-major order over the SMatrix (a fastest)
    ss = None
    for b in range(D):
        for a in range(D):
            term = S[a][b] * S[a][b]
            ss = term if ss is None else ss + term
    # gridsize: sqrt(sum(D[a][I[a]]^2))  operators.jl:1128
    d2 = None
    for a in range(D):
        term = _m1(N, a, g.Delta[a]) ** 2
        d2 = term if d2 is None else d2 + term
    d = np.sqrt(d2).astype(T)
    eddy = theta ** 2 * d ** 2 * np.sqrt(T(2) * ss)
    sigma = np.zeros(N + (D, D), dtype=T, order="F")
    for a in range(D):
        for b in range(D):
            sigma[_int(N) + (a, b)] = T(2) * eddy * S[a][b]
    return sigma


def divoftensor(sigma, setup):
    """divoftensor! src/operators.jl:1166-1238."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    s = vectorfield(setup)
    for a in range(D):
        acc = np.zeros(g.Np, dtype=T)
        ea = _e(D, a)
        for b in range(D):
            eb, mb = _e(D, b), _e(D, b, -1)
            Dab = _m1(N, b, g.Delta_u[b] if a == b else g.Delta[b])
            sg = sigma[..., a, b]
            if a == b:
                s2 = sg[_int(N, eb)]
                s1 = sg[_int(N)]
            else:
                s2 = (sg[_int(N)] + sg[_int(N, eb)] + sg[_int(N, _add(ea, eb))] + sg[_int(N, ea)]) / T(4)
                s1 = (sg[_int(N, mb)] + sg[_int(N)] + sg[_int(N, _add(ea, mb))] + sg[_int(N, ea)]) / T(4)
            acc = acc + (s2 - s1) / Dab
        s[_int(N) + (a,)] = acc
    return s


def smagorinsky_closure(setup):
    """smagorinsky_closure src/operators.jl:1294-1305: m(u, theta) -> s."""
    T = setup.grid.T

    def closure(u, theta):
        sigma = smagtensor(u, theta, setup)
        sigma = apply_bc_p(sigma, T(0), setup, copy=False)
        return divoftensor(sigma, setup)

    return closure


# --------------------------------------------------------------------------
# Pressure  (src/pressure.jl)
# --------------------------------------------------------------------------

def _sinpi(x):
    """sinpi for x in [0, 1] with the argument reduced first (Julia's sinpi is exact at
    multiples of 1/2)."""
    x = np.asarray(x, dtype=np.float64)
    r = np.where(x > 0.5, 1.0 - x, x)
    return np.sin(np.pi * r)


def psolver_spectral(setup, workers=-1):
    """psolver_spectral src/pressure.jl:289-351.  Returns psolve(p) -> new p (ghosts of the
    input are carried over untouched, as `copyto!(view(p, Ip), pI)` leaves them)."""
    g = setup.grid
    D, T = g.dimension, g.T
    Np = g.Np
    dx = [g.Delta[a][0] for a in range(D)]
    for a in range(D):
        assert np.allclose(g.Delta[a], dx[a]), "Spectral psolver requires uniform grid spacing."
        assert g.N[a] % 2 == 0, "Spectral psolver requires even number of volumes."
    kmax = tuple(Np[a] // 2 + 1 if a == 0 else Np[a] for a in range(D))
    omega = dx[0]
    for a in range(1, D):
        omega = T(omega * dx[a])
    ahat = []
    for a in range(D):
        k = np.arange(kmax[a])
        # `4 * O * sinpi(k / Np)^2 / dx^2`: k/Np is Float64, the product is stored as T
        val = np.float64(T(4) * omega) * _sinpi(k / Np[a]) ** 2 / np.float64(T(dx[a] * dx[a]))
        ahat.append(val.astype(T))
    axes = tuple(range(D - 1, -1, -1))  # last listed axis (0) is the halved one, as in Julia's rfft
    ctype = np.complex64 if T == np.float32 else np.complex128

    def psolve(p):
        pI = np.ascontiguousarray(p[g.Ip])
        phat = sfft.rfftn(pI, axes=axes, workers=workers).astype(ctype, copy=False)
        den = _ax(ahat[0], 0, D) + _ax(ahat[1], 1, D)
        if D == 3:
            den = den + _ax(ahat[2], 2, D)
        with np.errstate(divide="ignore", invalid="ignore"):
            # complex / real: both parts divided (Julia `-phat / r`)
            phat = (-phat.real / den + 1j * (-phat.imag / den)).astype(ctype)
        phat[(0,) * D] = 0
        pI = sfft.irfftn(phat, s=tuple(Np[a] for a in axes), axes=axes, workers=workers).astype(T, copy=False)
        out = p.copy(order="F")
        out[g.Ip] = pI
        return out

    psolve.ahat = ahat
    return psolve


def poisson(psolver, f):
    """poisson! src/pressure.jl:22"""
    return psolver(f)


def project(u, setup, psolver, return_p=False):
    """project! src/pressure.jl:69-82 (in-place arithmetic form)."""
    T = setup.grid.T
    p = divergence(u, setup)
    p = scalewithvolume(p, setup)
    p = poisson(psolver, p)
    p = apply_bc_p(p, T(0), setup, copy=False)
    u = applypressure(u, p, setup)
    return (u, p) if return_p else u


def pressure(u, temp, t, setup, psolver):
    """pressure! src/pressure.jl:41-49 (apply_bc_u!(F; dudt=true) is a periodic ghost fill)."""
    F = momentum(u, temp, t, setup)
    F = apply_bc_u(F, t, setup, copy=False)
    p = divergence(F, setup)
    p = scalewithvolume(p, setup)
    p = poisson(psolver, p)
    return apply_bc_p(p, t, setup, copy=False)


# --------------------------------------------------------------------------
# Reductions
# --------------------------------------------------------------------------

def kinetic_energy(u, setup):
    """kinetic_energy! (interpolate_first=false) src/operators.jl:1516-1545."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    ke = scalarfield(setup)
    k = np.zeros(g.Np, dtype=T)
    for a in range(D):
        k = k + (u[_int(N) + (a,)] ** 2 + u[_int(N, _e(D, a, -1)) + (a,)] ** 2)
    ke[g.Ip] = k / T(4)
    return ke


def total_kinetic_energy(u, setup):
    """total_kinetic_energy src/operators.jl:1551-1556 / energy_history_plot MakieExt.jl:281-295."""
    k = scalewithvolume(kinetic_energy(u, setup), setup)
    return k[setup.grid.Ip].sum()


def get_cfl_timestep(u, setup):
    """get_cfl_timestep! src/solver.jl:101-125."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    dt = T(np.inf)
    for a in range(D):
        dmin = g.Delta_u[a][1: N[a] - 1].min()
        dt_diff = setup.Re * dmin ** 2 / T(2)
        with np.errstate(divide="ignore"):
            buf = _ax(g.Delta_u[a], a, D) / np.abs(u[..., a])
        dt_conv = buf[g.Iu[a]].min()
        dt = min(dt, dt_diff, dt_conv)
    return T(dt)


def maxabs(u):
    """`maximum(abs, u)` src/processors.jl:67 (ghosts included)."""
    return np.abs(u).max()


# --------------------------------------------------------------------------
# Observers  (src/processors.jl:77-198, 303-332; src/utils.jl:49-108)
# --------------------------------------------------------------------------

def vorticity(u, setup):
    """vorticity! src/operators.jl:989-1021 over ndrange N-1 (Julia indices 1..N-1)."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    lo = tuple(slice(0, n - 1) for n in N)

    def sh(b):
        return tuple(slice(1, n) if a == b else slice(0, n - 1) for a, n in enumerate(N))

    def du(b):  # Delta_u[b][I[b]] for I[b] = 1..N-1
        return _ax(g.Delta_u[b][: N[b] - 1], b, D)

    if D == 2:
        w = scalarfield(setup)
        w[lo] = (u[sh(0) + (1,)] - u[lo + (1,)]) / du(0) - (u[sh(1) + (0,)] - u[lo + (0,)]) / du(1)
        return w
    w = vectorfield(setup)
    for a, ap, am in ((0, 1, 2), (1, 2, 0), (2, 0, 1)):
        w[lo + (a,)] = (u[sh(ap) + (am,)] - u[lo + (am,)]) / du(ap) - (u[sh(am) + (ap,)] - u[lo + (ap,)]) / du(am)
    return w


def interpolate_u_p(u, setup):
    """interpolate_u_p! src/operators.jl:1311-1326."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    up = vectorfield(setup)
    for a in range(D):
        up[_int(N) + (a,)] = (u[_int(N, _e(D, a, -1)) + (a,)] + u[_int(N) + (a,)]) / T(2)
    return up


def interpolate_w_p(w, setup):
    """interpolate_w_p! src/operators.jl:1338-1372."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    if D == 2:
        wp = scalarfield(setup)
        wp[_int(N)] = (w[_int(N, (-1, -1))] + w[_int(N)]) / T(2)
        return wp
    wp = vectorfield(setup)
    for a in range(D):
        back = _add(_e(D, (a + 1) % 3, -1), _e(D, (a + 2) % 3, -1))
        wp[_int(N) + (a,)] = (w[_int(N, back) + (a,)] + w[_int(N) + (a,)]) / T(2)
    return wp


def velocitynorm(up, setup):
    """observefield(:velocitynorm) src/processors.jl:141-151: sqrt(up1^2 + up2^2 [+ up3^2]), whole array."""
    D = setup.grid.dimension
    s = up[..., 0] ** 2 + up[..., 1] ** 2
    if D == 3:
        s = s + up[..., 2] ** 2
    return np.sqrt(s)


def observe_vorticity(u, t, setup):
    """observefield(:vorticity) src/processors.jl:152-155: apply_bc_u!, vorticity!, interpolate_w_p!, restricted to Ip."""
    u = apply_bc_u(u, t, setup)
    wp = interpolate_w_p(vorticity(u, setup), setup)
    return wp[setup.grid.Ip]


def spectral_stuff(setup, npoint=100, a=None):
    """spectral_stuff src/utils.jl:49-108.  Returns 0-based linear (column-major) index lists `inds`, kappa, K."""
    g = setup.grid
    D, T, Np = g.dimension, g.T, g.Np
    a = T((1 + np.sqrt(5)) / 2) if a is None else T(a)
    K = tuple(n // 2 for n in Np)
    ks = np.meshgrid(*[np.arange(k, dtype=np.int64) for k in K], indexing="ij")
    k = np.sqrt(sum(q.astype(np.float64) ** 2 for q in ks)).astype(T)
    k = k.reshape(-1, order="F")
    kmax = min(K) - 1
    isort = np.argsort(k, kind="stable")
    ksort = k[isort]
    kappa = np.exp(np.linspace(np.log(1.0), np.log(float(kmax)), npoint))  # logrange(T(1), T(kmax), npoint)
    kappa = np.array(sorted(set(int(v) for v in np.round(kappa))), dtype=np.int64)
    inds = []
    for ki in kappa:
        if D == 2:
            lo, hi = T(ki) / a, T(ki) * a
        else:
            tol = T(0.01)
            lo, hi = T(ki) - tol, T(ki) + T(1) - tol
        js = int(np.searchsorted(ksort, T(lo), side="left"))
        je = int(np.searchsorted(ksort, T(hi), side="left"))
        inds.append(isort[js:je])
    return SimpleNamespace(inds=inds, kappa=kappa, K=K)


def observespectrum(u, setup, npoint=100, a=None):
    """observespectrum src/processors.jl:303-332: per component FFT of u[Ip], |.|^2 / (2 prod(Np)^2) on the
    first K modes of every axis, summed over components and over the wavenumber shells."""
    import scipy.fft as sfft

    g = setup.grid
    D, T, Np = g.dimension, g.T, g.Np
    st = spectral_stuff(setup, npoint, a)
    ctype = np.complex64 if T == np.float32 else np.complex128
    e = None
    for al in range(D):
        uhat = sfft.fftn(u[g.Ip + (al,)].astype(ctype))
        half = uhat[tuple(slice(0, k) for k in st.K)]
        ea = (half.real ** 2 + half.imag ** 2).astype(T) / T(2 * int(np.prod(Np)) ** 2)
        e = ea if e is None else e + ea
    ef = e.reshape(-1, order="F")
    return np.array([ef[i].sum(dtype=T) for i in st.inds], dtype=T), st.kappa


def strain_gradient(u, setup):
    """`grad(u, I, D, Du)` src/operators.jl:1072-1088 on all I in Ip: G[a][b] = d_b u_a."""
    D = setup.grid.dimension
    return [[_dx(u, setup, a, b) for b in range(D)] for a in range(D)]


def dissipation_from_strain(u, setup):
    """dissipation_from_strain! src/operators.jl:819-836: eps[I] = 2 visc sum(S .* S), S = (G + G')/2."""
    g = setup.grid
    D, T = g.dimension, g.T
    G = strain_gradient(u, setup)
    ss = None
    for b in range(D):
        for a in range(D):
            S = (G[a][b] + G[b][a]) / T(2)
            term = S * S
            ss = term if ss is None else ss + term
    eps = scalarfield(setup)
    eps[g.Ip] = T(2) * (T(1) / setup.Re) * ss
    return eps


def Qfield(u, setup):
    """Qfield! src/operators.jl:1441-1460: Q[I] = -sum_{a,b} (d_b u_a)(d_a u_b) / 2 with one-sided differences
    (u[I,a]-u[I-e_b,a])/D[b] * (u[I,b]-u[I-e_a,b])/D[a] / 2, accumulated a outer, b inner."""
    g = setup.grid
    D, N, T = g.dimension, g.N, g.T
    q = np.zeros(g.Np, dtype=T)
    for a in range(D):
        for b in range(D):
            q = q - (u[_int(N) + (a,)] - u[_int(N, _e(D, b, -1)) + (a,)]) / _m1(N, b, g.Delta[b]) * (
                u[_int(N) + (b,)] - u[_int(N, _e(D, a, -1)) + (b,)]) / _m1(N, a, g.Delta[a]) / T(2)
    Q = scalarfield(setup)
    Q[g.Ip] = q
    return Q


def get_scale_numbers(u, setup):
    """get_scale_numbers src/operators.jl:1569-1619, statement by statement.  Note that `field = @. u^2 * Omega_u`
    (:1582-1583) runs over ALL components for every alpha, so uavg^2 is D times the volume mean of |u|^2."""
    g = setup.grid
    D, N, T, Np = g.dimension, g.N, g.T, g.Np
    visc = T(1) / setup.Re
    omega = scalewithvolume(np.ones(N, dtype=T, order="F"), setup)
    acc = T(0)
    for a in range(D):
        om_u = None
        for b in range(D):
            v = _ax(g.Delta_u[b] if a == b else g.Delta[b], b, D)
            om_u = v if om_u is None else om_u * v
        field = u ** 2 * om_u[..., None]
        acc = acc + field[g.Iu[0]].sum(dtype=T) / np.broadcast_to(om_u, N)[g.Iu[0]].sum(dtype=T)
    uavg = T(np.sqrt(acc))
    eps = dissipation_from_strain(u, setup)
    eps = T((omega * eps)[g.Ip].sum(dtype=T) / omega[g.Ip].sum(dtype=T))
    eta = T((visc ** 3 / eps) ** T(0.25))
    lam = T(np.sqrt(T(5) * visc / eps) * uavg)
    Relam = T(lam * uavg / np.sqrt(T(3)) / visc)
    K = tuple(n // 2 for n in Np)
    ctype = np.complex64 if T == np.float32 else np.complex128
    e = None
    ks = np.meshgrid(*[np.arange(k, dtype=np.float64) for k in K], indexing="ij")
    knorm = np.sqrt(sum(q ** 2 for q in ks)).astype(T)
    for a in range(D):
        uhat = sfft.fftn(u[g.Ip + (a,)].astype(ctype))
        half = uhat[tuple(slice(0, k) for k in K)]
        ea = (half.real ** 2 + half.imag ** 2).astype(T) / T(2 * int(np.prod(Np)) ** 2)
        with np.errstate(divide="ignore", invalid="ignore"):
            ea = ea / knorm
        e = ea if e is None else e + ea
    e[(0,) * D] = 0
    L = T(T(3 * np.pi) / T(2) / uavg ** 2 * e.sum(dtype=T))
    tau = T(L / uavg)
    Re_int = T(L * uavg / visc)
    return SimpleNamespace(uavg=uavg, eps=eps, eta=eta, lam=lam, Relam=Relam, L=L, tau=tau, Re_int=Re_int)


# --------------------------------------------------------------------------
# Time stepping  (src/time_steppers)
# --------------------------------------------------------------------------

def runge_kutta_method(A, b, c, r, T=np.float64):
    """runge_kutta_method src/time_steppers/methods.jl:222-240 (explicit branch: the tableau
    is shifted, A = [A[2:end,:]; b'], c = [c[2:end]; 1])."""
    A = np.array(A, dtype=np.float64)
    b = np.array(b, dtype=np.float64)
    c = np.array(c, dtype=np.float64)
    assert np.allclose(np.triu(A), 0), "only explicit tableaux are on the hot path"
    A = np.vstack((A[1:, :], b[None, :])).astype(T)
    c = np.concatenate((c[1:], [1.0])).astype(T)
    return SimpleNamespace(A=A, b=b.astype(T), c=c, r=T(r))


def RK44(T=np.float64):
    """RK44 src/time_steppers/RKMethods.jl:515-521."""
    A = [[0, 0, 0, 0], [0.5, 0, 0, 0], [0, 0.5, 0, 0], [0, 0, 1, 0]]
    b = [1 / 6, 1 / 3, 1 / 3, 1 / 6]
    c = np.sum(np.array(A), axis=1)
    return runge_kutta_method(A, b, c, 0, T=T)


def LMWray3(T=np.float64):
    """LMWray3 low-storage method (src/time_steppers/methods.jl; coefficients step_lmwray3.jl:77-79)."""
    return SimpleNamespace(lowstorage=True, a=(T(8 / 15), T(5 / 12), T(3 / 4)), b=(T(1 / 4), T(0)),
                           c=(T(0), T(8 / 15), T(2 / 3)), T=T)


def timestep_lmwray3(method, stepper, dt, theta=None):
    """timestep!(::LMWray3, ...) src/time_steppers/step_lmwray3.jl:4-104 (velocity only): two state
    registers (x, xstart) and one derivative register, exactly the reference's sequence of calls."""
    setup, psolver, u, t, n = stepper.setup, stepper.psolver, stepper.u, stepper.t, stepper.n
    T = setup.grid.T
    dt = T(dt)
    m = setup.closure_model
    tstart = t
    xstart = u.copy(order="F")                               # state_copyto!(xstart, x)  :61
    a, b, c = method.a, method.b, method.c
    nstage = len(a)
    for i in range(nstage):
        t = T(tstart + c[i] * dt)                            # :83
        u = apply_bc_u(u, t, setup, copy=False)              # f!: :17
        dx = momentum(u, None, t, setup)                     # :19
        if m is not None:
            dx = dx + m(u, theta)                            # :22
        u = xstart.copy(order="F")                           # :87
        u = u + T(a[i] * dt) * dx                            # :88  axpy!(a[i]*dt, dx, x)
        u = apply_bc_u(u, t, setup, copy=False)              # correct!: :38
        u = project(u, setup, psolver)                       # :39
        if i != nstage - 1:
            xstart = xstart + T(b[i] * dt) * dx              # :93
    t = T(tstart + dt)                                       # :97
    u = apply_bc_u(u, t, setup, copy=False)                  # :102
    return create_stepper(method, setup, psolver, u, None, t, n + 1)


# Explicit tableaux of src/time_steppers/RKMethods.jl used by the parity tests (file:line in the comments);
# (A, b, c or None = row sums, r) BEFORE the shift of runge_kutta_method.
_S6 = float(np.sqrt(6.0))
EXPLICIT_TABLEAUX = {
    "FE11": ([[0]], [1], [0], 1),                                                                   # :45-51
    "SSP22": ([[0, 0], [1, 0]], [0.5, 0.5], None, 1),                                               # :54-60
    "SSP33": ([[0, 0, 0], [1, 0, 0], [0.25, 0.25, 0]], [1 / 6, 1 / 6, 2 / 3], None, 1),             # :73-79
    "SSP43": ([[0, 0, 0, 0], [0.5, 0, 0, 0], [0.5, 0.5, 0, 0], [1 / 6, 1 / 6, 1 / 6, 0]], [1 / 6, 1 / 6, 1 / 6, 0.5], None, 2),  # :82-88
    "Wray3": ([[0, 0, 0], [8 / 15, 0, 0], [8 / 15 - 17 / 60, 5 / 12, 0]], [8 / 15 - 17 / 60, 0, 0.75],
              [0, 8 / 15, 8 / 15 - 17 / 60 + 5 / 12], 0),                                           # :137-146
    "RK56": ([[0] * 6, [0.25, 0, 0, 0, 0, 0], [0.125, 0.125, 0, 0, 0, 0], [0, 0, 0.5, 0, 0, 0],
              [3 / 16, -3 / 8, 3 / 8, 9 / 16, 0, 0], [-3 / 7, 8 / 7, 6 / 7, -12 / 7, 8 / 7, 0]],
             [7 / 90, 0, 16 / 45, 2 / 15, 16 / 45, 7 / 90], [0, 0.25, 0.25, 0.5, 0.75, 1], 0),       # :149-162
    "HEM5": ([[0] * 5, [0.3, 0, 0, 0, 0], [(1 + _S6) / 30, (11 - 4 * _S6) / 30, 0, 0, 0],
              [(-79 - 31 * _S6) / 150, (-1 - 4 * _S6) / 30, (24 + 11 * _S6) / 25, 0, 0],
              [(14 + 5 * _S6) / 6, (-8 + 7 * _S6) / 6, (-9 - 7 * _S6) / 4, (9 - _S6) / 4, 0]],
             [0, 0, (16 - _S6) / 36, (16 + _S6) / 36, 1 / 9], None, 0),                             # :253-265
    "RK33C2": ([[0, 0, 0], [2 / 3, 0, 0], [1 / 3, 1 / 3, 0]], [0.25, 0, 0.75], [0, 2 / 3, 2 / 3], 0),  # :497-503
    "RK44": ([[0, 0, 0, 0], [0.5, 0, 0, 0], [0, 0.5, 0, 0], [0, 0, 1, 0]], [1 / 6, 1 / 3, 1 / 3, 1 / 6], None, 0),  # :515-521
    "RK44P2": ([[0, 0, 0, 0], [1, 0, 0, 0], [3 / 8, 1 / 8, 0, 0], [-1 / 8, -3 / 8, 1.5, 0]], [1 / 6, -1 / 18, 2 / 3, 2 / 9], [0, 1, 0.5, 1], 0),  # :542-548
    "NSSP53": ([[0] * 5, [1 / 7, 0, 0, 0, 0], [0, 3 / 16, 0, 0, 0], [0, 0, 1 / 3, 0, 0], [0, 0, 0, 2 / 3, 0]],
               [0.25, 0, 0, 0, 0.75], [0, 1 / 7, 3 / 16, 1 / 3, 2 / 3], 0),                          # :627-641
}


def explicit_method(name, T=np.float64):
    A, b, c, r = EXPLICIT_TABLEAUX[name]
    return runge_kutta_method(A, b, np.sum(np.array(A, dtype=np.float64), axis=1) if c is None else c, r, T=T)


def create_stepper(method, setup, psolver, u, temp=None, t=0.0, n=0):
    """create_stepper src/time_steppers/step_explicit_runge_kutta.jl:1-2."""
    return SimpleNamespace(setup=setup, psolver=psolver, u=u, temp=temp, t=setup.grid.T(t), n=n)


def timestep(method, stepper, dt, theta=None, return_p=False):
    """timestep!(::ExplicitRungeKuttaMethod, ...) src/time_steppers/step_explicit_runge_kutta.jl:4-59."""
    setup, psolver, u, t, n = stepper.setup, stepper.psolver, stepper.u, stepper.t, stepper.n
    temp = getattr(stepper, "temp", None)
    T = setup.grid.T
    dt = T(dt)
    A, b, c = method.A, method.b, method.c
    nstage = len(b)
    m = setup.closure_model
    tstart = t
    ustart = u.copy(order="F")
    tempstart = None if temp is None else temp.copy(order="F")
    ku, ktemp = [], []
    p = None
    for i in range(nstage):
        u = apply_bc_u(u, t, setup, copy=False)
        if temp is not None:
            temp = apply_bc_temp(temp, t, setup, copy=False)
        k = momentum(u, temp, t, setup)
        if temp is not None:
            kt = convection_diffusion_temp(scalarfield(setup), u, temp, setup)   # :23-24
            if setup.temperature.dodissipation:
                kt = dissipation(kt, u, setup)                                    # :25
            ktemp.append(kt)
        if m is not None:
            k = k + m(u, theta)
        ku.append(k)
        t = T(tstart + c[i] * dt)
        u = ustart.copy(order="F")
        for j in range(i + 1):
            u = u + T(dt * A[i, j]) * ku[j]
        if temp is not None:
            temp = tempstart.copy(order="F")
            for j in range(i + 1):
                temp = temp + T(dt * A[i, j]) * ktemp[j]
        u = apply_bc_u(u, t, setup, copy=False)
        u, p = project(u, setup, psolver, return_p=True)
    u = apply_bc_u(u, t, setup, copy=False)
    if temp is not None:
        temp = apply_bc_temp(temp, t, setup, copy=False)
    out = create_stepper(method, setup, psolver, u, temp, t, n + 1)
    return (out, p) if return_p else out


def right_hand_side(u, t, setup, psolver):
    """right_hand_side! src/sciml.jl:37-49 (== create_right_hand_side :13-19): apply_bc_u!(copy(u)),
    momentum!, apply_bc_u!(dudt; dudt=true) (periodic: ghost fill), project!."""
    v = apply_bc_u(u, t, setup)
    F = momentum(v, None, t, setup)
    F = apply_bc_u(F, t, setup, copy=False)
    return project(F, setup, psolver)


def solve_unsteady(setup, tlims, ustart, method=None, psolver=None, dt=None, dt_min=None,
                   cfl=0.9, n_adapt_dt=1, theta=None, processors=None, nstep_max=None):
    """solve_unsteady src/solver.jl:18-92.  `processors` is a dict name -> callable(state)
    invoked after every step (the Observable of src/processors.jl:39-40)."""
    T = setup.grid.T
    method = RK44(T) if method is None else method
    psolver = psolver_spectral(setup) if psolver is None else psolver
    tstart, tend = T(tlims[0]), T(tlims[1])
    stepper = create_stepper(method, setup, psolver, ustart.copy(order="F"), None, tstart)
    processors = processors or {}
    outputs = {k: [] for k in processors}

    def fire():
        for k, f in processors.items():
            outputs[k].append(f(stepper))

    fire()
    if dt is None:
        cfl = T(cfl)
        d = None
        while stepper.t < tend:
            if stepper.n % n_adapt_dt == 0:
                d = T(cfl * get_cfl_timestep(stepper.u, setup))
                d = d if dt_min is None else max(d, T(dt_min))
            d = min(d, T(tend - stepper.t))
            stepper = timestep(method, stepper, d, theta=theta)
            fire()
            if nstep_max is not None and stepper.n >= nstep_max:
                break
    else:
        nstep = int(round(float(tend - tstart) / float(dt)))
        d = T((tend - tstart) / T(nstep))
        for _ in range(nstep):
            stepper = timestep(method, stepper, d, theta=theta)
            fire()
    return SimpleNamespace(u=stepper.u, t=stepper.t, n=stepper.n), outputs


# --------------------------------------------------------------------------
# Initial conditions (src/initializers.jl)
# --------------------------------------------------------------------------

def velocityfield(setup, ufunc, t=0.0, psolver=None, doproject=True):
    """velocityfield src/initializers.jl:13-46.  `ufunc(dim, x...)` is vectorised over
    broadcastable coordinate arrays, dim is 1-based."""
    g = setup.grid
    D, T = g.dimension, g.T
    u = vectorfield(setup)
    for a in range(D):
        xin = [_ax(g.xu[a][b][g.Iu[a][b]], b, D) for b in range(D)]
        u[g.Iu[a] + (a,)] = np.broadcast_to(np.asarray(ufunc(a + 1, *xin), dtype=T), g.Np)
    u = apply_bc_u(u, t, setup, copy=False)
    if doproject:
        psolver = psolver_spectral(setup) if psolver is None else psolver
        # allocating `project` (pressure.jl:52-66): ((div*dx)*dy)*dz, then u .- G
        p = divergence(u, setup)
        p = scalewithvolume(p, setup, inplace_form=False)
        p = poisson(psolver, p)
        p = apply_bc_p(p, T(0), setup, copy=False)
        u = u - pressuregradient(p, setup)
        u = apply_bc_u(u, t, setup, copy=False)
    return u


def create_spectrum(setup, kp, rng):
    """create_spectrum src/initializers.jl:82-181.  Random numbers come from `rng`
    (numpy Generator) -- Julia's Xoshiro stream cannot be reproduced here, parity runs feed
    identical *arrays* to both implementations."""
    g = setup.grid
    D, T, N = g.dimension, g.T, g.N
    ctype = np.complex64 if T == np.float32 else np.complex128
    tau = T(2 * np.pi)
    assert all(n % 2 == 0 for n in N)
    K = tuple((n - 2) // 2 for n in N)
    kk = [_ax(np.arange(K[a]).astype(T), a, D) for a in range(D)]
    k = np.zeros(K, dtype=T)
    for a in range(D):
        k = k + kk[a] ** 2
    k = np.sqrt(k)
    A = T(T(8) * tau / T(3)) / T(kp) ** 5
    a_ = np.sqrt(A * k ** 4 * np.exp(-tau * (k / T(kp)) ** 2)).astype(ctype)
    a_ = a_ * T(np.prod(N))
    xi = [rng.random(K).astype(T) for _ in range(D)]
    for al in range(D):
        a_ = np.concatenate((a_, np.flip(a_, axis=al)), axis=al)
        xi = [
            np.concatenate((xi[b], np.flip((-1 if al == b else 1) * xi[b], axis=al)), axis=al)
            for b in range(D)
        ]
    xis = xi[0]
    for b in range(1, D):
        xis = xis + xi[b]
    a_ = (np.exp(1j * (tau * xis).astype(T)).astype(ctype) * a_).astype(ctype)
    KK = tuple(2 * kk_ for kk_ in K)
    kkkk = [_ax(np.arange(KK[a]).astype(T), a, D) for a in range(D)]
    knorm = np.zeros(KK, dtype=T)
    for a in range(D):
        knorm = knorm + kkkk[a] ** 2
    knorm = np.sqrt(knorm)
    if D == 2:
        th = rng.random(KK).astype(T)
        e = [np.cos(np.pi * 2 * th).astype(T), np.sin(np.pi * 2 * th).astype(T)]
    else:
        th = rng.random(KK).astype(T)
        ph = rng.random(KK).astype(T)
        e = [
            (np.sin(np.pi * th) * np.cos(2 * np.pi * ph)).astype(T),
            (np.sin(np.pi * th) * np.sin(2 * np.pi * ph)).astype(T),
            np.cos(np.pi * th).astype(T),
        ]
    ke = sum(e[a] * kkkk[a] for a in range(D))
    with np.errstate(divide="ignore", invalid="ignore"):
        for a in range(D):
            e0 = e[a][(0,) * D]
            e[a] = (e[a] - kkkk[a] * ke / knorm ** 2).astype(T)
            e[a][(0,) * D] = e0
    enorm = np.sqrt(sum(e[a] ** 2 for a in range(D)))
    e = [e[a] / enorm for a in range(D)]
    return np.stack([a_ * e[a] for a in range(D)], axis=-1).astype(ctype)


def random_field(setup, t=0.0, A=1, kp=10, psolver=None, rng=None):
    """random_field src/initializers.jl:189-219."""
    g = setup.grid
    D, T = g.dimension, g.T
    rng = np.random.default_rng(0) if rng is None else rng
    psolver = psolver_spectral(setup) if psolver is None else psolver
    uhat = create_spectrum(setup, kp, rng)
    u = sfft.ifftn(uhat, axes=tuple(range(D)))
    u = (T(A) * u.real).astype(T)
    pad = [(1, 1)] * D + [(0, 0)]
    u = np.asfortranarray(np.pad(u, pad, mode="wrap"))
    u = apply_bc_u(u, t, setup, copy=False)
    p = divergence(u, setup)
    p = scalewithvolume(p, setup, inplace_form=False)
    p = poisson(psolver, p)
    p = apply_bc_p(p, T(0), setup, copy=False)
    u = u - pressuregradient(p, setup)
    return apply_bc_u(u, t, setup, copy=False)
