"""The FAST arithmetic of the fused stage kernel (csrc/fast_stage.cuh), restated in NumPy and checked against the
oracle's `momentum` and `smagorinsky_closure` (src/operators.jl:647-690, 1135-1151, 1205-1238).

The fast kernel evaluates the same operators in a different association order (allowed by north_star's tolerance,
rel-L2 <= 1e-5 / 1e-12; the bit-exact kernels stay selectable with set_param("exact", 1)):
  * convection-diffusion in FLUX form: the flux through a cell face is formed once and differenced,
    diffusion as (up + um - 2 uI), constants visc/h^2 and 1/(4h) folded;
  * the velocity gradient of the closure with the four one-sided differences telescoped into
    ((u(+b) - u(-b)) + (u(-a+b) - u(-a-b))) / (4 h_b);
  * the divergence of the tensor with the corner averages telescoped the same way.
In Float64 the reformulation agrees with the oracle to a few ulp, which is what this test pins; the CUDA kernel is
compared with the oracle on the GPU (tests/test_gpu_fast.py).
"""
import numpy as np
import pytest

from oracle import ins_oracle as o


def sh(a, dx=0, dy=0, dz=0):
    """a[I + (dx, dy, dz)] on the periodic interior."""
    return np.roll(a, (-dx, -dy, -dz), axis=(0, 1, 2))


def fast_convdiff(u, h, visc):
    """F_a = sum_b [ visc (up + um - 2 uI) / h_b^2 - (x2 - x1) / (4 h_b) ] with the products of fast_stage.cuh."""
    ux, uy, uz = u
    rh = [1.0 / v for v in h]
    vrr = [visc * r * r for r in rh]
    q = [0.25 * r for r in rh]
    e = [(1, 0, 0), (0, 1, 0), (0, 0, 1)]
    F = []
    for a in range(3):
        ua = u[a]
        f = 0.0
        for b in range(3):
            ub = u[b]
            eb, ea = e[b], e[a]
            neg = tuple(-v for v in eb)
            amb = tuple(ea[i] - eb[i] for i in range(3))
            up, um = sh(ua, *eb), sh(ua, *neg)
            lap = up + um - 2.0 * ua
            x2 = (ua + up) * (ub + sh(ub, *ea))
            x1 = (um + ua) * (sh(ub, *neg) + sh(ub, *amb))
            f = f + vrr[b] * lap - q[b] * (x2 - x1)
        F.append(f)
    return F


def fast_sigma(u, h, theta):
    ux, uy, uz = u
    rh = [1.0 / v for v in h]
    e = [(1, 0, 0), (0, 1, 0), (0, 0, 1)]
    G = [[None] * 3 for _ in range(3)]
    for a in range(3):
        ua = u[a]
        ea = e[a]
        nea = tuple(-v for v in ea)
        ut = ua + sh(ua, *nea)  # sum over the two a-faces of the p-cell
        for b in range(3):
            eb = e[b]
            neb = tuple(-v for v in eb)
            if a == b:
                G[a][b] = (ua - sh(ua, *nea)) * rh[a]
            else:
                G[a][b] = (sh(ut, *eb) - sh(ut, *neb)) * (0.25 * rh[b])
    S = [[0.5 * (G[a][b] + G[b][a]) for b in range(3)] for a in range(3)]
    ss = S[0][0] ** 2 + S[1][1] ** 2 + S[2][2] ** 2 + 2.0 * (S[0][1] ** 2 + S[0][2] ** 2 + S[1][2] ** 2)
    d2 = h[0] ** 2 + h[1] ** 2 + h[2] ** 2
    eddy2 = 2.0 * theta ** 2 * d2 * np.sqrt(2.0 * ss)
    return [[eddy2 * S[a][b] for b in range(3)] for a in range(3)]


def fast_divsigma(sg, h):
    rh = [1.0 / v for v in h]
    e = [(1, 0, 0), (0, 1, 0), (0, 0, 1)]
    out = []
    for a in range(3):
        ea = e[a]
        s = 0.0
        for b in range(3):
            eb = e[b]
            t = sg[a][b]
            if a == b:
                s = s + (sh(t, *eb) - t) * rh[b]
            else:
                neb = tuple(-v for v in eb)
                apb = tuple(ea[i] + eb[i] for i in range(3))
                amb = tuple(ea[i] - eb[i] for i in range(3))
                s = s + ((sh(t, *eb) + sh(t, *apb)) - (sh(t, *neb) + sh(t, *amb))) * (0.25 * rh[b])
        out.append(s)
    return out


@pytest.mark.parametrize("n", [(16, 12, 10), (8, 8, 8)])
def test_fast_formulas_match_oracle(n):
    T = np.float64
    x = tuple(np.linspace(0.0, L, m + 1).astype(T) for m, L in zip(n, (1.0, 0.75, 1.25)))
    so = o.Setup(x, Re=T(300))
    rng = np.random.default_rng(3)
    u = o.apply_bc_u(np.asfortranarray(rng.standard_normal(so.grid.N + (3,))), 0.0, so)
    h = [float(so.grid.Delta[a][1]) for a in range(3)]
    ui = [np.array(u[1:-1, 1:-1, 1:-1, a]) for a in range(3)]
    F = o.momentum(u, None, 0.0, so)
    Ff = fast_convdiff(ui, h, 1.0 / 300.0)
    for a in range(3):
        ref = F[1:-1, 1:-1, 1:-1, a]
        assert np.abs(Ff[a] - ref).max() <= 1e-12 * np.abs(ref).max()
    theta = 0.17
    s = o.smagorinsky_closure(so)(u, T(theta))
    sg = fast_sigma(ui, h, theta)
    sf = fast_divsigma(sg, h)
    for a in range(3):
        ref = s[1:-1, 1:-1, 1:-1, a]
        assert np.abs(sf[a] - ref).max() <= 1e-12 * np.abs(ref).max()
