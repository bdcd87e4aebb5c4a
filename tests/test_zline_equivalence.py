"""The transpose-free last-axis solve (csrc/zline.cuh) against the ORACLE's psolver_spectral (src/pressure.jl:289-351), in
NumPy and Float64: transform the in-plane axes, then per in-plane mode solve the periodic tridiagonal system by the two
recursions over chunks of the last axis exactly as the kernels do (pass A from a zero state with the chunk sums A and B,
chain over the chunks, pass B backwards with the neighbour correction, Green's function for the constant mode), transform
back.  The result must be the oracle's pressure to round-off: dividing by the reference's symbol and solving this system
are the same operation.  (The kernels' own code is checked by tests/test_zline_emul.py; this test ties the algorithm to the
oracle, which is pinned to the reference's known answers.)"""
import numpy as np
import pytest

from oracle import ins_oracle as o
from tests.util import axes, rand_p


def line_solve(h, lam, V):
    """h[mode, n]: right-hand sides of p[j-1] - (2 + lam) p[j] + p[j+1] = h[j] (periodic), lam > 0 per mode; V chunks."""
    nm, n = h.shape
    mv = n // V
    lr = -np.log1p(lam / 2 + np.sqrt(lam * (1 + lam / 4)))
    rho = np.exp(lr)[:, None]
    hc = h.reshape(nm, V, mv)
    g0 = np.zeros_like(hc)
    A = np.zeros((nm, V), complex)
    B = np.zeros((nm, V), complex)
    s = np.zeros((nm, V), complex)
    pw = np.ones((nm, 1))
    for j in range(mv):  # pass A
        s = hc[:, :, j] + rho * s
        g0[:, :, j] = s
        B += pw * hc[:, :, j]
        pw = pw * rho
    A = s
    rm = np.exp(mv * lr)[:, None]
    inv1n = (-1 / np.expm1(n * lr))[:, None]
    inv1r2 = (-1 / np.expm1(2 * lr))[:, None]
    G = np.zeros((nm, V), complex)   # true causal state at the end of chunk v
    GA = np.zeros((nm, V), complex)  # anticausal filter of h at the first point of chunk v
    for v in range(V):
        for k in range(V):
            G[:, v] += (rm[:, 0] ** k) * A[:, (v - k) % V]
            GA[:, v] += (rm[:, 0] ** k) * B[:, (v + k) % V]
    G *= inv1n
    GA *= inv1n
    Gin = np.roll(G, 1, axis=1)                      # state entering chunk v from below
    q = (rho * G + np.roll(GA, -1, axis=1)) * inv1r2  # state entering chunk v from above
    p = np.zeros_like(hc)
    for j in range(mv - 1, -1, -1):  # pass B
        q = g0[:, :, j] + rho ** (j + 1) * Gin + rho * q
        p[:, :, j] = -rho * q
    return p.reshape(nm, n)


def zero_mode(h):
    n = len(h)
    d = (np.arange(n)[:, None] - np.arange(n)[None, :]) % n
    return (d * (n - d) / (2.0 * n)) @ (h - h.mean())


@pytest.mark.parametrize("n,V", [((16, 12, 32), 1), ((16, 12, 32), 4), ((12, 8, 48), 6), ((24, 36), 4)], ids=str)
def test_line_solve_is_the_oracle_pressure_solve(n, V):
    so = o.Setup(axes(n, np.float64, (0.0, 1.5)), Re=1000.0)
    g = so.grid
    D = len(n)
    f = rand_p(so, 5, np.float64)
    psol = o.psolver_spectral(so)
    want = o.poisson(psol, f)[g.Ip]
    fI = f[g.Ip]
    F = np.fft.rfft(fI, axis=0)
    if D == 3:
        F = np.fft.fft(F, axis=1)
    ahat = [a.astype(np.float64) for a in psol.ahat]
    axy = ahat[0][:, None] + ahat[1][None, :] if D == 3 else ahat[0]
    dx = [float(g.Delta[a][0]) for a in range(D)]
    kappa = dx[-1] ** 2 / float(np.prod(dx))
    H = (kappa * F).reshape(-1, n[-1])   # cuFFT leaves the in-plane transforms unnormalised; here numpy's inverse divides
    lam = (kappa * axy).reshape(-1)
    P = np.zeros_like(H)
    P[1:] = line_solve(H[1:], lam[1:], V)
    P[0] = zero_mode(H[0])
    P = P.reshape(F.shape)
    if D == 3:
        P = np.fft.ifft(P, axis=1)
    got = np.fft.irfft(P, n=n[0], axis=0)
    err = np.linalg.norm(got - want) / np.linalg.norm(want)
    assert err < 1e-12, err
