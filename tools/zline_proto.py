"""Prototype of the distributed last-axis line solve (recursive filters instead of a distributed FFT).

    p[j-1] - (2 + lam) p[j] + p[j+1] = h[j]   (periodic, n = V * mv points in V chunks)

is solved chunk by chunk: pass 0 forms two weighted sums per chunk, the chunks exchange them, pass 1 (causal) and
pass 2 (anticausal) run a first-order recursion each.  Checked here against the circulant solve by FFT.
"""
import numpy as np

def solve_fft(h, lam):
    n = len(h)
    k = np.arange(n)
    sym = -(lam + 4 * np.sin(np.pi * k / n) ** 2)
    H = np.fft.fft(h)
    if lam == 0:
        sym[0] = 1; H[0] = 0
    return np.fft.ifft(H / sym)

def solve_chunks(h, lam, V):
    n = len(h); mv = n // V
    hc = h.reshape(V, mv)
    lr = -np.log1p(lam / 2 + np.sqrt(lam * (1 + lam / 4)))
    rho = np.exp(lr)
    A = np.zeros(V, complex); B = np.zeros(V, complex)
    for v in range(V):
        a = 0; b = 0; pw = 1.0
        for i in range(mv):
            a = a * rho + hc[v, i]; b += pw * hc[v, i]; pw *= rho
        A[v] = a; B[v] = b
    rm = np.exp(mv * lr); inv1n = -1 / np.expm1(n * lr); inv1r2 = -1 / np.expm1(2 * lr)
    def G(v):   # true causal state at the end of chunk v
        s = 0; w = 1.0
        for k in range(V):
            if w < 1e-30: break
            s += w * A[(v - k) % V]; w *= rm
        return s * inv1n
    def GA(v):  # anticausal filter of h at the first point of chunk v
        s = 0; w = 1.0
        for k in range(V):
            if w < 1e-30: break
            s += w * B[(v + k) % V]; w *= rm
        return s * inv1n
    g = np.zeros((V, mv), complex); p = np.zeros((V, mv), complex)
    for v in range(V):
        s = G(v - 1)
        for j in range(mv):
            s = hc[v, j] + rho * s; g[v, j] = s
    for v in range(V):
        q = (rho * G(v) + GA((v + 1) % V)) * inv1r2
        for j in range(mv - 1, -1, -1):
            q = g[v, j] + rho * q; p[v, j] = -rho * q
    return p.reshape(n)

def solve_zero(h):
    n = len(h)
    r = h - h.mean()
    d = np.cumsum(r)            # p[j+1] - p[j] = d[j] + c
    c = -d.sum() / n
    p = np.concatenate([[0], np.cumsum(d[:-1] + c)])
    return p - p.mean()

rng = np.random.default_rng(1)
for n, V in [(64, 1), (64, 4), (128, 8), (1024, 8), (1024, 32)]:
    h = rng.standard_normal(n) + 1j * rng.standard_normal(n)
    for lam in [3.8e-5, 1e-3, 0.05, 1.0, 8.0, 200.0]:
        ref = solve_fft(h, lam); got = solve_chunks(h, lam, V)
        print(n, V, lam, np.abs(got - ref).max() / np.abs(ref).max())
    ref = solve_fft(h, 0.0); got = solve_zero(h)
    print(n, V, 'zero', np.abs(got - ref).max() / np.abs(ref).max())
