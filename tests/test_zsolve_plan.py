"""Host-side restatement of the last-axis solve of `k_zsolve` (csrc/zsolve.cuh, `zsolve_prepare` in csrc/insb200.cu):
the radix plan, the in-place DIF passes (natural order in, digit-reversed out), the symbol looked up at
digit-reversed positions and the adjoint DIT passes.  Pins the index logic of the kernel -- in particular the
mixed-radix digit reversal with the radix-16 passes -- against numpy's FFT on the CPU; the kernel itself is compared
with cuFFT and the oracle in tests/test_gpu_operators.py::test_zsolve_matches_cufft_and_oracle.
"""
import numpy as np
import pytest


def radix_plan(nz, f32=True, radix16=True):
    """log2 of the radices of the forward passes (zsolve_prepare)."""
    m = int(np.log2(nz))
    assert 1 << m == nz
    if f32 and radix16:
        nd = (m + 3) // 4
        base, rem = divmod(m, nd)
        return [base + (1 if i >= nd - rem else 0) for i in range(nd)]
    plan = [3] * (m // 3)
    if m % 3:
        plan.append(m % 3)
    return plan


def digit_reversed_mode(t, plan, lnz):
    """mode kz held at position t after the forward passes (zsolve_tables)."""
    kz, lspan, lw = 0, lnz, 0
    for r in plan:
        lspan -= r
        kz += ((t >> lspan) & ((1 << r) - 1)) << lw
        lw += r
    return kz


def zpass(s, nz, lr, lsub, inv):
    """one in-place pass over s[nz, cols] for blocks of length (2^lr) << lsub (zpass in zsolve.cuh)."""
    R, sub = 1 << lr, 1 << lsub
    L = R * sub
    tstep = nz // L
    tw = np.exp(-2j * np.pi * np.arange(nz) / nz)
    dft = np.exp((2j if inv else -2j) * np.pi * np.outer(np.arange(R), np.arange(R)) / R)
    out = s.copy()
    for b in range(nz // L):
        for j in range(sub):
            idx = b * L + j + sub * np.arange(R)
            v = s[idx]
            w = tw[(j * tstep * np.arange(R)) % nz][:, None]
            if inv:
                v = dft @ (v * np.conj(w))
            else:
                v = (dft @ v) * w
            out[idx] = v
    return out


def zsolve_tile(s, sym, plan):
    """forward passes, symbol at digit-reversed positions, inverse passes (zsolve_tile); sym[kz] multiplies mode kz."""
    nz = s.shape[0]
    lnz = int(np.log2(nz))
    lL = lnz
    for r in plan:  # the kernel fuses the last forward pass, the symbol and the first inverse pass in registers
        s = zpass(s, nz, r, lL - r, False)
        lL -= r
    pos_mode = np.array([digit_reversed_mode(t, plan, lnz) for t in range(nz)])
    s = s * sym[pos_mode][:, None]
    for r in reversed(plan):
        s = zpass(s, nz, r, lL, True)
        lL += r
    return s


def test_plans():
    assert radix_plan(512) == [3, 3, 3]
    assert radix_plan(1024) == [3, 3, 4]
    assert radix_plan(256) == [4, 4]
    assert radix_plan(2048) == [3, 4, 4]
    assert radix_plan(16) == [4] and radix_plan(8) == [3]
    assert radix_plan(1024, f32=False) == [3, 3, 3, 1]
    assert radix_plan(128, f32=False) == [3, 3, 1]
    for m in range(1, 13):
        for f32 in (True, False):
            p = radix_plan(1 << m, f32)
            assert sum(p) == m and all(1 <= r <= (4 if f32 else 3) for r in p)


@pytest.mark.parametrize("nz", [2, 4, 8, 16, 32, 64, 128, 256, 1024])
@pytest.mark.parametrize("f32", [True, False], ids=["plan_f32", "plan_f64"])
def test_dif_symbol_dit_equals_fft_filter(nz, f32):
    rng = np.random.default_rng(nz)
    cols = 3
    x = rng.standard_normal((nz, cols)) + 1j * rng.standard_normal((nz, cols))
    sym = rng.standard_normal(nz)
    plan = radix_plan(nz, f32)
    got = zsolve_tile(x, sym, plan)
    want = np.fft.ifft(np.fft.fft(x, axis=0) * sym[:, None], axis=0) * nz  # the kernel folds 1/N into the symbol
    assert np.abs(got - want).max() <= 1e-11 * nz * np.abs(want).max()
    # the forward passes alone leave mode kz at its digit-reversed position
    s, lL = x.copy(), int(np.log2(nz))
    for r in plan:
        s = zpass(s, nz, r, lL - r, False)
        lL -= r
    pos_mode = np.array([digit_reversed_mode(t, plan, int(np.log2(nz))) for t in range(nz)])
    assert sorted(pos_mode) == list(range(nz))
    assert np.abs(s - np.fft.fft(x, axis=0)[pos_mode]).max() <= 1e-11 * nz * np.abs(x).max()
