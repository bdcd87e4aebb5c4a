"""World-size-2 (gloo, CPU) tests of the slab decomposition the multi-GPU path uses:
  * ghost planes of the decomposed axis by ring send/recv == the reference's periodic apply_bc on the whole field
  * the distributed spectral Poisson solve -- in-slab transforms, pack by kx chunk, all-to-all, last-axis
    transform, symbol scaling with global wavenumbers, and back -- == psolver_spectral on the whole field.
  * the transpose-free last-axis solve (csrc/zline.cuh, the default of the GPU path): pass A per rank and sub-chunk, an
    all_gather of the two chunk sums per mode (the only exchange), the chain over the chunks, pass B -- == psolver_spectral.
The array arithmetic here is NumPy (the product's is CUDA + NVLink peer stores / NCCL, tested on GPUs); the
DECOMPOSITION -- slab bounds and the kx ranges of the transposed spectrum -- is the product's own
`insb200_decomposition` (pure host arithmetic of libinsb200.so, the function `insb200_create` uses), so that this test
pins the message layout and the global index offsets the CUDA path relies on."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ins_oracle as o


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def slab_bounds(n_last, rank, nranks):
    """The PRODUCT's decomposition (insb200_decomposition, the host arithmetic insb200_create uses; no GPU needed)."""
    import insb200 as ins

    return ins.slab_bounds(n_last, rank, nranks)


def kx_chunk(nxh, rank, nranks):
    """(first kx, count) of `rank` in the transposed spectrum -- again the product's own arithmetic."""
    import insb200 as ins

    _, _, kx0, kxl = ins.decomposition(2 * (nxh - 1), nranks, rank, nranks)
    return kx0, kxl


def _worker(rank, world, port, D, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        T = np.float64
        x = tuple(np.linspace(0.0, 1.0, m + 1) for m in n)
        setup = o.Setup(x, Re=T(100))
        g = setup.grid
        rng = np.random.default_rng(7)
        u = np.asfortranarray(rng.standard_normal(g.N + (D,)))
        f = np.asfortranarray(rng.standard_normal(g.N))
        lo, hi = slab_bounds(n[D - 1], rank, world)
        prev, nxt = (rank - 1) % world, (rank + 1) % world
        # ---- halo: local slab with ghosts; x/y ghosts filled locally, last-axis ghosts over the ring
        idx = [slice(None)] * (D + 1)
        idx[D - 1] = slice(lo, hi + 2)
        ul = np.ascontiguousarray(np.moveaxis(u[tuple(idx)], D - 1, 0))  # last axis first: planes contiguous
        for b in range(D - 1):  # local axes (axis b of the field = axis b+1 here)
            sl = lambda k: tuple([slice(None)] * (b + 1) + [k])
            ul[sl(0)] = ul[sl(-2)]
            ul[sl(-1)] = ul[sl(1)]
        send_top, send_bot = torch.from_numpy(ul[-2].copy()), torch.from_numpy(ul[1].copy())
        recv_bot, recv_top = torch.empty_like(send_top), torch.empty_like(send_bot)
        reqs = [dist.isend(send_top, nxt), dist.isend(send_bot, prev), dist.irecv(recv_bot, prev), dist.irecv(recv_top, nxt)]
        for r in reqs:
            r.wait()
        ul[0], ul[-1] = recv_bot.numpy(), recv_top.numpy()
        want = np.moveaxis(o.apply_bc_u(u, 0.0, setup)[tuple(idx)], D - 1, 0)
        ok_halo = np.array_equal(ul, want)
        # ---- distributed spectral solve of f[Ip]
        fI = f[g.Ip]
        idx = [slice(None)] * D
        idx[D - 1] = slice(lo, hi)
        fl = np.moveaxis(fI[tuple(idx)], D - 1, 0)           # (zl, x[, y]) Julia axis order after the first
        fl = np.ascontiguousarray(np.transpose(fl, [0] + list(range(D - 1, 0, -1))))  # (zl, [y,] x): x fastest
        S = np.fft.rfftn(fl, axes=tuple(range(1, D)))        # in-slab transform, kx halved & fastest
        nxh = n[0] // 2 + 1
        kx0, kxl = kx_chunk(nxh, rank, world)
        send = []
        for qk in range(world):
            qlo, qw = kx_chunk(nxh, qk, world)
            send.append(torch.from_numpy(np.ascontiguousarray(S[..., qlo:qlo + qw])))
        recv = [torch.empty((hi - lo,) + S.shape[1:-1] + (kxl,), dtype=send[0].dtype) for _ in range(world)]
        ops = []
        for qk in range(world):
            if qk == rank:
                recv[qk].copy_(send[qk])
            else:
                ops += [dist.isend(send[qk], qk), dist.irecv(recv[qk], qk)]
        for r in ops:
            r.wait()
        R = np.concatenate([t.numpy() for t in recv], axis=0)  # (z global, [ky,] kxl)
        R = np.fft.fft(R, axis=0)
        ps = o.psolver_spectral(setup)
        ah = [a.astype(np.float64) for a in ps.ahat]
        den = ah[0][kx0:kx0 + kxl]
        if D == 3:
            den = den[None, None, :] + ah[1][None, :, None] + ah[2][:, None, None]
        else:
            den = den[None, :] + ah[1][:, None]
        with np.errstate(divide="ignore", invalid="ignore"):
            R = -R / den
        if kx0 == 0:
            R[(0,) * D] = 0
        R = np.fft.ifft(R, axis=0)
        back_send = [torch.from_numpy(np.ascontiguousarray(R[slab_bounds(n[D - 1], qk, world)[0]:slab_bounds(n[D - 1], qk, world)[1]]))
                     for qk in range(world)]
        back_recv = [torch.empty((hi - lo,) + S.shape[1:-1] + (kx_chunk(nxh, qk, world)[1],), dtype=send[0].dtype) for qk in range(world)]
        ops = []
        for qk in range(world):
            if qk == rank:
                back_recv[qk].copy_(back_send[qk])
            else:
                ops += [dist.isend(back_send[qk], qk), dist.irecv(back_recv[qk], qk)]
        for r in ops:
            r.wait()
        S2 = np.concatenate([t.numpy() for t in back_recv], axis=-1)
        pl = np.fft.irfftn(S2, s=fl.shape[1:], axes=tuple(range(1, D)))
        pw = ps(f)[g.Ip]
        wl = np.moveaxis(pw[tuple(idx)], D - 1, 0)
        wl = np.transpose(wl, [0] + list(range(D - 1, 0, -1)))
        err = float(np.abs(pl - wl).max() / np.abs(pw).max())
        q.put((rank, ok_halo, err))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("D,n", [(2, (16, 12)), (3, (8, 6, 8))])
def test_slab_decomposition_world2(D, n):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, D, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok_halo, err in res:
        assert ok_halo, f"rank {rank}: ring halo != periodic apply_bc"
        assert err < 1e-12, f"rank {rank}: distributed Poisson solve error {err}"


# ---- the transpose-free last-axis solve (csrc/zline.cuh) on two processes ------------------------------------------------
def _line_worker(rank, world, port, D, n, C, q):
    """Each rank holds its slab of the in-plane spectrum, runs pass A on its C sub-chunks, the ranks exchange ONLY the chunk
    sums A, B (and the constant mode's column) -- here an all_gather over gloo, in the product peer loads -- and pass B
    finishes the solve locally.  Must equal psolver_spectral on the whole field."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        x = tuple(np.linspace(0.0, 1.5, m + 1) for m in n)
        setup = o.Setup(x, Re=100.0)
        g = setup.grid
        f = np.asfortranarray(np.random.default_rng(11).standard_normal(g.N))
        ps = o.psolver_spectral(setup)
        want = ps(f)[g.Ip]
        lo, hi = slab_bounds(n[D - 1], rank, world)
        sl = (slice(None),) * (D - 1) + (slice(lo, hi),)
        fl = f[g.Ip][sl]                                   # this rank's planes only
        F = np.fft.rfft(fl, axis=0)
        if D == 3:
            F = np.fft.fft(F, axis=1)
        ah = [a.astype(np.float64) for a in ps.ahat]
        axy = (ah[0][:, None] + ah[1][None, :]) if D == 3 else ah[0]
        dx = [float(g.Delta[a][0]) for a in range(D)]
        kappa = dx[-1] ** 2 / float(np.prod(dx))
        nl, ntot = hi - lo, n[D - 1]
        mv, V = nl // C, world * C
        H = (kappa * F).reshape(-1, nl)
        lam = (kappa * axy).reshape(-1)
        lam[0] = 1.0                                        # placeholder: mode 0 is solved separately
        lr = -np.log1p(lam / 2 + np.sqrt(lam * (1 + lam / 4)))
        rho = np.exp(lr)[:, None]
        hc = H.reshape(-1, C, mv)
        g0 = np.zeros_like(hc)
        s = np.zeros(hc.shape[:2], complex)
        B = np.zeros(hc.shape[:2], complex)
        pw = np.ones((hc.shape[0], 1))
        for j in range(mv):                                 # pass A
            s = hc[:, :, j] + rho * s
            g0[:, :, j] = s
            B = B + pw * hc[:, :, j]
            pw = pw * rho
        mine = torch.from_numpy(np.ascontiguousarray(np.stack([s, B])))          # [2][modes][C]
        allab = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allab, mine)                        # the ONLY data exchange besides the constant mode's column
        AB = np.concatenate([t.numpy() for t in allab], axis=2)                  # [2][modes][V]
        col = torch.from_numpy(np.ascontiguousarray(H[0]))
        cols = [torch.empty_like(col) for _ in range(world)]
        dist.all_gather(cols, col)
        rm = np.exp(mv * lr)
        inv1n = -1 / np.expm1(ntot * lr)
        inv1r2 = (-1 / np.expm1(2 * lr))[:, None]
        Gall = np.zeros(AB.shape[1:], complex)
        GAall = np.zeros(AB.shape[1:], complex)
        for v in range(V):
            for k in range(V):
                Gall[:, v] += rm ** k * AB[0][:, (v - k) % V]
                GAall[:, v] += rm ** k * AB[1][:, (v + k) % V]
        Gall *= inv1n[:, None]
        GAall *= inv1n[:, None]
        mych = [rank * C + c for c in range(C)]
        Gin = Gall[:, [(v - 1) % V for v in mych]]
        qs = (rho * Gall[:, mych] + GAall[:, [(v + 1) % V for v in mych]]) * inv1r2
        P = np.zeros_like(hc)
        for j in range(mv - 1, -1, -1):                     # pass B
            qs = g0[:, :, j] + rho ** (j + 1) * Gin + rho * qs
            P[:, :, j] = -rho * qs
        P = P.reshape(-1, nl)
        hcol = np.concatenate([t.numpy() for t in cols])
        d = (np.arange(lo, hi)[:, None] - np.arange(ntot)[None, :]) % ntot
        P[0] = (d * (ntot - d) / (2.0 * ntot)) @ (hcol - hcol.mean())
        P = P.reshape(F.shape)
        if D == 3:
            P = np.fft.ifft(P, axis=1)
        got = np.fft.irfft(P, n=n[0], axis=0)
        q.put((rank, float(np.abs(got - want[sl]).max() / np.abs(want).max())))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("D,n,C", [(2, (16, 24), 1), (3, (8, 6, 16), 2), (3, (12, 8, 12), 3)])
def test_line_solve_world2(D, n, C):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_line_worker, args=(r, 2, port, D, n, C, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err in res:
        assert err < 1e-12, f"rank {rank}: line solve on slabs differs from psolver_spectral by {err}"
