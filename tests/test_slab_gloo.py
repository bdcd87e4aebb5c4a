"""World-size-2 (gloo, CPU) tests of the slab decomposition the multi-GPU path uses:
  * ghost planes of the decomposed axis by ring send/recv == the reference's periodic apply_bc on the whole field
  * the distributed spectral Poisson solve -- in-slab transforms, pack by kx chunk, all-to-all, last-axis
    transform, symbol scaling with global wavenumbers, and back -- == psolver_spectral on the whole field.
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
