#!/usr/bin/env python
"""Multi-GPU parity script (run under torchrun on N GPUs of one box; not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/mgpu_parity.py

Every rank owns one slab of the last axis; the slab-decomposed CUDA path (NCCL halos, pack + all-to-all
distributed FFT) is compared with the single-domain oracle on identical initial arrays.  Exit code 0 = pass."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import insb200 as ins  # noqa: E402
from oracle import ins_oracle as o  # noqa: E402

CASES = [
    # n, dtype, force, theta, nsteps, dt
    ((32, 32), np.float64, True, None, 6, 2e-3),
    ((32, 16, 16), np.float64, False, None, 4, 2e-3),        # TMA path, slabs of 8 (or fewer) planes
    ((32, 16, 32), np.float32, True, 0.17, 4, 1e-3),         # TMA path + closure + force
    ((24, 20, 16), np.float64, False, 0.17, 3, 2e-3),        # generic path (not tile aligned)
    ((64, 32, 64), np.float32, False, 0.17, 2, 1e-3),        # 8 planes per rank at 8 GPUs, closure kernel v2
    ((96, 64), np.float32, True, None, 6, 5e-4),             # 2-D Float32 + Kolmogorov force: fused 2-D stage kernel on slabs
]


def kolmogorov(D):
    if D == 2:
        return lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x
    return lambda dim, x, y, z, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x + 0 * z


def tile_box(a, reps, D):
    """ghosted small field (n+2, .., [D]) -> ghosted field of the box tiled reps[a] times along axis a."""
    inner = a[(slice(1, -1),) * D]
    big = np.tile(inner, tuple(reps) + (1,) * (a.ndim - D))
    return np.asfortranarray(np.pad(big, [(1, 1)] * D + [(0, 0)] * (a.ndim - D), mode="wrap"))


def big_case(backend):
    """256 x 256 x 512 Float32 + Smagorinsky on the slab path against the oracle on 64^3: a field of period 1 on the
    box [0,4) x [0,4) x [0,8) with the same spacing evolves exactly like the field on [0,1)^3 (tests/test_gpu_fullsize.py),
    so the oracle steps the small box and every rank its slab of the tiled copy -- 64 or more planes per rank, the
    pipelined transposes, the 4-plane unrolled march and real chunk boundaries of the stage kernels."""
    rank, world = backend.rank, backend.nranks
    T, ns, reps, nsteps, dt, theta = np.float32, (64, 64, 64), (4, 4, 8), 2, 1e-3, 0.17
    xs = tuple(np.linspace(0.0, 1.0, m + 1).astype(T) for m in ns)
    xb = tuple(np.linspace(0.0, float(r), m * r + 1).astype(T) for m, r in zip(ns, reps))
    so = o.Setup(xs, Re=T(2000))
    sc = ins.Setup(xb, Re=T(2000), backend=backend)
    so.closure_model = o.smagorinsky_closure(so)
    sc.closure_model = ins.smagorinsky_closure(sc)
    for k in ("MGPU_P2P", "MGPU_FASTBAR", "MGPU_EXACT", "MGPU_TMODE", "MGPU_ZLINE", "MGPU_ZLCHUNKS"):
        if k in os.environ:
            sc.ctx.call("insb200_set_param", {"MGPU_P2P": b"p2p", "MGPU_FASTBAR": b"fast_barrier", "MGPU_EXACT": b"exact",
                                              "MGPU_TMODE": b"transpose_mode", "MGPU_ZLINE": b"zline", "MGPU_ZLCHUNKS": b"zline_chunks"}[k], float(os.environ[k]))
    u0 = o.random_field(so, kp=4, rng=np.random.default_rng(3))
    ul = ins.to_device(ins.local_slab(tile_box(u0, reps, 3), 3, rank, world), sc)
    m = ins.RKMethods.RK44(T=T)
    _, ke = ins.api._rk_steps(m, sc, ul, 0.0, dt, nsteps, theta, ke_hist=True)
    parts = [torch.empty_like(ul) for _ in range(world)]
    dist.all_gather(parts, ul.contiguous())
    ok = True
    if rank == 0:
        ug = ins.gather_slabs([ins.to_host(p) for p in parts], 3)
        mo = o.RK44(T)
        st = o.create_stepper(mo, so, o.psolver_spectral(so), u0.copy(order="F"))
        ke_o = [o.total_kinetic_energy(u0, so)]
        for _ in range(nsteps):
            st = o.timestep(mo, st, T(dt), theta=theta)
            ke_o.append(o.total_kinetic_energy(st.u, so))
        want = tile_box(st.u, reps, 3)
        err = o.rel_l2(ug, want)
        kerr = float(np.max(np.abs(ke / (np.array(ke_o, dtype=np.float64) * np.prod(reps)) - 1)))
        ok = err <= 1e-5 and kerr <= 1e-5
        print(f"[mgpu x{world}] BIG 256x256x512 float32 + Smagorinsky (tiled 64^3 oracle): rel_l2(u)={err:.3e} ke_err={kerr:.3e} "
              f"-> {'ok' if ok else 'FAIL'}", flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    sc.ctx.close()
    return 0 if ok else 1


def main():
    backend = ins.init_distributed()
    rank, world = backend.rank, backend.nranks
    torch.cuda.set_device(backend.device)
    failures = 0
    import faulthandler
    sel = os.environ.get("MGPU_CASES")
    cases = [CASES[int(i)] for i in sel.split(",")] if sel else CASES
    for n, T, force, theta, nsteps, dt in cases:
        faulthandler.dump_traceback_later(90, exit=True)  # a hung collective must not burn the GPU budget
        D = len(n)
        if n[D - 1] % world or n[D - 1] // world < 4:
            continue
        x = tuple(np.linspace(0.0, 1.0, m + 1).astype(T) for m in n)
        bf = kolmogorov(D) if force else None
        so = o.Setup(x, Re=T(800), bodyforce=bf)
        sc = ins.Setup(x, Re=T(800), bodyforce=bf, backend=backend)
        if "MGPU_P2P" in os.environ:  # A/B: 0 = NCCL send/recv transposes, 1 = peer-memory scatter
            sc.ctx.call("insb200_set_param", b"p2p", float(os.environ["MGPU_P2P"]))
        if "MGPU_FASTBAR" in os.environ:  # A/B: 0 = ncclAllReduce barriers, 1 = flag words in peer memory
            sc.ctx.call("insb200_set_param", b"fast_barrier", float(os.environ["MGPU_FASTBAR"]))
        if "MGPU_EXACT" in os.environ:  # 1 = the bit-exact kernels instead of k_stage3d_fast
            sc.ctx.call("insb200_set_param", b"exact", float(os.environ["MGPU_EXACT"]))
        if "MGPU_ZLINE" in os.environ:  # last axis of the pressure solve: 1 / -1 = without transposes (zline.cuh, default), 0 = distributed FFT
            sc.ctx.call("insb200_set_param", b"zline", float(os.environ["MGPU_ZLINE"]))
        if "MGPU_ZLCHUNKS" in os.environ:
            sc.ctx.call("insb200_set_param", b"zline_chunks", float(os.environ["MGPU_ZLCHUNKS"]))
        if "MGPU_TMODE" in os.environ:  # transposes: 2 = pack + contiguous peer copies (default), 1 = pitched peer copies, 0 = scatter kernels
            sc.ctx.call("insb200_set_param", b"transpose_mode", float(os.environ["MGPU_TMODE"]))
        sc.ctx.call("insb200_set_param", b"fft_pipe_min_ranks", 2.0)  # exercise the pipelined transposes at any world size
        if theta is not None:
            so.closure_model = o.smagorinsky_closure(so)
            sc.closure_model = ins.smagorinsky_closure(sc)
        u0 = o.random_field(so, kp=3, rng=np.random.default_rng(21))
        ul = ins.to_device(ins.local_slab(u0, D, rank, world), sc)
        m = ins.RKMethods.RK44(T=T)
        t, ke = ins.api._rk_steps(m, sc, ul, 0.0, dt, nsteps, theta, ke_hist=True)
        div = ins.to_host(ins.divergence(ul, sc))
        # gather the slabs on every rank
        parts = [torch.empty_like(ul) for _ in range(world)]
        dist.all_gather(parts, ul.contiguous())
        ug = ins.gather_slabs([ins.to_host(p) for p in parts], D)
        if rank == 0:
            mo = o.RK44(T)
            st = o.create_stepper(mo, so, o.psolver_spectral(so), u0.copy(order="F"))
            ke_o = [o.total_kinetic_energy(u0, so)]
            for _ in range(nsteps):
                st = o.timestep(mo, st, dt, theta=theta)
                ke_o.append(o.total_kinetic_energy(st.u, so))
            tol = 1e-12 if T == np.float64 else 1e-5
            err = o.rel_l2(ug, st.u)
            kerr = float(np.max(np.abs(ke / np.array(ke_o, dtype=np.float64) - 1)))
            ok = err <= tol and kerr <= tol
            print(f"[mgpu x{world}] n={n} {np.dtype(T).name} force={force} theta={theta}: rel_l2(u)={err:.3e} "
                  f"ke_err={kerr:.3e} max|div|={np.abs(div[sc.grid.Ip]).max():.2e} -> {'ok' if ok else 'FAIL'}", flush=True)
            failures += not ok
        torch.cuda.synchronize()
        dist.barrier()
        sc.ctx.close()  # same program point on every rank
    if os.environ.get("MGPU_BIG", "1") != "0" and 512 % world == 0:
        faulthandler.dump_traceback_later(240, exit=True)
        failures += big_case(backend)
    faulthandler.cancel_dump_traceback_later()
    f = torch.tensor([failures], device=f"cuda:{backend.device}")
    dist.broadcast(f, 0)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(int(f.item()) != 0)


if __name__ == "__main__":
    main()
