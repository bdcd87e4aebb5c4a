#!/usr/bin/env python
"""Parameter sweep of the slab path under torchrun: ms per RK44 step (512^3 cells per GPU, no closure unless --smag) for a
list of insb200_set_param settings, one process group for all of them.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/sweep_mgpu.py \\
        "transpose_ce=1" "transpose_ce=0,scatter_blocks=64" ... [--n=256] [--steps=5] [--smag]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("NCCL_DEBUG", "WARN")
import bench  # noqa: E402
import insb200 as ins  # noqa: E402


def main():
    specs = [a for a in sys.argv[1:] if not a.startswith("--")]
    opt = {a.split("=")[0]: (a.split("=") + ["1"])[1] for a in sys.argv[1:] if a.startswith("--")}
    n_base, steps, smag = int(opt.get("--n", 512)), int(opt.get("--steps", 5)), "--smag" in opt
    backend = ins.init_distributed()
    rank, world = backend.rank, backend.nranks
    import torch.distributed as dist
    torch.cuda.set_device(backend.device)
    torch.cuda.set_stream(torch.cuda.Stream(device=backend.device))
    wl = dict(bench.WORKLOADS["dt3d_smag" if smag else "tgv3d"])
    n = bench.weak_grid(3, n_base, world)
    method = ins.RKMethods.RK44(T=np.float32)
    theta = wl["theta"]
    u0 = dt = None
    for spec in specs or [""]:
        setup = bench.make_setup(ins, backend, wl, n, [kv for kv in spec.split(",") if kv])
        if u0 is None:
            u0 = bench.synthetic_field(ins, setup)
            dt = min(wl["dt"] * wl["n"] / n_base, 0.5 * float(ins.get_cfl_timestep_(None, u0, setup)))
        u = u0.clone()
        ins.api._rk_steps(method, setup, u, 0.0, dt, 2, theta)
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ins.api._rk_steps(method, setup, u, 0.0, dt, steps, theta)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=u.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        setup.ctx.call("insb200_reset_timers")
        setup.ctx.call("insb200_set_profiling", 1)
        ins.api._rk_steps(method, setup, u, 0.0, dt, 2, theta)
        torch.cuda.synchronize()
        tm = bench.read_timers(setup)
        setup.ctx.call("insb200_set_profiling", 0)
        ok = bool(torch.isfinite(u).all())
        if rank == 0:
            sel = ("cufft_forward", "cufft_inverse", "poisson_scale", "fft_transpose", "fft_alltoall", "halo_exchange", "fill_ghosts")
            print(f"[sweep x{world}] {spec or 'default':48s} {t.item():8.3f} ms/step  finite={ok}  "
                  + " ".join(f"{k}={tm[k][0] / 2:.2f}" for k in sel if k in tm), flush=True)
        dist.barrier()
        setup.ctx.close()
        del u
        torch.cuda.empty_cache()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
