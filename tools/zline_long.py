"""Long-run probe of the pressure solvers: N blocks of K RK44 steps of a bench workload, kinetic energy and max |u| after
every block, for the parameter sets given on the command line (e.g. "zline=0" "zline=1,zline_chunks=2").
    python tools/zline_long.py [--workload tgv3d] [--blocks 10] [--steps 10] [--n 512] "k=v,k=v" ...
Under torch.distributed.run it runs on the slab path (every rank prints its own max |u|)."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="tgv3d")
    ap.add_argument("--blocks", type=int, default=10)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--n", type=int, default=0)
    ap.add_argument("sets", nargs="*", default=["zline=0", "zline=1"])
    a = ap.parse_args()
    import torch

    import insb200 as ins

    backend = ins.init_distributed()
    rank, world = backend.rank, backend.nranks
    torch.cuda.set_device(backend.device)
    torch.cuda.set_stream(torch.cuda.Stream(device=backend.device))
    wl = bench.WORKLOADS[a.workload]
    base_n = a.n or wl["n"]
    T = np.float32 if wl["T"] == "float32" else np.float64
    n = bench.weak_grid(wl["D"], base_n, world)
    for s in a.sets:
        setup = bench.make_setup(ins, backend, wl, n, [kv for kv in s.split(",") if kv])
        u = bench.synthetic_field(ins, setup, amp=1e-2 if wl["force"] else 1.0)
        dt = bench.stable_dt(wl, n, base_n, 0.5 * float(ins.get_cfl_timestep_(None, u, setup)))
        m = ins.RKMethods.RK44(T=T)
        t = 0.0
        for b in range(a.blocks):
            _, ke = ins.api._rk_steps(m, setup, u, t, dt, a.steps, wl["theta"], ke_hist=True)
            t += a.steps * dt
            umax = float(u.abs().max())
            div = float(ins.divergence(u, setup).abs().max())
            print(f"[{s} r{rank}] block {b} (step {(b + 1) * a.steps}): ke {float(ke[-1]):.9e}  max|u| {umax:.6e}  max|div| {div:.3e}", flush=True)
            if not np.isfinite(umax):
                break
        setup.ctx.close()
        del u
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
