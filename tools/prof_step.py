#!/usr/bin/env python
"""Short driver for ncu captures: a few RK44 steps of one bench workload, nothing else.

    python tools/prof_step.py [--workload dt3d_smag] [--n 512] [--steps 2] [--variant 2]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="dt3d_smag")
    ap.add_argument("--n", type=int, default=0)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--variant", type=int, default=2)
    ap.add_argument("--zchunk", type=int, default=0)
    args = ap.parse_args()
    import torch

    import insb200 as ins

    wl = bench.WORKLOADS[args.workload]
    T = np.float32 if wl["T"] == "float32" else np.float64
    n = args.n or wl["n"]
    x = tuple(np.linspace(0.0, 1.0, n + 1).astype(T) for _ in range(wl["D"]))
    bf = (lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x) if wl["force"] else None
    setup = ins.Setup(x, Re=T(wl["Re"]), bodyforce=bf)
    if wl["theta"] is not None:
        setup.closure_model = ins.smagorinsky_closure(setup)
    setup.ctx.call("insb200_set_param", b"variant", float(args.variant))
    if args.zchunk:
        setup.ctx.call("insb200_set_param", b"zchunk", float(args.zchunk))
    u = bench.synthetic_field(ins, setup, amp=1e-2 if wl["force"] else 1.0, nmodes=4)
    m = ins.RKMethods.RK44(T=T)
    ins.api._rk_steps(m, setup, u, 0.0, wl["dt"] * wl["n"] / n, args.steps, wl["theta"])
    torch.cuda.synchronize()
    assert bool(torch.isfinite(u).all())
    print("ok", args.workload, n, args.steps)


if __name__ == "__main__":
    main()
