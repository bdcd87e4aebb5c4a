#!/usr/bin/env python
"""Tuning sweep of the fused last-axis solve kernel (k_zsolve): tile width x threads per block."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import insb200 as ins  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
T = np.float32
x = tuple(np.linspace(0.0, 1.0, n + 1).astype(T) for _ in range(3))
setup = ins.Setup(x, Re=T(1000))
p = torch.randn(tuple(reversed(setup.grid.N)), dtype=torch.float32, device=setup.device)
ps = ins.psolver_spectral(setup)
for zsolve, cols, thr in [(0, 0, 256), (1, 16, 256), (1, 16, 512), (1, 16, 384), (1, 8, 256), (1, 8, 128), (1, 8, 512), (1, 4, 128), (1, 4, 256)]:
    for k, v in (("zsolve", zsolve), ("zcols", cols), ("zthreads", thr)):
        setup.ctx.call("insb200_set_param", k.encode(), float(v))
    for _ in range(3):
        ps(p.clone())
    setup.ctx.call("insb200_reset_timers")
    setup.ctx.call("insb200_set_profiling", 1)
    for _ in range(10):
        ps(p.clone())
    tm = bench.read_timers(setup)
    setup.ctx.call("insb200_set_profiling", 0)
    print(f"zsolve={zsolve} cols={cols} threads={thr}: " + "  ".join(f"{k}={v[0] / v[1]:.3f}ms" for k, v in tm.items() if k.startswith(("cufft", "poisson"))), flush=True)
