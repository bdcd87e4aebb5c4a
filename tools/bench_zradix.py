#!/usr/bin/env python
"""k_zsolve with and without radix-16 passes (set_param zradix16) on an nx x ny x nz Float32 grid."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import insb200 as ins  # noqa: E402

n = tuple(int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (256, 256, 1024)
T = np.float32
x = tuple((np.arange(m + 1) / float(max(n))).astype(T) for m in n)
setup = ins.Setup(x, Re=T(1000))
p = torch.randn(tuple(reversed(setup.grid.N)), dtype=torch.float32, device=setup.device)
ps = ins.psolver_spectral(setup)
for r16, thr in [(0, 256), (1, 256), (1, 384), (0, 384)]:
    setup.ctx.call("insb200_set_param", b"zradix16", float(r16))
    setup.ctx.call("insb200_set_param", b"zthreads", float(thr))
    for _ in range(3):
        ps(p.clone())
    setup.ctx.call("insb200_reset_timers")
    setup.ctx.call("insb200_set_profiling", 1)
    for _ in range(10):
        ps(p.clone())
    tm = bench.read_timers(setup)
    setup.ctx.call("insb200_set_profiling", 0)
    print(f"n={n} zradix16={r16} threads={thr}: " + "  ".join(f"{k}={v[0] / v[1]:.4f}ms" for k, v in tm.items() if k.startswith("poisson")), flush=True)
