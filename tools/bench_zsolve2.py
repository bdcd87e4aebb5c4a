#!/usr/bin/env python
"""A/B of k_zsolve on an nx x ny x nz Float32 grid (default 256 x 256 x 1024: config 5's column length): exact build
(-fmad=false) against the contracted build of fast.cu, tile width 8 / 16 columns, 256 / 384 / 512 threads.
Prints ms per launch and the fraction of the HBM peak (half spectrum read + written once)."""
import json
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
peak, _ = bench.peaks()
byts = 2.0 * (n[0] // 2 + 1) * n[1] * n[2] * 8  # complex half spectrum, read + write
for exact in (1, 0):
    for cols, thr in [(0, 256), (8, 256), (8, 384), (16, 256), (16, 512)]:
        setup = ins.Setup(x, Re=T(1000))
        p = torch.randn(tuple(reversed(setup.grid.N)), dtype=torch.float32, device=setup.device)
        ps = ins.psolver_spectral(setup)
        for k, v in (("exact", exact), ("zcols", cols), ("zthreads", thr)):
            setup.ctx.call("insb200_set_param", k.encode(), float(v))
        for _ in range(3):
            ps(p.clone())
        setup.ctx.call("insb200_reset_timers")
        setup.ctx.call("insb200_set_profiling", 1)
        for _ in range(10):
            ps(p.clone())
        tm = bench.read_timers(setup)
        setup.ctx.call("insb200_set_profiling", 0)
        ms = tm["poisson_scale"][0] / tm["poisson_scale"][1]
        print(f"n={n} exact={exact} zcols={cols or 'auto'} zthreads={thr}: k_zsolve {ms:.3f} ms = {byts / ms / 1e6:.0f} GB/s = {byts / ms / 1e6 / peak:.3f} of peak"
              f"   (cuFFT 2-D fwd {tm['cufft_forward'][0] / tm['cufft_forward'][1]:.3f} ms)", flush=True)
        setup.ctx.close()
