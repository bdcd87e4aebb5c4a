#!/usr/bin/env python
"""Large-slab check on one GPU: the TMA plane-pipeline kernels (variant 2) against the LDG-tiled kernels
(variant 1) on the same start field -- the variants are bit-identical by construction
(tests/test_gpu_stepping.py::test_kernel_variants_bit_identical at small sizes); this runs the comparison
where 32-bit offsets could wrap, e.g.   python tools/check_large.py 1024 1024 768 1   (nx ny nz, Smagorinsky on/off).
(The stored-sigma closure of variant 1 needs 6 more scalar fields: 1024^3 with the closure does not fit next to the
copies this script holds, 1024 x 1024 x 768 does and still crosses 2^31 elements per vector field.)

Only one extra copy of the field is held: the start field is regenerated (deterministic) for the second run.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch

    import bench
    import insb200 as ins

    n = tuple(int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (1024, 1024, 768)
    smag = bool(int(sys.argv[4])) if len(sys.argv) > 4 else True
    nsteps = int(sys.argv[5]) if len(sys.argv) > 5 else 2
    T = np.float32
    torch.cuda.set_stream(torch.cuda.Stream())
    x = tuple((np.arange(m + 1) / float(max(n))).astype(T) for m in n)  # one spacing 1/max(n) on every axis
    setup = ins.Setup(x, Re=T(4e3))
    theta = None
    if smag:
        setup.closure_model = ins.smagorinsky_closure(setup)
        theta = 0.17
    method = ins.RKMethods.RK44(T=T)
    dt = 2.5e-4 * 512 / max(n)

    def run(variant):
        setup.ctx.call("insb200_set_param", b"exact", 1.0)  # bit-exact kernels on both sides
        setup.ctx.call("insb200_set_variant", variant)
        u = bench.synthetic_field(ins, setup)
        setup.ctx.call("insb200_reset_timers")
        setup.ctx.call("insb200_set_profiling", 1)
        ins.api._rk_steps(method, setup, u, 0.0, dt, nsteps, theta)
        torch.cuda.synchronize()
        tm = bench.read_timers(setup)
        setup.ctx.call("insb200_set_profiling", 0)
        per = {k: round(v[0] / max(v[1], 1), 3) for k, v in tm.items() if k in ("momentum_stage", "momentum_stage_last", "smagorinsky_fused")}
        return u, per

    ua, pa = run(2)
    ke = ins.total_kinetic_energy(ua, setup)
    ub, pb = run(1)
    same = bool(torch.equal(ua, ub))
    finite = bool(torch.isfinite(ua).all())
    dmax = 0.0
    if not same:
        for a in range(3):
            dmax = max(dmax, float((ua[a] - ub[a]).abs().max()))
    print(f"n={n} smag={smag} steps={nsteps}: variant 2 vs 1 bit-identical={same} finite={finite} max|diff|={dmax:.3e} "
          f"KE={ke:.9e} umax={float(ua.abs().max()):.6f}")
    print("  ms per launch, variant 2:", pa)
    print("  ms per launch, variant 1:", pb)
    print("  workspace GB:", setup.ctx.lib.insb200_workspace_bytes(setup.ctx.h) / 1e9)
    sys.exit(0 if same and finite else 1)


if __name__ == "__main__":
    main()
