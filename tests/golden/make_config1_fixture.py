#!/usr/bin/env python
"""Golden fixture of BASELINE config 1 at its stated size: DecayingTurbulence2D periodic 256^2 Float64, Re = 4e3,
random_field(kp = 10, A = 1), tlims = (0, 1), adaptive dt with cfl = 0.9, RK44, psolver_spectral
(examples/DecayingTurbulence2D.jl:22-29, src/solver.jl:25-28, 52-80).

Written by oracle/ins_oracle.py (the NumPy restatement of the reference; the reference itself is Julia and cannot run
here -- see DESIGN.md section 0).  Stores the start field, the end field, the kinetic-energy history and the step count,
and -- from a twin run whose start field is perturbed by one ulp per entry -- how much the flow itself amplifies
round-off over the 805 steps: no implementation can agree with another one better than that.

    python tests/golden/make_config1_fixture.py        (about 4 minutes on 8 cores)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ins_oracle as o  # noqa: E402


def main():
    T = np.float64
    x = tuple(np.linspace(0.0, 1.0, 257) for _ in range(2))
    so = o.Setup(x, Re=T(4e3))
    u0 = o.random_field(so, kp=10, A=1, rng=np.random.default_rng(5))
    ke = lambda s: float(o.total_kinetic_energy(s.u, so))  # noqa: E731
    st, out = o.solve_unsteady(so, (0.0, 1.0), u0, cfl=0.9, processors={"ke": ke})
    rng = np.random.default_rng(99)
    u0p = np.asfortranarray(u0 * (1.0 + np.finfo(T).eps * rng.choice([-1.0, 1.0], size=u0.shape)))
    stp, outp = o.solve_unsteady(so, (0.0, 1.0), o.apply_bc_u(u0p, 0.0, so), cfl=0.9, processors={"ke": ke})
    growth = o.rel_l2(stp.u, st.u) if stp.n == st.n else float("nan")
    ke_growth = float(np.max(np.abs(np.array(outp["ke"][: len(out["ke"])]) / np.array(out["ke"][: len(outp["ke"])]) - 1)))
    print(f"steps {st.n} (twin {stp.n}), t = {float(st.t)}, KE {out['ke'][0]:.6f} -> {out['ke'][-1]:.6f}, "
          f"round-off amplification: rel-L2(u) = {growth:.3e}, KE = {ke_growth:.3e}")
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "config1_256sq_f64.npz"), u0=u0, u_end=st.u, ke=np.array(out["ke"]),
                        nsteps=st.n, t_end=float(st.t), roundoff_growth_u=growth, roundoff_growth_ke=ke_growth)


if __name__ == "__main__":
    main()
