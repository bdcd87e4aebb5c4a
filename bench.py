#!/usr/bin/env python
"""bench.py -- RK44 time-stepping throughput of the B200 core (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload ...]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one RK44 `timestep!` (4 stages, each momentum + projection) of the whole grid.
Workloads (BASELINE.json configs):
  dt3d_smag  DecayingTurbulence3D periodic 512^3 Float32 + Smagorinsky closure (config 4; N=1
             default).  Weak scaling: the grid doubles z, y, x in turn with N, so that every GPU
             keeps 512^3 cells: N=2 512x512x1024, N=4 512x1024x1024, N=8 1024^3.
  tgv3d      TaylorGreenVortex3D periodic, Float32, no closure (config 5 is its 8-GPU 1024^3 case).
  kolmo2d    Kolmogorov2D forced periodic 4096^2 Float32 (config 3).
  tgv3d_f64  TaylorGreenVortex3D periodic 128^3 Float64 (config 2).
  dt2d_f64   DecayingTurbulence2D periodic 256^2 Float64 (config 1; the parity run is adaptive, the timed one uses dt = 1e-3).
One JSON line on rank 0:
  value     Mcell-updates/s, state resident in HBM, K steps inside ONE insb200_rk_steps call
  e2e       the same metric through insb200_solve_unsteady_host: HOST buffer in, H2D, K steps with a
            kinetic-energy read-back after every step, D2H of the result -- all inside the timed region
  roofline  dominant own kernel: algorithmic bytes per launch / CUDA-event duration per launch
  cpu_baseline  the CPU restatement of the reference algorithm timed on this box's host cores
`--impl reference` times only that CPU restatement (the reference is Julia; no Julia runtime exists
in this image) and prints the line with "impl": "reference".
"""
import argparse
import json
import os

# The caller's NCCL_DEBUG / NCCL_DEBUG_FILE are honoured (the driver reads the communicator sizes from NCCL's INFO
# log); only when nothing is set NCCL is kept quiet so that stdout carries the ONE JSON line and nothing else.
os.environ.setdefault("NCCL_DEBUG", "WARN")
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # NCCL logs to stdout by default: keep them, but on stderr
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (D, base n per axis, dtype, Re, dt, smagorinsky theta or None, forced)
    "dt3d_smag": dict(D=3, n=512, T="float32", Re=4e3, dt=2.5e-4, theta=0.17, force=False,
                      title="DecayingTurbulence3D periodic 512^3/GPU Float32 + Smagorinsky, RK44, psolver_spectral"),
    "tgv3d": dict(D=3, n=512, T="float32", Re=1e3, dt=2.5e-4, theta=None, force=False,
                  title="TaylorGreenVortex3D periodic 512^3/GPU Float32, RK44, psolver_spectral"),
    "tgv3d_f64": dict(D=3, n=128, T="float64", Re=1e3, dt=1e-3, theta=None, force=False,
                      title="TaylorGreenVortex3D periodic 128^3 Float64, RK44, psolver_spectral"),
    "dt2d_f64": dict(D=2, n=256, T="float64", Re=4e3, dt=1e-3, theta=None, force=False,
                     title="DecayingTurbulence2D periodic 256^2 Float64, RK44, psolver_spectral (fixed dt for timing)"),
    "kolmo2d": dict(D=2, n=4096, T="float32", Re=2e3, dt=2.5e-5, theta=None, force=True,
                    title="Kolmogorov2D forced periodic 4096^2/GPU Float32, RK44, psolver_spectral"),
}


def weak_grid(D, n, nranks):
    """Global cells per axis: the base n^D per GPU, doubling the last, then the previous, ... axis."""
    dims = [n] * D
    a, r = D - 1, nranks
    while r > 1:
        if r % 2:
            raise SystemExit("--gpus must be a power of two")
        dims[a] *= 2
        r //= 2
        a = (a - 1) % D
    return tuple(dims)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.lines:
            if t0 is not None and not (t0 <= ts <= t1 + 0.15):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# synthetic random-spectrum initial field, evaluated analytically on each rank's slab (device)
# ------------------------------------------------------------------------------------------------
def synthetic_modes(D, kp, nmodes, seed):
    """nmodes random Fourier modes with an Orlandi-type amplitude a(k) ~ k^2 exp(-pi (k/kp)^2)
    (the shape of create_spectrum, src/initializers.jl:82-181), directions projected ⟂ k."""
    rng = np.random.default_rng(seed)
    ks, es, ph = [], [], []
    while len(ks) < nmodes:
        k = rng.integers(-2 * kp, 2 * kp + 1, size=D)
        if not k.any():
            continue
        e = rng.standard_normal(D)
        e = e - k * (e @ k) / (k @ k)
        if np.linalg.norm(e) < 1e-3:
            continue
        kn = np.linalg.norm(k)
        a = kn ** 2 * np.exp(-np.pi * (kn / kp) ** 2)
        ks.append(k); es.append(a * e / np.linalg.norm(e)); ph.append(rng.random())
    es = np.array(es)
    es *= 1.0 / np.sqrt(0.5 * (es ** 2).sum())  # rms |u| = 1
    return np.array(ks), es, np.array(ph)


def synthetic_field(ins, setup, amp=1.0, kp=10, nmodes=24, seed=0):
    import torch

    g = setup.grid
    D = g.dimension
    ks, es, ph = synthetic_modes(D, kp, nmodes, seed)
    u = ins.vectorfield(setup)
    two_pi = 2.0 * np.pi
    for a in range(D):
        acc = u[a]
        for m in range(len(ks)):
            # phase = 2 pi (k . x + phi), separable 1-D pieces in float64, summed by broadcasting
            parts = []
            for b in range(D):
                v = torch.as_tensor(two_pi * (ks[m, b] * np.asarray(g.xu[a][b], dtype=np.float64) + (ph[m] if b == 0 else 0.0)),
                                    device=setup.device)
                c, s = torch.cos(v), torch.sin(v)
                sh = [1] * D
                sh[D - 1 - b] = -1  # tensor dims are reversed Julia dims
                parts.append((c.reshape(sh), s.reshape(sh)))
            # cos(p0+p1[+p2]) by angle addition on 1-D tables: one full-size temporary per mode
            c01 = parts[0][0] * parts[1][0] - parts[0][1] * parts[1][1]
            if D == 3:
                s01 = parts[0][1] * parts[1][0] + parts[0][0] * parts[1][1]
                val = c01 * parts[2][0] - s01 * parts[2][1]
            else:
                val = c01
            acc.add_(val.to(acc.dtype), alpha=float(amp * es[m, a]))
    ins.apply_bc_u_(u, 0.0, setup)
    ins.project_(u, setup)
    return ins.apply_bc_u_(u, 0.0, setup)


# ------------------------------------------------------------------------------------------------
# CPU baseline: the reference algorithm restated for the host cores (oracle/), bounded sample
# ------------------------------------------------------------------------------------------------
def host_cores():
    """Host cores this process may use (affinity mask), NOT the inherited OMP_NUM_THREADS: torch.distributed.run
    exports OMP_NUM_THREADS=1 to its workers, which made the N > 1 reference arm of round 1 single-threaded."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def cpu_default_n(wl):
    """Grid of the CPU arm: the workload's own per-GPU grid (512^3, 4096^2, 128^3) when the host has the memory for
    the reference's launch structure (about 40 words per cell incl. the 9-word tensor of the closure), else half of it."""
    D, n = wl["D"], wl["n"]
    need = (n ** D) * 48 * (4 if wl["T"] == "float32" else 8)
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 64 << 30
    return n if need < 0.6 * avail else n // 2


class CpuArm:
    """The reference's CPU algorithm (oracle/cpu_ref.c: C + OpenMP, the reference's launch structure; FFTs through
    pocketfft) on all host cores.  One instance keeps its field, so that consecutive step() calls are consecutive
    RK44 steps of one run; every step() is timed on its own."""

    def __init__(self, wl, n=None):
        from oracle import cpu_ref
        from oracle import ins_oracle as o

        self.o, self.cpu_ref, self.wl = o, cpu_ref, wl
        D = wl["D"]
        T = np.float32 if wl["T"] == "float32" else np.float64
        self.n = int(n or cpu_default_n(wl))
        self.cores = cpu_ref.set_num_threads(host_cores())
        x = tuple(np.linspace(0.0, 1.0, self.n + 1).astype(T) for _ in range(D))
        bf = None
        if wl["force"]:
            bf = (lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x)
        self.so = o.Setup(x, Re=T(wl["Re"]), bodyforce=bf)
        rng = np.random.default_rng(0)
        u = np.empty(self.so.grid.N + (D,), dtype=T, order="F")
        for a in range(D):  # component by component: no float64 temporary of the whole field
            u[..., a] = (0.1 * rng.standard_normal(self.so.grid.N, dtype=np.float32)).astype(T)
        self.u = o.apply_bc_u(u, 0.0, self.so)
        self.dt = wl["dt"] * wl["n"] / self.n
        self.cells = float(self.n) ** D
        self.desc = (f"RK44 step at {'x'.join([str(self.n)] * D)} {wl['T']}" + (" + Smagorinsky" if wl["theta"] is not None else "")
                     + (" + Kolmogorov force" if wl["force"] else "")
                     + f", oracle/cpu_ref.c (C + OpenMP, {self.cores} threads; FFT: pocketfft, {self.cores} workers) on {cpu_model()}")

    def step(self):
        t0 = time.perf_counter()
        self.cpu_ref.rk44_steps(self.u, self.so, self.dt, 1, self.wl["theta"], workers=self.cores)
        return time.perf_counter() - t0


# words (sizeof(T)) per cell per LAUNCH of each library kernel: the algorithmic minimum, every
# logical field read / written once (DESIGN.md "Kernels and rooflines")
def kernel_words(D, smag, force, nstages=4, tma=False, zsolve=False, fused2d=False, closure_fused=False):
    """Algorithmic words per cell per LAUNCH of each library kernel (every logical field read / written once)."""
    ntens = D * (D + 1) // 2
    extra = (D if force else 0)
    w = {
        # 2-D / generic path: stages 1..s-1: R u, R ustart, W k, W u' (4D); last stage: R u, R ustart, R k_1..k_{s-1}, W u'
        "momentum_stage": ((nstages - 1) * 4.0 * D + (3.0 + nstages - 1) * D) / nstages + extra + (ntens if smag else 0),
        "smagtensor": D + ntens,
        "divergence": D + 1.0,
        "pressure_gradient": 2.0 * D + 1,
        "poisson_scale": 2.0,          # half spectrum of complex words, read + written
        "poisson_forward": 2.0,        # decomposed grid (zline.cuh): the causal pass, read + write of the half spectrum; "poisson_scale" is the anticausal one
        "cufft_forward": 2.0 * D,      # one read + one write per transformed axis
        "cufft_inverse": 2.0 * D,
        "import_ghosted": 2.0 * D,
        "export_ghosted": 2.0 * D,
    }
    if tma:
        # 3-D TMA path: the stage kernel also writes the Poisson right-hand side (W div) and applies the previous
        # projection on load (R p).  first stage: R u, W k, W u', W div; middle stages: + R ustart, R p
        first, mid = 3.0 * D + 1, 4.0 * D + 2
        sadd = (D if smag and not closure_fused else 0)  # closure term s read by the stage kernel (separate closure pass only)
        w["momentum_stage"] = (first + (nstages - 2) * mid) / (nstages - 1) + extra + sadd
        # last stage: R u, R p, R ustart, R k_1..k_{s-1}, W u', W div
        w["momentum_stage_last"] = 3.0 * D + 2 + D * (nstages - 1) + extra + sadd
        # R u, R p (all but the first stage), W s; sigma never leaves shared memory
        w["smagorinsky_fused"] = 2.0 * D + (nstages - 1.0) / nstages
    if fused2d:
        # 2-D fused warp-marching stage kernel (stage2d.cuh): as the TMA kernels, one timer for all stages
        first, mid, last = 3.0 * D + 1, 4.0 * D + 2, 3.0 * D + 2 + D * (nstages - 1)
        w["momentum_stage"] = (first + (nstages - 2) * mid + last) / nstages + extra
    if zsolve:
        # k_zsolve does the last-axis transforms and the symbol in one read + write of the half spectrum
        # ("poisson_scale" timer); cuFFT only transforms the D-1 in-plane axes
        w["cufft_forward"] = w["cufft_inverse"] = 2.0 * (D - 1)
    return w


# DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of the committed `ncu --set full` capture of each
# kernel: profiles/ncu_traffic.json, written by tools/ncu_traffic.py from a capture of THIS command.  The value is
# static (ncu cannot run inside a timed bench); the line says which capture and which library revision it belongs to.
def ncu_traffic(workload, n, kernel):
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None, None
    with open(p) as f:
        tab = json.load(f)
    e = tab.get(f"{workload}/{n}/{kernel}")
    if not e:
        return None, None
    return float(e["dram_bytes_per_launch"]), f"static: {e['source']}@{e.get('git', '?')}"


def run_reference(args, wl):
    """`--impl reference`: the reference is Julia and cannot run in this image; its CPU algorithm restated under
    oracle/ (C + OpenMP, the reference's launch structure) is timed on ALL host cores.  One "step" = one RK44 step of
    ONE GPU's share of the workload (512^3 cells for the default workload): at N = 1 that is the configured grid, at
    N > 1 a bounded sample of it (per-cell rate; the 1024^3 grid of N = 8 would take an hour of CPU time)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm(wl, n=args.cpu_n)
    K, W = args.steps, args.warmup
    t_first = arm.step()  # first warm-up step (page faults, FFT plans)
    # keep the whole run within a few minutes: drop warm-up steps beyond the first when they do not fit
    w_done = 1
    budget = float(os.environ.get("BENCH_CPU_BUDGET_S", "420"))
    while w_done < W and (time.perf_counter() - T0) + (K + 1) * t_first < budget:
        arm.step()
        w_done += 1
    times = [arm.step() for _ in range(K)]
    el = float(np.mean(times))
    v = arm.cells / el / 1e6
    n = weak_grid(wl["D"], wl["n"], args.gpus)
    same = args.gpus == 1 and arm.n == wl["n"]
    line = {
        "impl": "reference", "metric": "RK44 Mcell-updates/s", "value": v, "unit": "Mcell-updates/s",
        "n_gpus": args.gpus, "steps": K, "warmup": w_done, "warmup_requested": W,
        "ms_per_step": el * 1e3, "steps_per_s": 1.0 / el, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if wl["T"] == "float32" else "f64", "data": "synthetic",
        "config": {"workload": wl["title"], "grid": list(n), "stepper": "RK44",
                   "cpu_grid": [arm.n] * wl["D"], "cpu_grid_is_configured_grid": same},
        "cpu_baseline": {"value": v, "unit": "Mcell-updates/s", "cores": arm.cores, "kind": "port", "host_cores": host_cores(),
                         "cpu_model": cpu_model(), "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS"),
                         "sample": f"{K} x " + arm.desc + ("" if same else " -- one GPU's share of the grid, per-cell rate")},
        "e2e": {"value": v, "unit": "Mcell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference is Julia (no runtime in this image): CPU restatement of its algorithm and launch structure",
    }
    print(json.dumps(line), flush=True)


T0 = time.perf_counter()


def read_timers(setup):
    import ctypes as C

    cap = 64
    names = (C.c_char_p * cap)()
    ms = (C.c_double * cap)()
    cnt = (C.c_longlong * cap)()
    n = C.c_int()
    setup.ctx.call("insb200_timers", names, ms, cnt, cap, C.byref(n))
    return {names[i].decode(): (ms[i], cnt[i]) for i in range(n.value)}


def make_setup(ins, backend, wl, n, params=()):
    T = np.float32 if wl["T"] == "float32" else np.float64
    x = tuple(np.linspace(0.0, 1.0, m + 1).astype(T) for m in n)
    bf = None
    if wl["force"]:
        bf = (lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x) if wl["D"] == 2 else \
             (lambda dim, x, y, z, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x + 0 * z)
    setup = ins.Setup(x, Re=T(wl["Re"]), bodyforce=bf, backend=backend)
    for kv in params:
        k, v = kv.split("=")
        setup.ctx.call("insb200_set_param", k.encode(), float(v))
    if wl["theta"] is not None:
        setup.closure_model = ins.smagorinsky_closure(setup)
    return setup


def parity_check(ins, backend, wl, dist, params=()):
    """Before anything is timed: 2 RK44 steps of a small grid on the SAME slab path (every rank one slab of the last
    axis: halos, distributed FFT, peer transposes) against oracle/ins_oracle.py on rank 0, from identical initial
    arrays.  3-D: 128 x 128 x 32 N, 2-D: 256 x 64 N.  Returns the record for the JSON line; `ok` is all-reduced."""
    import torch
    from oracle import ins_oracle as o

    rank, world = backend.rank, backend.nranks
    D = wl["D"]
    T = np.float32 if wl["T"] == "float32" else np.float64
    n = (128, 128, 32 * world) if D == 3 else (256, 64 * world)
    if os.environ.get("BENCH_PARITY_SMALL"):  # builder's own multi-GPU runs: all GPUs idle while rank 0 steps the oracle
        n = (64, 64, 16 * world) if D == 3 else (128, 32 * world)
    dt, nsteps = (1e-3 if D == 3 else 5e-4), 2
    pw = dict(wl)
    sc = make_setup(ins, backend, pw, n, params)
    x = tuple(np.linspace(0.0, 1.0, m + 1).astype(T) for m in n)
    bf = None
    if wl["force"]:
        bf = (lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x) if D == 2 else \
             (lambda dim, x, y, z, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x + 0 * z)
    so = o.Setup(x, Re=T(wl["Re"]), bodyforce=bf)
    theta = wl["theta"]
    if theta is not None:
        so.closure_model = o.smagorinsky_closure(so)
    u0 = o.random_field(so, kp=4, A=(1e-2 if wl["force"] else 1), rng=np.random.default_rng(7))  # every rank: same stream
    ul = ins.to_device(ins.local_slab(u0, D, rank, world), sc)
    m = ins.RKMethods.RK44(T=T)
    _, ke = ins.api._rk_steps(m, sc, ul, 0.0, dt, nsteps, theta, ke_hist=True)
    div = ins.divergence(ul, sc)
    dmax = float(div[tuple(slice(1, -1) for _ in range(D))].abs().max())
    if world > 1:
        parts = [torch.empty_like(ul) for _ in range(world)]
        dist.all_gather(parts, ul.contiguous())
        dm = torch.tensor([dmax], dtype=torch.float64, device=ul.device)
        dist.all_reduce(dm, op=dist.ReduceOp.MAX)
        dmax = float(dm.item())
    else:
        parts = [ul]
    tol = 1e-12 if T == np.float64 else 1e-5
    rec = None
    ok = 1
    if rank == 0:
        ug = ins.gather_slabs([ins.to_host(q) for q in parts], D)
        mo = o.RK44(T)
        st = o.create_stepper(mo, so, o.psolver_spectral(so), u0.copy(order="F"))
        ke_o = [o.total_kinetic_energy(u0, so)]
        for _ in range(nsteps):
            st = o.timestep(mo, st, T(dt), theta=theta)
            ke_o.append(o.total_kinetic_energy(st.u, so))
        err = float(o.rel_l2(ug, st.u))
        kerr = float(np.max(np.abs(np.asarray(ke) / np.asarray(ke_o, dtype=np.float64) - 1)))
        ok = int(err <= tol and kerr <= tol and np.isfinite(err))
        rec = {"grid": list(n), "ranks": world, "steps": nsteps, "dtype": wl["T"], "closure": theta is not None,
               "force": bool(wl["force"]), "rel_l2_u": err, "ke_err": kerr, "max_abs_div": dmax, "tol": tol, "ok": bool(ok),
               "against": "oracle/ins_oracle.py (NumPy restatement of the reference), identical initial arrays"}
    if world > 1:
        f = torch.tensor([ok], device=ul.device)
        dist.broadcast(f, 0)
        ok = int(f.item())
        dist.barrier()
    torch.cuda.synchronize()
    sc.ctx.close()
    return rec, bool(ok)


class BlowUp(Exception):
    """the timed solution became non-finite (decided collectively: every rank raises it at the same point)"""


def measure_with_fallback(run, args, world, rank, remeasured, name):
    """A non-finite run is invalid and is re-measured ONCE, on a decomposed grid with the other pressure solver (the
    distributed FFT instead of the transpose-free line solve; `run` reads args.param): the JSON line then says so in
    "remeasured".  One rank, an explicit solver choice, or a second failure end the run."""
    try:
        return run()
    except BlowUp as e:
        flat = [kv.replace(" ", "") for kv in args.param]
        if world == 1 or "zline=0" in flat or "p2p=0" in flat:
            raise SystemExit(f"bench.py: the solution blew up ({e}) -- invalid run")
        if rank == 0:
            print(f"bench.py: {e}; measuring again with --param zline=0", file=sys.stderr, flush=True)
        remeasured[name] = {"why": str(e), "with": "zline=0 (distributed FFT)"}
        args.param = list(args.param) + ["zline=0"]
        try:
            return run()
        except BlowUp as e2:
            raise SystemExit(f"bench.py: the solution blew up again ({e2}) -- invalid run")


def stable_dt(wl, n, base_n, dt_cfl):
    """Fixed step of a throughput run: the configured one, capped by half the reference's CFL bound (get_cfl_timestep!,
    solver.jl:101-125) and by the RK44 stability limit of the diffusion operator.  The reference's bound is per axis
    (Re dx^2 / 2); in 3-D the stiffest mode sees the SUM of the axes' eigenvalues, -nu * sum_d 4 / dx_d^2, and RK44 is stable
    on the negative real axis only up to 2.785: at 1024^3, Re = 1000 half the CFL bound gives z = -3.0, amplification 1.375
    per step -- the (pi, pi, pi) mode grows from round-off to overflow within about a hundred steps (seen once in the
    8-GPU run of round 2; throughput does not depend on dt, the solution must stay finite)."""
    nu_sum = sum(4.0 * float(m) ** 2 for m in n) / wl["Re"]   # unit domain: dx_d = 1 / n_d
    return min(wl["dt"] * wl["n"] / base_n, dt_cfl, 0.8 * 2.785 / nu_sum)


def nvlink_model(D, n, world, esz):
    """Algorithmic NVLink bytes sent per GPU per RK44 step: distributed-FFT transposes and halo planes."""
    if world == 1:
        return 0.0, 0.0
    nxh = n[0] // 2 + 1
    rows = float(np.prod([float(m) for m in n[1:]])) / world          # (ky, local z) rows of the in-slab spectrum
    transpose = 4 * 2 * (world - 1) / world * nxh * rows * 2 * esz     # 4 solves x (there + back), complex words
    plane = float(np.prod([float(m) for m in n[:-1]]))
    halo = 4 * (3 * D + 1) * plane * esz                               # per stage: 2 planes x D (u), 1 x D (u'), 1 (p)
    return transpose, halo


def measure(ins, backend, wl, args, base_n, dist, workload_name, legs=("timers", "per_step", "e2e")):
    """Time K RK44 steps of one workload on the slab path; returns the fields of the JSON line."""
    import torch

    rank, world = backend.rank, backend.nranks
    D = wl["D"]
    T = np.float32 if wl["T"] == "float32" else np.float64
    n = weak_grid(D, base_n, world)
    setup = make_setup(ins, backend, wl, n, args.param)
    theta = wl["theta"]
    method = ins.RKMethods.RK44(T=T)
    amp = 1e-2 if wl["force"] else 1.0
    if world == 1 and not args.modes_init:
        # random_field of the reference (src/initializers.jl:189-219): spectrum, inverse FFT and projection on the device
        u = ins.random_field(setup, A=amp, kp=10, rng=np.random.default_rng(0), ondevice=True)
        init = "random_field(kp=10) (src/initializers.jl:189-219), spectrum + inverse FFT + projection on the device"
    else:
        u = synthetic_field(ins, setup, amp=amp)
        init = "24 random Fourier modes with Orlandi-type amplitudes evaluated per slab, projected on the device (the random_field spectrum needs the global grid on one device)"
    # fixed step for the throughput run: the configured one, capped at half the reference's own CFL bound
    # (get_cfl_timestep!, solver.jl:101-125) of the start field -- random_field(A = 1, kp = 10) reaches |u| = 16, for which
    # the examples step adaptively; throughput does not depend on dt
    dt_cfl = 0.5 * float(ins.get_cfl_timestep_(None, u, setup))
    dt = stable_dt(wl, n, base_n, dt_cfl)
    lib, h = setup.ctx.lib, setup.ctx.h
    K, W = args.steps, args.warmup
    cells = float(np.prod([float(m) for m in n]))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=setup.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- value: state resident in HBM, K steps in one insb200_rk_steps call --------------------
    t = 0.0
    ins.api._rk_steps(method, setup, u, t, dt, W, theta)
    sampler = ClockSampler(backend.device) if rank == 0 else None
    barrier()
    if sampler:
        sampler.start()
        time.sleep(0.3)
    l0, f0, g0 = lib.insb200_launch_count(h), lib.insb200_fft_exec_count(h), lib.insb200_graph_replays(h)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    w0 = time.perf_counter()
    e0.record()
    torch.cuda.nvtx.range_push("timed")  # ncu --nvtx --nvtx-include "timed/" selects this region
    ins.api._rk_steps(method, setup, u, t, dt, K, theta)
    torch.cuda.nvtx.range_pop()
    e1.record()
    barrier()
    w1 = time.perf_counter()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches, ffts = lib.insb200_launch_count(h) - l0, lib.insb200_fft_exec_count(h) - f0
    graph_replays = lib.insb200_graph_replays(h) - g0
    if ms < 1000.0:
        # the 100 ms clock sampler needs a region of >= 1 s: repeat the same K steps (untimed for `value`) until it has
        # samples.  EVERY rank repeats (the steps are collective), the count follows from the rank-maximum `ms`
        reps = int(np.ceil(1200.0 / max(ms, 1.0)))
        for _ in range(reps):
            ins.api._rk_steps(method, setup, u, t, dt, K, theta)
        torch.cuda.synchronize()
        w1 = time.perf_counter()
    if dist is not None:
        dist.barrier()
    clocks = sampler.stop(w0, w1) if sampler else None
    ms_step = ms / K
    value = cells / (ms_step * 1e-3) / 1e6
    if max_over_ranks(0.0 if bool(torch.isfinite(u).all()) else 1.0) > 0.5:  # the same verdict on every rank
        barrier()
        setup.ctx.close()  # collective teardown, as at the end of a good run
        del u
        torch.cuda.empty_cache()
        raise BlowUp(f"{workload_name}: non-finite u after {W + K} steps and the clock-sampling repeats, dt = {dt:g}")
    out = {"n": n, "cells": cells, "ms_step": ms_step, "value": value, "launches": int(launches), "ffts": int(ffts),
           "graph_replays": int(graph_replays), "clocks": clocks, "dt": dt, "init": init, "theta": theta}
    esz = np.dtype(T).itemsize
    cells_local = cells / world
    peak, peak_src = peaks()
    model_bytes = lib.insb200_model_bytes_per_cell_step(h, 4)
    step_gbs = model_bytes * cells_local / (ms_step * 1e-3) / 1e9
    out["step_roofline"] = {"model_bytes_per_cell_step": model_bytes, "achieved": step_gbs, "peak": peak, "unit": "GB/s",
                            "frac": step_gbs / peak, "aggregate_peak": peak * world,
                            "note": "SURVEY 8d model, 46D+8 words per cell and step (+D per stage with a stored force); the "
                                    "closure adds no words when it is fused into the stage kernel (closure_fused below)"}

    # ---- per-kernel timers over an identical region (CUDA events around every launch) ----------
    if "timers" in legs:
        setup.ctx.call("insb200_reset_timers")
        setup.ctx.call("insb200_set_profiling", 1)
        barrier()
        ins.api._rk_steps(method, setup, u, t, dt, K, theta)
        barrier()
        timers = read_timers(setup)
        setup.ctx.call("insb200_set_profiling", 0)
        if os.environ.get("BENCH_RANK_TIMERS"):  # per-rank view of all timers (stderr), for load-balance checks
            print(f"[rank {rank}] " + "  ".join(f"{k}={v[0] / K:.3f}ms/step({v[1] // K}x)" for k, v in timers.items()),
                  file=sys.stderr, flush=True)
        nz_glob = n[-1]
        closure_fused = theta is not None and "smagorinsky_fused" not in timers and "smagtensor" not in timers
        words = kernel_words(D, theta is not None, wl["force"], tma=(D == 3 and base_n % 32 == 0),
                             zsolve=(D == 3 and 8 <= nz_glob <= 2048 and nz_glob & (nz_glob - 1) == 0),
                             fused2d=(D == 2 and theta is None and not any(kv.replace(" ", "") == "fuse2d=0" for kv in args.param)),
                             closure_fused=closure_fused)
        kern = {}
        for name, (tms, cnt) in timers.items():
            per = tms / max(cnt, 1)
            gbs = words[name] * esz * cells_local / (per * 1e-3) / 1e9 if name in words and per > 0 else None
            kern[name] = {"ms_per_step": tms / K, "launches_per_step": cnt / K, "ms_per_launch": per,
                          "alg_GBs": gbs, "frac": gbs / peak if gbs else None}
        own = {k: v for k, v in kern.items() if not k.startswith("cufft") and k in words}
        dom = max(own, key=lambda k: own[k]["ms_per_step"]) if own else None
        if dom:
            traffic, tsrc = ncu_traffic(workload_name, base_n, dom)
            out["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": kern[dom]["alg_GBs"], "peak": peak, "unit": "GB/s",
                               "frac": kern[dom]["frac"], "traffic": traffic, "traffic_source": tsrc,
                               "peak_source": peak_src,
                               "alg_bytes_per_launch": words[dom] * esz * cells_local,
                               "alg_words_per_cell_launch": words[dom],
                               "ms_per_launch": kern[dom]["ms_per_launch"],
                               "share_of_step": kern[dom]["ms_per_step"] / sum(v["ms_per_step"] for v in kern.values())}
        out["kernels"] = kern
        out["closure_fused"] = closure_fused
        if world > 1:
            tb, hb = nvlink_model(D, n, world, esz)
            ht = kern.get("halo_exchange", {}).get("ms_per_step", 0.0)
            flat = [kv.replace(" ", "") for kv in args.param]
            if "zline=0" not in flat and "p2p=0" not in flat:
                # default on a decomposed grid: the last axis is solved by two recursions through the slab (zline.cuh); a GPU
                # reads two complex words per in-plane mode and solve from its neighbours instead of transposing its spectrum
                modes = float(n[0] // 2 + 1) * (float(n[1]) if D == 3 else 1.0)
                out["nvlink"] = {"pressure_solve": "transpose-free last-axis solve (zline.cuh)",
                                 "exchange_bytes_per_gpu_step": 4 * 2 * modes * 2 * esz, "halo_bytes_per_gpu_step": hb,
                                 "distributed_fft_bytes_per_gpu_step_avoided": tb, "halo_ms_per_step": ht,
                                 "barrier_ms_per_step": kern.get("fft_alltoall", {}).get("ms_per_step"),
                                 "achieved": None, "unit": "GB/s per GPU, one direction", "peak": 900.0, "frac": None,
                                 "note": "the neighbours' sums are read by peer loads inside the solve kernels (timer poisson_scale); "
                                         "no separate transfer exists to time"}
            else:
                tt = sum(kern[k]["ms_per_step"] for k in ("fft_transpose", "fft_pack_unpack", "fft_alltoall") if k in kern)
                out["nvlink"] = {"pressure_solve": "distributed FFT with transposes", "transpose_bytes_per_gpu_step": tb, "halo_bytes_per_gpu_step": hb,
                                 "transpose_ms_per_step": tt, "halo_ms_per_step": ht,
                                 "achieved": tb / (tt * 1e-3) / 1e9 if tt > 0 else None, "unit": "GB/s per GPU, one direction",
                                 "peak": 900.0, "peak_source": "NVLink 5 nominal per direction; a plain peer copy measured 690 GB/s "
                                                               "(profiles/r1_p2p_probe_2xB200.txt)",
                                 "frac": (tb / (tt * 1e-3) / 1e9 / 900.0) if tt > 0 else None,
                                 "note": "transposes run on other streams under the in-slab cuFFT; their time is the transpose timers' sum "
                                         "(stream time including waits: a lower bound of the rate)"}

    # ---- per-step API: K calls of insb200_rk_step, the loop solve_unsteady runs in Julia (solver.jl:65,75) ----
    if "per_step" in legs:
        st = ins.create_stepper(method, setup, None, u, None, 0.0)
        for _ in range(2):
            st = ins.timestep_(method, st, dt, theta=theta)
        barrier()
        e0.record()
        for _ in range(K):
            st = ins.timestep_(method, st, dt, theta=theta)
        e1.record()
        barrier()
        ms_api = max_over_ranks(e0.elapsed_time(e1)) / K
        out["per_step_api"] = {"value": cells / (ms_api * 1e-3) / 1e6, "unit": "Mcell-updates/s", "ms_per_step": ms_api,
                               "what": "K calls of insb200_rk_step on the caller's reference-layout device array (timestep!, "
                                       "step_explicit_runge_kutta.jl:4-59): layout import + export every step"}

    # ---- e2e: host buffer -> insb200_solve_unsteady_host (H2D, K steps + KE read-back, D2H) ----
    if "e2e" in legs:
        uh = torch.empty(u.shape, dtype=u.dtype, pin_memory=True)
        uh.copy_(u)
        u_start = uh.clone()  # pageable host copy of the start field
        del u                 # the library stages the host field itself: do not hold a second device copy (1024^3 on one GPU)
        torch.cuda.empty_cache()
        ins.solve_unsteady_host(setup, (0.0, dt * 1), uh, method=method, dt=dt, theta=theta, ke_hist=2)  # warm
        uh.copy_(u_start)
        del u_start
        barrier()
        c0 = time.perf_counter()
        ns, tf, ke = ins.solve_unsteady_host(setup, (0.0, dt * K), uh, method=method, dt=dt, theta=theta, ke_hist=K + 1)
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - c0)
        assert ns == K, (ns, K)
        field_bytes = uh.numel() * uh.element_size() * world
        out["e2e"] = {"value": cells * K / e2e_s / 1e6, "unit": "Mcell-updates/s", "h2d_bytes_per_step": field_bytes / K,
                      "d2h_bytes_per_step": field_bytes / K + 8 * (K + 1) / K, "ms_per_step": e2e_s * 1e3 / K,
                      "what": "insb200_solve_unsteady_host: pinned host u -> H2D -> K RK44 steps, kinetic energy read back "
                              "after every step -> D2H of u; field copies amortised over K steps",
                      "ke_first_last": [float(ke[0]), float(ke[-1])]}
    out["workspace_bytes"] = int(lib.insb200_workspace_bytes(h))
    barrier()
    setup.ctx.close()  # collective teardown at the same program point on every rank
    torch.cuda.empty_cache()
    return out


def run_ours(args, wl):
    import torch

    import insb200 as ins

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 core has no CPU fallback")
    backend = ins.init_distributed()
    rank, world = backend.rank, backend.nranks
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    dist = None
    if world > 1:
        import torch.distributed as dist
    torch.cuda.set_device(backend.device)
    # a non-default stream: the library replays whole steps as CUDA graphs on small grids, and capture is not
    # allowed on the legacy default stream (Julia's CUDA.jl runs on a task-local stream as well)
    torch.cuda.set_stream(torch.cuda.Stream(device=backend.device))
    T = np.float32 if wl["T"] == "float32" else np.float64
    base_n = args.n or wl["n"]
    K, W = args.steps, args.warmup

    # ---- parity self-check on the slab path, before anything is timed ---------------------------
    parity = None
    if not args.no_parity:
        parity, ok = parity_check(ins, backend, wl, dist, args.param)
        if not ok:
            if rank == 0:
                print(json.dumps({"metric": "RK44 Mcell-updates/s", "value": None, "parity_check": parity,
                                  "error": "parity self-check against the oracle FAILED: nothing was timed"}), flush=True)
            raise SystemExit(3)

    remeasured = {}

    def measured(wl_, name, **kw):
        return measure_with_fallback(lambda: measure(ins, backend, wl_, args, base_n, dist, name, **kw), args, world, rank, remeasured, name)

    m = measured(wl, args.workload)

    # ---- config 5 (north_star's target): TaylorGreenVortex3D 1024^3 Float32 on 8 GPUs ------------
    config5 = None
    if (world == 8 and args.workload == "dt3d_smag" and not args.n) or args.also_tgv:
        wl5 = WORKLOADS["tgv3d"]
        p5, ok5 = (None, True) if args.no_parity else parity_check(ins, backend, wl5, dist, args.param)
        m5 = measured(wl5, "tgv3d", legs=("timers",))
        if rank == 0:
            config5 = {"workload": wl5["title"], "grid": list(m5["n"]), "n_gpus": world, "ms_per_step": m5["ms_step"],
                       "steps_per_s": 1e3 / m5["ms_step"], "value": m5["value"], "unit": "Mcell-updates/s",
                       "step_roofline": m5["step_roofline"], "nvlink": m5.get("nvlink"), "parity_check": p5,
                       "kernels": {k: {"ms_per_step": v["ms_per_step"], "frac": v["frac"]} for k, v in m5["kernels"].items()},
                       "target": "north_star: >= 0.60 of the aggregate HBM roofline (<= 20.2 ms/step at 1024^3 on 8 B200)",
                       "init": m5["init"]}
        if not ok5:
            if rank == 0:
                print(json.dumps({"metric": "RK44 Mcell-updates/s", "value": None, "parity_check": p5,
                                  "error": "config-5 parity self-check FAILED"}), flush=True)
            raise SystemExit(3)
    if dist is not None:
        dist.destroy_process_group()
    if rank != 0:
        return
    # ---- CPU baseline beside it (rank 0, N = 1 only): a bounded sample of the same workload ----------
    cpu = None
    if world == 1 and not args.no_cpu:
        arm = CpuArm(wl, n=args.cpu_n)
        arm.step()  # warm-up
        ts = [arm.step() for _ in range(max(1, args.cpu_steps))]
        el = float(np.mean(ts))
        cpu = {"value": arm.cells / el / 1e6, "unit": "Mcell-updates/s", "cores": arm.cores, "kind": "port",
               "sample": f"{len(ts)} x " + arm.desc, "host_cores": host_cores(), "cpu_model": cpu_model(), "ms_per_step": el * 1e3}
    n = m["n"]
    line = {
        "metric": "RK44 Mcell-updates/s", "value": m["value"], "unit": "Mcell-updates/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": m["ms_step"], "steps_per_s": 1e3 / m["ms_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32" if T == np.float32 else "f64", "data": "synthetic",
        "config": {"workload": wl["title"] if base_n == wl["n"] else wl["title"].replace(f"{wl['n']}^", f"{base_n}^"),
                   "grid": list(n), "cells_per_gpu": m["cells"] / world, "stepper": "RK44",
                   "dt": m["dt"], "Re": wl["Re"], "smagorinsky_theta": m["theta"], "parallelism": f"slab{world}(last axis)",
                   "l2": "inputs larger than L2 (state >> 126 MB), no flush", "init": m["init"]},
        "clocks": m["clocks"], "e2e": m.get("e2e"), "per_step_api": m.get("per_step_api"),
        "gpu_launches": m["launches"], "cufft_execs": m["ffts"],
        "graph_replays": m["graph_replays"], "workspace_bytes_per_gpu": m["workspace_bytes"],
        "roofline": m.get("roofline"), "step_roofline": m["step_roofline"], "closure_fused": m.get("closure_fused"),
        "nvlink": m.get("nvlink"), "parity_check": parity, "config5": config5, "remeasured": remeasured or None,
        "kernels": m.get("kernels"), "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="dt3d_smag", choices=sorted(WORKLOADS))
    ap.add_argument("--n", type=int, default=0, help="override the per-GPU base grid size (smoke runs)")
    ap.add_argument("--cpu-n", type=int, default=None, help="grid size of the CPU baseline sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--param", action="append", default=[], help="key=value forwarded to insb200_set_param (A/B runs)")
    ap.add_argument("--cpu-steps", type=int, default=2, help="timed steps of the cpu_baseline leg of the default arm")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity self-check (profiling runs only)")
    ap.add_argument("--also-tgv", action="store_true", help="add the config-5 sub-record (tgv3d on the same ranks) at any N")
    ap.add_argument("--modes-init", action="store_true", help="N = 1: start from the 24-mode synthetic field instead of random_field")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours" and not args.n:
        print("bench.py: warmup < 3 is not a valid measurement; using 3", file=sys.stderr)
        args.warmup = 3
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
