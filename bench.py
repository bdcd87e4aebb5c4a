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

# stdout carries exactly ONE JSON line: NCCL's banner ("NCCL version ...") would otherwise land on it
os.environ["NCCL_DEBUG"] = os.environ.get("BENCH_NCCL_DEBUG", "WARN")
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
def cpu_sample(wl, steps=1, n=None):
    """Time `steps` RK44 steps of the CPU restatement at a bounded size of the same workload
    (same D, dtype, closure, force).  Returns (Mcell-updates/s, steps/s at that size, description).
    Uses oracle/cpu_ref (C + OpenMP, launch structure of the reference) when it is built, else the
    NumPy oracle."""
    from oracle import ins_oracle as o

    D = wl["D"]
    T = np.float32 if wl["T"] == "float32" else np.float64
    try:
        from oracle import cpu_ref
        have_c = cpu_ref.available()
    except Exception:
        have_c = False
    if n is None:
        n = (256 if D == 3 else 2048) if have_c else (96 if D == 3 else 1024)
    x = tuple(np.linspace(0.0, 1.0, n + 1).astype(T) for _ in range(D))
    bf = None
    if wl["force"]:
        bf = (lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x)
    so = o.Setup(x, Re=T(wl["Re"]), bodyforce=bf)
    rng = np.random.default_rng(0)
    u = np.asfortranarray((0.1 * rng.standard_normal(so.grid.N + (D,))).astype(T))
    dt = wl["dt"] * (512 if D == 3 else 4096) / n
    if have_c:
        cores = cpu_ref.num_threads()
        u = o.apply_bc_u(u, 0.0, so)
        cpu_ref.rk44_steps(u, so, dt, 1, wl["theta"])  # warm-up (page faults, FFT tables)
        t0 = time.perf_counter()
        cpu_ref.rk44_steps(u, so, dt, steps, wl["theta"])
        el = time.perf_counter() - t0
        kind_desc = f"oracle/cpu_ref.c (C+OpenMP, {cores} threads)"
    else:
        cores = 1
        if wl["theta"] is not None:
            so.closure_model = o.smagorinsky_closure(so)
        ps = o.psolver_spectral(so)
        u = o.apply_bc_u(o.project(o.apply_bc_u(u, 0.0, so), so, ps), 0.0, so)
        m = o.RK44(T)
        st = o.create_stepper(m, so, ps, u)
        t0 = time.perf_counter()
        for _ in range(steps):
            st = o.timestep(m, st, T(dt), theta=wl["theta"])
        el = time.perf_counter() - t0
        kind_desc = "oracle/ins_oracle.py (NumPy; scipy.fft on all cores, stencils single-threaded)"
    cells = float(n) ** D
    desc = f"{steps} RK44 step(s) at {'x'.join([str(n)] * D)} {wl['T']}" + (" + Smagorinsky" if wl["theta"] is not None else "") + f", {kind_desc}"
    return cells * steps / el / 1e6, steps / el, cores, desc


# words (sizeof(T)) per cell per LAUNCH of each library kernel: the algorithmic minimum, every
# logical field read / written once (DESIGN.md "Kernels and rooflines")
def kernel_words(D, smag, force, nstages=4, tma=False, zsolve=False, fused2d=False):
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
        "cufft_forward": 2.0 * D,      # one read + one write per transformed axis
        "cufft_inverse": 2.0 * D,
        "import_ghosted": 2.0 * D,
        "export_ghosted": 2.0 * D,
    }
    if tma:
        # 3-D TMA path: the stage kernel also writes the Poisson right-hand side (W div) and applies the previous
        # projection on load (R p).  first stage: R u, W k, W u', W div; middle stages: + R ustart, R p
        first, mid = 3.0 * D + 1, 4.0 * D + 2
        sadd = (D if smag else 0)  # closure term s read by the stage kernel
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


# DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of one `ncu --set full` capture at the
# default workload size; see profiles/r1_ncu_full_kernels_v3.txt
NCU_TRAFFIC = {
    ("dt3d_smag", 512, "smagorinsky_fused"): 2.450518e9 + 1.580941e9,
    ("dt3d_smag", 512, "momentum_stage"): 5.952208e9 + 3.730599e9,
    ("dt3d_smag", 512, "momentum_stage_last"): 11.164976e9 + 2.139155e9,
    ("dt3d_smag", 512, "poisson_scale"): 0.539117e9 + 0.502182e9,
    ("dt3d_smag", 512, "pressure_gradient"): 2.283003e9 + 1.561083e9,
}


def run_reference(args, wl):
    """`--impl reference`: the reference is Julia and cannot run in this image; its CPU algorithm
    restated under oracle/ is timed on the host cores, one bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    times = []
    desc = cores = None
    for i in range(args.warmup + args.steps):
        mc, sps, cores, desc = cpu_sample(wl, steps=1, n=args.cpu_n)
        if i >= args.warmup:
            times.append(mc)
        if sum(1 for _ in times) and i >= args.warmup and time.perf_counter() - T0 > 240:
            break
    v = float(np.mean(times))
    n = weak_grid(wl["D"], wl["n"], args.gpus)
    line = {
        "impl": "reference", "metric": "RK44 Mcell-updates/s", "value": v, "unit": "Mcell-updates/s",
        "n_gpus": args.gpus, "steps": len(times), "warmup": args.warmup,
        "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if wl["T"] == "float32" else "f64", "data": "synthetic",
        "config": {"workload": wl["title"], "grid": list(n), "stepper": "RK44"},
        "cpu_baseline": {"value": v, "unit": "Mcell-updates/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": v, "unit": "Mcell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference is Julia (no runtime in this image): CPU restatement of its algorithm, per-cell rate of a bounded sample",
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
    D = wl["D"]
    T = np.float32 if wl["T"] == "float32" else np.float64
    base_n = args.n or wl["n"]
    n = weak_grid(D, base_n, world)
    x = tuple(np.linspace(0.0, 1.0, m + 1).astype(T) for m in n)
    bf = None
    if wl["force"]:
        bf = (lambda dim, x, y, t: (dim == 1) * 5 * np.sin(np.pi * 8 * y) + 0 * x)
    setup = ins.Setup(x, Re=T(wl["Re"]), bodyforce=bf, backend=backend)
    for kv in args.param:
        k, v = kv.split("=")
        setup.ctx.call("insb200_set_param", k.encode(), float(v))
    theta = wl["theta"]
    if theta is not None:
        setup.closure_model = ins.smagorinsky_closure(setup)
    method = ins.RKMethods.RK44(T=T)
    u = synthetic_field(ins, setup, amp=1e-2 if wl["force"] else 1.0)
    dt = wl["dt"] * wl["n"] / base_n
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
    clocks = sampler.stop(w0, w1) if sampler else None
    ms_step = ms / K
    value = cells / (ms_step * 1e-3) / 1e6
    if not bool(torch.isfinite(u).all()):
        raise SystemExit("bench.py: the solution blew up (non-finite u) -- invalid run")

    # ---- per-kernel timers over an identical region (CUDA events around every launch) ----------
    setup.ctx.call("insb200_reset_timers")
    setup.ctx.call("insb200_set_profiling", 1)
    barrier()
    ins.api._rk_steps(method, setup, u, t, dt, K, theta)
    barrier()
    timers = read_timers(setup)
    setup.ctx.call("insb200_set_profiling", 0)
    if os.environ.get("BENCH_RANK_TIMERS"):  # per-rank view of the communication timers (stderr), for load-balance checks
        sel = ("fft_pack_unpack", "fft_alltoall", "halo_exchange", "cufft_forward", "poisson_scale")
        print(f"[rank {rank}] " + "  ".join(f"{k}={timers[k][0] / max(timers[k][1], 1):.3f}ms" for k in sel if k in timers),
              file=sys.stderr, flush=True)
    esz = np.dtype(T).itemsize
    cells_local = cells / world
    nz_glob = n[-1]
    words = kernel_words(D, theta is not None, wl["force"], tma=(D == 3 and base_n % 32 == 0),
                         zsolve=(D == 3 and 8 <= nz_glob <= 2048 and nz_glob & (nz_glob - 1) == 0),
                         fused2d=(D == 2 and world == 1 and theta is None and "divergence" not in timers))
    peak, peak_src = peaks()
    kern = {}
    for name, (tms, cnt) in timers.items():
        per = tms / max(cnt, 1)
        gbs = words[name] * esz * cells_local / (per * 1e-3) / 1e9 if name in words and per > 0 else None
        kern[name] = {"ms_per_step": tms / K, "launches_per_step": cnt / K, "ms_per_launch": per,
                      "alg_GBs": gbs, "frac": gbs / peak if gbs else None}
    own = {k: v for k, v in kern.items() if not k.startswith("cufft") and k in words}
    dom = max(own, key=lambda k: own[k]["ms_per_step"]) if own else None
    roofline = None
    if dom:
        roofline = {"bound": "hbm", "kernel": dom, "achieved": kern[dom]["alg_GBs"], "peak": peak, "unit": "GB/s",
                    "frac": kern[dom]["frac"], "traffic": NCU_TRAFFIC.get((args.workload, base_n, dom)),
                    "peak_source": peak_src,
                    "alg_bytes_per_launch": words[dom] * esz * cells_local,
                    "ms_per_launch": kern[dom]["ms_per_launch"],
                    "share_of_step": kern[dom]["ms_per_step"] / sum(v["ms_per_step"] for v in kern.values())}
    model_bytes = lib.insb200_model_bytes_per_cell_step(h, 4)
    step_gbs = model_bytes * cells_local / (ms_step * 1e-3) / 1e9

    # ---- e2e: host buffer -> insb200_solve_unsteady_host (H2D, K steps + KE read-back, D2H) ----
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
    e2e = {"value": cells * K / e2e_s / 1e6, "unit": "Mcell-updates/s", "h2d_bytes_per_step": field_bytes / K,
           "d2h_bytes_per_step": field_bytes / K + 8 * (K + 1) / K, "ms_per_step": e2e_s * 1e3 / K,
           "what": "insb200_solve_unsteady_host: pinned host u -> H2D -> K RK44 steps, kinetic energy read back "
                   "after every step -> D2H of u; field copies amortised over K steps",
           "ke_first_last": [float(ke[0]), float(ke[-1])]}

    ws_bytes = int(lib.insb200_workspace_bytes(h))
    barrier()
    setup.ctx.close()  # collective teardown at the same program point on every rank
    if dist is not None:
        dist.destroy_process_group()
    if rank != 0:
        return
    # ---- CPU baseline beside it (rank 0, N = 1 only) ----------------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu:
        mc, sps, cores, desc = cpu_sample(wl, steps=1, n=args.cpu_n)
        cpu = {"value": mc, "unit": "Mcell-updates/s", "cores": cores, "kind": "port", "sample": desc,
               "host_cores": os.cpu_count()}
    line = {
        "metric": "RK44 Mcell-updates/s", "value": value, "unit": "Mcell-updates/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": ms_step, "steps_per_s": 1e3 / ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32" if T == np.float32 else "f64", "data": "synthetic",
        "config": {"workload": wl["title"] if base_n == wl["n"] else wl["title"].replace(f"{wl['n']}^", f"{base_n}^"),
                   "grid": list(n), "cells_per_gpu": cells_local, "stepper": "RK44",
                   "dt": dt, "Re": wl["Re"], "smagorinsky_theta": theta, "parallelism": f"slab{world}(last axis)",
                   "l2": "inputs larger than L2 (state >> 126 MB), no flush",
                   "init": "24 random Fourier modes, Orlandi-type amplitudes, projected on the device"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "cufft_execs": int(ffts),
        "graph_replays": int(graph_replays), "workspace_bytes_per_gpu": ws_bytes,
        "roofline": roofline,
        "step_roofline": {"model_bytes_per_cell_step": model_bytes, "achieved": step_gbs, "peak": peak, "unit": "GB/s",
                          "frac": step_gbs / peak, "note": "SURVEY 8d model (46D+8 words, closure fused) per GPU"},
        "kernels": kern, "cpu_baseline": cpu,
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
