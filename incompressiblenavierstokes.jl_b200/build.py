"""In-tree build of libinsb200.so (sm_100a only, no multi-arch, no JIT cache).

`python incompressiblenavierstokes.jl_b200/build.py [--force]` or `build()` from
`__graft_entry__.py`.  nvcc cross-compiles without a GPU.  The library links cuFFT (transforms of
the spectral pressure solve) and NCCL (halos, FFT transpose); the CUDA runtime is linked statically.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libinsb200.so")
SOURCES = [os.path.join(CSRC, "insb200.cu"), os.path.join(CSRC, "temperature.cu"), os.path.join(CSRC, "fast.cu")]
# translation units built WITH FMA contraction: the default (tolerance-parity) kernels; everything else keeps the
# reference's operation order bit for bit (-fmad=false)
CONTRACTED = {"fast.cu"}
OBJDIR = os.path.join(HERE, "build")  # git-ignored object cache: only stale translation units are recompiled
HEADERS = [os.path.join(CSRC, f) for f in ("ctx.cuh", "stencil.cuh", "tiled.cuh", "tma_stage.cuh", "tma_stage_x2.cuh", "f32x2.cuh", "stage2d.cuh", "zsolve.cuh", "zsolve_launch.cuh", "comm.cuh", "fast.cuh", "fast_stage.cuh", "zline.cuh")] + [
    os.path.join(os.path.dirname(HERE), "include", "insb200.h")
]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false",                      # arithmetic contract of stencil.cuh: no FMA contraction
    "-Xcompiler", "-fPIC",
]


def _rpaths():
    paths = ["/usr/local/cuda/lib64"]
    try:  # prefer the NCCL / cuFFT that torch ships when they are present (same image on the GPU box)
        import importlib.util
        for mod in ("nvidia.nccl", "nvidia.cufft"):
            spec = importlib.util.find_spec(mod)
            if spec and spec.submodule_search_locations:
                paths.append(os.path.join(list(spec.submodule_search_locations)[0], "lib"))
    except Exception:
        pass
    return paths


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(f) > t for f in deps)


def needs_build():
    return _stale(LIB, SOURCES + HEADERS + [os.path.abspath(__file__)])


def build(force=False, verbose=False, extra_flags=()):
    """extra_flags=("-DINSB200_EXPERIMENTAL",) also compiles the recorded negative results (packed-FP32 stage kernel
    k_stage3d_tma_x2, fused peer z-solve k_zsolve_peer); the shipped library leaves them out.  extra_flags=("-DINSB200_DEBUG",)
    bounds every mbarrier wait and traps with a message instead of hanging (tma_stage.cuh)."""
    if not force and not needs_build():
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    # one nvcc -c per translation unit (run concurrently), then one link
    jobs, objs = [], []
    for src in SOURCES:
        obj = os.path.join(OBJDIR, os.path.splitext(os.path.basename(src))[0] + ".o")
        objs.append(obj)
        if not force and not extra_flags and not _stale(obj, [src] + HEADERS + [os.path.abspath(__file__)]):
            continue
        flags = [f for f in NVCC_FLAGS if not (os.path.basename(src) in CONTRACTED and f == "-fmad=false")]
        cmd = [nvcc] + flags + list(extra_flags) + ["-c", "-o", obj, src]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd))
        jobs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = []
    for src, pr in jobs:
        out, _ = pr.communicate()
        if verbose:
            print(out)
        if pr.returncode != 0:
            failed.append("nvcc failed on " + src + ":\n" + out)
    if failed:
        raise RuntimeError("\n".join(failed))
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC", "-o", LIB] + objs + ["-lcufft", "-lnccl"]
    for p in _rpaths():
        cmd += ["-Xlinker", "-rpath=" + p]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose:
        print(" ".join(cmd))
        print(res.stdout, res.stderr)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + res.stdout + res.stderr)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
