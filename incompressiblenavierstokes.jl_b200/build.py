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
SOURCES = [os.path.join(CSRC, "insb200.cu")]
HEADERS = [os.path.join(CSRC, f) for f in ("ctx.cuh", "stencil.cuh", "tiled.cuh", "tma_stage.cuh", "zsolve.cuh")] + [
    os.path.join(os.path.dirname(HERE), "include", "insb200.h")
]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false",                      # arithmetic contract of stencil.cuh: no FMA contraction
    "-Xcompiler", "-fPIC", "-shared",
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


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(f) > t for f in SOURCES + HEADERS + [os.path.abspath(__file__)])


def build(force=False, verbose=False, extra_flags=()):
    if not force and not needs_build():
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + list(extra_flags) + ["-o", LIB] + SOURCES + ["-lcufft", "-lnccl"]
    for p in _rpaths():
        cmd += ["-Xlinker", "-rpath=" + p]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose:
        print(res.stdout, res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
