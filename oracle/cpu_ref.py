"""ctypes wrapper of oracle/_build/libcpuref.so (oracle/cpu_ref.c): the reference's CPU launch structure in
C + OpenMP, FFTs through scipy.fft (pocketfft, all workers).

TEST / BASELINE INFRASTRUCTURE ONLY: imported by tests/ (checked against oracle/ins_oracle.py) and by the
`cpu_baseline` / `--impl reference` legs of bench.py.  The product never loads it.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import scipy.fft as sfft

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libcpuref.so")
_lib = None
_CB = C.CFUNCTYPE(None, C.c_void_p)


def build():
    subprocess.run(["make", "-s", "-C", HERE], check=True)
    return LIB


def available():
    return os.path.exists(LIB)


def _load():
    global _lib
    if _lib is None:
        if not available():
            build()
        _lib = C.CDLL(LIB)
        for name, T in (("cpuref_rk_steps_f32", C.c_float), ("cpuref_rk_steps_f64", C.c_double)):
            fn = getattr(_lib, name)
            fn.restype = C.c_int
            fn.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_double, C.c_int,
                           C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_double, C.c_void_p, _CB]
        _lib.cpuref_num_threads.restype = C.c_int
        _lib.cpuref_set_num_threads.argtypes = [C.c_int]
        _lib.cpuref_set_num_threads.restype = None
    return _lib


def num_threads():
    return int(_load().cpuref_num_threads())


def set_num_threads(n):
    """OpenMP threads of the C loops (the scipy.fft workers follow `workers=` of rk_steps).  Callers that want
    all host cores must say so: torch.distributed.run exports OMP_NUM_THREADS=1 to its workers."""
    _load().cpuref_set_num_threads(int(n))
    return num_threads()


def _psolver_inplace(setup, workers):
    """psolve!(pI) of psolver_spectral (src/pressure.jl:318-350) on the ghost-stripped array, in place.
    Same symbol / operation order as oracle.ins_oracle.psolver_spectral."""
    from . import ins_oracle as o

    g = setup.grid
    D, T, Np = g.dimension, g.T, g.Np
    ahat = o.psolver_spectral(setup).ahat
    den = ahat[0].reshape([-1] + [1] * (D - 1)) + ahat[1].reshape([1, -1] + [1] * (D - 2))
    if D == 3:
        den = den + ahat[2].reshape(1, 1, -1)
    den = np.ascontiguousarray(den.T)  # C order of the reversed-axes view
    ctype = np.complex64 if T == np.float32 else np.complex128
    shape_c = tuple(reversed(Np))      # Julia (n1,..,nD) column-major == C array (nD,..,n1)
    axes = tuple(range(D))             # last C axis = first Julia axis = the halved one

    def cb(ptr):
        a = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_float if T == np.float32 else C.c_double)), shape=shape_c)
        phat = sfft.rfftn(a, axes=axes, workers=workers).astype(ctype, copy=False)
        with np.errstate(divide="ignore", invalid="ignore"):
            phat = (-phat.real / den + 1j * (-phat.imag / den)).astype(ctype)
        phat[(0,) * D] = 0
        a[...] = sfft.irfftn(phat, s=shape_c, axes=axes, workers=workers).astype(T, copy=False)

    return cb


def rk_steps(u, setup, dt, nsteps, method, theta=None, workers=-1):
    """Advance the Julia-shaped Fortran-order array `u` (N1,..,ND,D) in place by `nsteps` explicit RK steps.
    `setup` is an oracle Setup (grid, Re, stored body force); `method` an oracle RK method (shifted tableau)."""
    lib = _load()
    g = setup.grid
    D, T = g.dimension, g.T
    assert u.flags.f_contiguous and u.dtype == T
    n = (C.c_int * 3)(*(list(g.Np) + [1] * (3 - D)))
    h = (C.c_double * 3)(*([float(g.Delta[a][0]) for a in range(D)] + [1.0] * (3 - D)))
    s = len(method.b)
    A = np.ascontiguousarray(np.asarray(method.A, dtype=np.float64)).ravel()
    c = np.ascontiguousarray(np.asarray(method.c, dtype=np.float64))
    force = None
    if setup.bodyforce is not None:
        force = np.asfortranarray(setup.bodyforce, dtype=T)
    cb = _CB(_psolver_inplace(setup, workers))
    fn = lib.cpuref_rk_steps_f32 if T == np.float32 else lib.cpuref_rk_steps_f64
    rc = fn(u.ctypes.data, D, n, h, float(setup.Re), float(dt), int(nsteps), s, A.ctypes.data_as(C.POINTER(C.c_double)),
            c.ctypes.data_as(C.POINTER(C.c_double)), float("nan") if theta is None else float(theta),
            force.ctypes.data if force is not None else None, cb)
    assert rc == 0
    return u


def rk44_steps(u, setup, dt, nsteps, theta=None, workers=-1):
    from . import ins_oracle as o

    return rk_steps(u, setup, dt, nsteps, o.RK44(setup.grid.T), theta, workers=workers)
