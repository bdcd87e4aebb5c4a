"""Host-side mirror of the reference's operator API for the periodic RK hot path.

Same names, argument meaning and error behaviour as IncompressibleNavierStokes.jl (Julia `f!` is
spelled `f_` here); every function body is a call into libinsb200.so through the C ABI of
include/insb200.h -- the Python code holds no arithmetic of the path.  torch is used for device
memory and streams only.

Field representation: a torch CUDA tensor whose memory is the Julia array, i.e. the Julia array
`(N1,..,ND[,D])` (column-major) is a C-contiguous tensor of shape `([D,] ND,..,N1)`.
`to_device(a)` / `to_host(t)` convert from / to Julia-shaped NumPy arrays.
"""
from __future__ import annotations

import ctypes as C
import math
import os
from types import SimpleNamespace

import numpy as np
import torch

from . import _lib

__all__ = [
    "B200Backend", "Setup", "scalarfield", "vectorfield", "to_device", "to_host",
    "apply_bc_u_", "apply_bc_p_", "apply_bc_u", "apply_bc_p", "momentum_", "momentum", "convection",
    "diffusion", "convectiondiffusion", "divergence_", "divergence", "scalewithvolume_",
    "pressuregradient_", "pressuregradient", "applypressure_", "applypressure", "psolver_spectral",
    "default_psolver", "poisson_", "poisson", "project_", "project", "pressure_", "pressure",
    "smagorinsky_closure", "total_kinetic_energy", "get_cfl_timestep_", "maxabs", "RKMethods",
    "runge_kutta_method", "ode_method_cache", "create_stepper", "timestep_", "solve_unsteady",
    "solve_unsteady_host", "init_distributed", "local_slab", "gather_slabs",
]


# --------------------------------------------------------------------------------------------
# backend / distributed bring-up
# --------------------------------------------------------------------------------------------
class B200Backend:
    """The `backend` value stored in `Setup` (reference: `CPU()` / `CUDABackend()`,
    src/setup.jl:9).  One backend = one GPU = one slab of the last axis."""

    def __init__(self, device=0, rank=0, nranks=1, nccl_id=None):
        self.device, self.rank, self.nranks, self.nccl_id = int(device), int(rank), int(nranks), nccl_id

    def __repr__(self):
        return f"B200Backend(device={self.device}, rank={self.rank}, nranks={self.nranks})"


def init_distributed():
    """Read RANK / LOCAL_RANK / WORLD_SIZE (torchrun), create the NCCL unique id on rank 0 and
    broadcast it with torch.distributed.  Returns a B200Backend."""
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world == 1:
        return B200Backend(device=local)
    torch.cuda.set_device(local)
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _lib.load()
    buf = (C.c_ubyte * 128)()
    if rank == 0:
        _lib.check(lib, None, lib.insb200_comm_unique_id(buf))
    t = torch.tensor(list(bytes(buf)), dtype=torch.uint8, device=f"cuda:{local}")
    dist.broadcast(t, 0)
    return B200Backend(device=local, rank=rank, nranks=world, nccl_id=bytes(t.cpu().tolist()))


# --------------------------------------------------------------------------------------------
# Setup / Grid   (src/setup.jl:2-46, src/grid.jl:100-276)
# --------------------------------------------------------------------------------------------
class _Ctx:
    def __init__(self, D, n, dtype, h, Re, backend):
        self.lib = _lib.load()
        self.h = C.c_void_p()
        nn = (C.c_int * 3)(*([int(v) for v in n] + [1] * (3 - D)))
        hh = (C.c_double * 3)(*([float(v) for v in h] + [1.0] * (3 - D)))
        idbuf = None
        if backend.nccl_id is not None:
            idbuf = (C.c_ubyte * 128).from_buffer_copy(backend.nccl_id)
        rc = self.lib.insb200_create(C.byref(self.h), D, nn, 0 if dtype == np.float32 else 1, hh, float(Re),
                                     backend.device, backend.rank, backend.nranks, idbuf)
        if rc != 0:
            self.h = None
            raise _lib.INSB200Error(self.lib.insb200_last_error(None).decode())

    def call(self, name, *args):
        rc = getattr(self.lib, name)(self.h, *args)
        _lib.check(self.lib, self.h, rc)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.lib.insb200_destroy(self.h)
                self.h = None
        except Exception:
            pass


def _padghost_periodic(x):
    x = np.concatenate(([x[0] - (x[-1] - x[-2])], x))
    return np.concatenate((x, [x[-1] + (x[1] - x[0])]))


def Setup(x, boundary_conditions=None, bodyforce=None, issteadybodyforce=True, closure_model=None,
          backend=None, Re=None, temperature=None, workgroupsize=64):
    """`Setup(; x, boundary_conditions, bodyforce, issteadybodyforce, closure_model, backend, Re)`
    (src/setup.jl:2-46).  `x`: tuple of GLOBAL 1-D face coordinate vectors; their dtype selects
    Float32 / Float64 (docs/src/manual/precision.md).  Only the hot path's admissible domain is
    accepted -- periodic, uniform, even n (src/utils.jl:1-13); anything else raises like the
    reference's `assert_uniform_periodic`."""
    if boundary_conditions is not None:
        for bc in boundary_conditions:
            if tuple(bc) != ("periodic", "periodic"):
                raise AssertionError("B200 backend requires periodic boundary conditions.")
    if temperature is not None:
        raise NotImplementedError("temperature equation is outside the B200 hot path (SURVEY 8f)")
    backend = backend or B200Backend()
    x = tuple(np.asarray(xa) for xa in x)
    T = x[0].dtype.type
    if T not in (np.float32, np.float64):
        raise TypeError("grid coordinates must be float32 or float64")
    D = len(x)
    if D not in (2, 3):
        raise AssertionError("dimension must be 2 or 3")
    n = tuple(len(xa) - 1 for xa in x)
    for a in range(D):
        d = np.diff(x[a])
        if not np.allclose(d, d[0], rtol=1e-5 if T == np.float32 else 1e-8):
            raise AssertionError("B200 backend requires uniform grid spacing.")
        if n[a] % 2:
            raise AssertionError("B200 backend requires even number of volumes.")
    Re = T(1000) if Re is None else T(Re)
    xg = tuple(_padghost_periodic(xa).astype(T) for xa in x)
    h = tuple(T(np.diff(xa)[0]) for xa in xg)  # Delta[a][1]
    ctx = _Ctx(D, n, T, h, Re, backend)
    nloc = (C.c_int * 3)()
    off = C.c_int()
    ctx.lib.insb200_local_geometry(ctx.h, nloc, C.byref(off))
    nloc = tuple(nloc[a] for a in range(D))
    # local coordinates (the slab of the last axis owned by this rank, plus its ghosts)
    xl = list(xg)
    xl[D - 1] = xg[D - 1][off.value: off.value + nloc[D - 1] + 3]
    xl = tuple(xl)
    N = tuple(m + 2 for m in nloc)
    grid = SimpleNamespace(
        dimension=D, N=N, Np=nloc, n_global=n, offset_last=off.value, T=T,
        Ip=tuple(slice(1, m - 1) for m in N), Iu=tuple(tuple(slice(1, m - 1) for m in N) for _ in range(D)),
        x=xl, xp=tuple(((v[:-1] + v[1:]) / T(2)).astype(T) for v in xl),
        xu=tuple(tuple(xl[b][1:] if a == b else ((xl[b][:-1] + xl[b][1:]) / T(2)).astype(T) for b in range(D)) for a in range(D)),
        Delta=tuple(np.full(N[a], h[a], dtype=T) for a in range(D)), h=h,
        xlims=tuple((xa[0], xa[-1]) for xa in x),
    )
    setup = SimpleNamespace(grid=grid, boundary_conditions=boundary_conditions, Re=Re, bodyforce=None,
                            issteadybodyforce=False, closure_model=closure_model, backend=backend,
                            workgroupsize=workgroupsize, temperature=None, ctx=ctx,
                            device=torch.device("cuda", backend.device),
                            tdtype=torch.float32 if T == np.float32 else torch.float64)
    if bodyforce is not None:
        if not issteadybodyforce:
            raise NotImplementedError("only steady body forces are on the B200 hot path (setup.jl:25-32)")
        # setup.jl:25-32 / operators.jl:870-877: F[.., a] = bodyforce(a, xu[a]..., 0) over the whole array
        F = np.zeros(N + (D,), dtype=T, order="F")
        for a in range(D):
            xin = []
            for b in range(D):
                sh = [1] * D
                sh[b] = -1
                xin.append(grid.xu[a][b].reshape(sh))
            F[..., a] = np.broadcast_to(np.asarray(bodyforce(a + 1, *xin, T(0)), dtype=T), N)
        setup.bodyforce = to_device(F, setup)
        setup.issteadybodyforce = True
        ctx.call("insb200_set_bodyforce", _ptr(setup.bodyforce))
    return setup
