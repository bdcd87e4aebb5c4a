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
    "smagorinsky_closure", "vorticity_", "vorticity", "interpolate_u_p_", "interpolate_u_p", "interpolate_w_p_",
    "interpolate_w_p", "velocitynorm_", "total_kinetic_energy", "get_cfl_timestep_", "maxabs", "RKMethods", "LMWray3",
    "runge_kutta_method", "ode_method_cache", "create_stepper", "timestep_", "solve_unsteady",
    "solve_unsteady_host", "init_distributed", "local_slab", "gather_slabs",
    "temperature_equation", "apply_bc_temp_", "apply_bc_temp", "convection_diffusion_temp_", "convection_diffusion_temp",
    "dissipation_", "dissipation", "gravity_", "gravity", "right_hand_side_", "create_right_hand_side", "Qfield_", "Qfield",
    "dissipation_from_strain_", "dissipation_from_strain", "get_scale_numbers",
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
    return B200Backend(device=local, rank=rank, nranks=world, nccl_id=_fresh_nccl_id(rank, local))


def _fresh_nccl_id(rank, device):
    """A new NCCL unique id from rank 0, broadcast over torch.distributed: every communicator (one per
    context) needs its own."""
    import torch.distributed as dist

    lib = _lib.load()
    buf = (C.c_ubyte * 128)()
    if rank == 0:
        _lib.check(lib, None, lib.insb200_comm_unique_id(buf))
    t = torch.tensor(list(bytes(buf)), dtype=torch.uint8, device=f"cuda:{device}")
    dist.broadcast(t, 0)
    return bytes(t.cpu().tolist())


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
        if backend.nranks > 1:
            if getattr(backend, "_id_used", False):  # a unique id serves exactly one communicator
                backend.nccl_id = _fresh_nccl_id(backend.rank, backend.device)
            backend._id_used = True
            idbuf = (C.c_ubyte * 128).from_buffer_copy(backend.nccl_id)
        rc = self.lib.insb200_create(C.byref(self.h), D, nn, 0 if dtype == np.float32 else 1, hh, float(Re),
                                     backend.device, backend.rank, backend.nranks, idbuf)
        if rc != 0:
            self.h = None
            raise _lib.INSB200Error(self.lib.insb200_last_error(None).decode())

    def call(self, name, *args):
        rc = getattr(self.lib, name)(self.h, *args)
        _lib.check(self.lib, self.h, rc)

    def close(self):
        """insb200_destroy: frees the workspace, plans and the NCCL communicator.  With nranks > 1 call it on
        every rank at the same point of the program (communicator teardown)."""
        if getattr(self, "h", None):
            self.lib.insb200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
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
    backend = backend or B200Backend()
    if temperature is not None and backend.nranks > 1:
        raise NotImplementedError("the temperature equation runs on one rank only (SURVEY 8f rank 4)")
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
    if Re is None:  # setup.jl:12: Re = isnothing(temperature) ? 1000 : 1 / temperature.α1
        Re = T(1000) if temperature is None else T(1) / T(temperature.alpha1)
    Re = T(Re)
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
                            workgroupsize=workgroupsize, temperature=temperature, ctx=ctx,
                            device=torch.device("cuda", backend.device),
                            tdtype=torch.float32 if T == np.float32 else torch.float64)
    if temperature is not None:
        q = temperature
        if getattr(q, "boundary_conditions", None) is not None:
            for bc in q.boundary_conditions:
                if tuple(bc) != ("periodic", "periodic"):
                    raise AssertionError("B200 backend requires periodic temperature boundary conditions.")
        ctx.call("insb200_set_temperature", float(T(q.alpha1)), float(T(q.alpha2)), float(T(q.alpha3)), float(T(q.alpha4)),
                 float(T(q.gamma)), int(bool(q.dodissipation)), int(q.gdir))
    if bodyforce is not None:
        def force_field(t):
            # setup.jl:25-32 / operators.jl:870-877: F[.., a] = bodyforce(a, xu[a]..., t) over the whole array
            F = np.zeros(N + (D,), dtype=T, order="F")
            for a in range(D):
                xin = []
                for b in range(D):
                    sh = [1] * D
                    sh[b] = -1
                    xin.append(grid.xu[a][b].reshape(sh))
                # scalar broadcast `bodyforce.(a, x..., t)` like Julia: one scalar call per point (an
                # array call could take NumPy's SIMD sin/cos path, which differs in the last bit)
                val = np.vectorize(lambda *xs, _a=a: bodyforce(_a + 1, *xs, T(t)), otypes=[T])(*np.broadcast_arrays(*xin))
                F[..., a] = val.astype(T)
            return to_device(F, setup)

        if issteadybodyforce:
            setup.bodyforce = force_field(0)
            setup.issteadybodyforce = True
            ctx.call("insb200_set_bodyforce", _ptr(setup.bodyforce))
        else:
            # unsteady force (operators.jl:870-877): the library asks for bodyforce.(a, x..., t) at every stage time;
            # this host mirror evaluates the Python callable on the host and uploads the field (a Julia caller
            # broadcasts on the device)
            setup.bodyforce = bodyforce
            setup.issteadybodyforce = False

            def _cb(user, t, field_out):
                try:
                    setup._force_now = force_field(t)  # kept alive until the next invocation; the upload is synchronous
                    field_out[0] = setup._force_now.data_ptr()
                    return 0
                except Exception as e:  # noqa: BLE001 -- never raise across the C boundary
                    setup._force_error = e
                    return 1

            setup._force_cb = _lib.FORCE_FN(_cb)  # keep the trampoline alive as long as the setup
            ctx.call("insb200_set_bodyforce_callback", setup._force_cb, None)
    return setup


# --------------------------------------------------------------------------------------------
# fields
# --------------------------------------------------------------------------------------------
def _ptr(t):
    if t is None:
        return None
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.is_contiguous()):
        raise TypeError("fields must be contiguous CUDA tensors (see to_device)")
    return C.c_void_p(t.data_ptr())


def _check(t, setup, vector):
    g = setup.grid
    shape = tuple(reversed(g.N)) if not vector else (g.dimension,) + tuple(reversed(g.N))
    if tuple(t.shape) != shape or t.dtype != setup.tdtype:
        raise ValueError(f"field has shape {tuple(t.shape)} / {t.dtype}, expected {shape} / {setup.tdtype}")
    return t


def scalarfield(setup):
    """src/initializers.jl:1-2"""
    return torch.zeros(tuple(reversed(setup.grid.N)), dtype=setup.tdtype, device=setup.device)


def vectorfield(setup):
    """src/initializers.jl:4-6"""
    g = setup.grid
    return torch.zeros((g.dimension,) + tuple(reversed(g.N)), dtype=setup.tdtype, device=setup.device)


def to_device(a, setup):
    """Julia-shaped NumPy array `(N1,..,ND[,D])` -> device field (same memory image)."""
    a = np.asfortranarray(np.asarray(a, dtype=setup.grid.T))
    return torch.from_numpy(np.ascontiguousarray(a.T)).to(setup.device)


def to_host(t):
    """device field -> Julia-shaped NumPy array (Fortran order)."""
    return t.detach().cpu().numpy().T


def _sync_stream(setup):
    # run the library on torch's current stream so that torch ops and kernels are ordered
    s = torch.cuda.current_stream(setup.device).cuda_stream
    setup.ctx.call("insb200_set_stream", C.c_void_p(s))


# --------------------------------------------------------------------------------------------
# operators (in-place `f_` = Julia `f!`, allocating twins without the underscore)
# --------------------------------------------------------------------------------------------
def apply_bc_u_(u, t, setup):
    """apply_bc_u!(u, t, setup)  src/boundary_conditions.jl:159-167"""
    _sync_stream(setup)
    setup.ctx.call("insb200_apply_bc_u", _ptr(_check(u, setup, True)))
    return u


def apply_bc_p_(p, t, setup):
    """apply_bc_p!(p, t, setup)  src/boundary_conditions.jl:197-206"""
    _sync_stream(setup)
    setup.ctx.call("insb200_apply_bc_p", _ptr(_check(p, setup, False)))
    return p


def apply_bc_u(u, t, setup):
    return apply_bc_u_(u.clone(), t, setup)


def apply_bc_p(p, t, setup):
    return apply_bc_p_(p.clone(), t, setup)


def momentum_(F, u, temp, t, setup):
    """momentum!(F, u, temp, t, setup)  src/operators.jl:967-976"""
    _sync_stream(setup)
    if temp is None:
        setup.ctx.call("insb200_momentum", _ptr(_check(F, setup, True)), _ptr(_check(u, setup, True)), float(t))
    else:
        setup.ctx.call("insb200_momentum_temp", _ptr(_check(F, setup, True)), _ptr(_check(u, setup, True)),
                       _ptr(_check(temp, setup, False)), float(t))
    return F


def momentum(u, temp, t, setup):
    return momentum_(vectorfield(setup), u, temp, t, setup)


def _convdiff(u, setup, terms):
    F = vectorfield(setup)
    _sync_stream(setup)
    setup.ctx.call("insb200_convectiondiffusion", _ptr(F), _ptr(_check(u, setup, True)), terms)
    return F


def convection(u, setup):
    """convection(u, setup)  src/operators.jl:366-415"""
    return _convdiff(u, setup, 1)


def diffusion(u, setup):
    """diffusion(u, setup)  src/operators.jl:537-573"""
    return _convdiff(u, setup, 2)


def convectiondiffusion(u, setup):
    """convectiondiffusion!(zero(u), u, setup)  src/operators.jl:634-690"""
    return _convdiff(u, setup, 3)


def divergence_(div, u, setup):
    """divergence!(div, u, setup)  src/operators.jl:106-115"""
    _sync_stream(setup)
    setup.ctx.call("insb200_divergence", _ptr(_check(div, setup, False)), _ptr(_check(u, setup, True)))
    return div


def divergence(u, setup):
    return divergence_(scalarfield(setup), u, setup)


def _wfield(setup):
    """vorticity is a scalar field in 2-D and a vector field in 3-D (operators.jl:978-983)"""
    return scalarfield(setup) if setup.grid.dimension == 2 else vectorfield(setup)


def vorticity_(w, u, setup):
    """vorticity!(w, u, setup)  src/operators.jl:985-1021 (ghosts of u must be valid)"""
    _sync_stream(setup)
    setup.ctx.call("insb200_vorticity", _ptr(_check(w, setup, setup.grid.dimension == 3)), _ptr(_check(u, setup, True)))
    return w


def vorticity(u, setup):
    return vorticity_(_wfield(setup), u, setup)


def interpolate_u_p_(up, u, setup):
    """interpolate_u_p!(up, u, setup)  src/operators.jl:1311-1326"""
    _sync_stream(setup)
    setup.ctx.call("insb200_interpolate_u_p", _ptr(_check(up, setup, True)), _ptr(_check(u, setup, True)))
    return up


def interpolate_u_p(u, setup):
    return interpolate_u_p_(vectorfield(setup), u, setup)


def interpolate_w_p_(wp, w, setup):
    """interpolate_w_p!(wp, w, setup)  src/operators.jl:1338-1372"""
    vec = setup.grid.dimension == 3
    _sync_stream(setup)
    setup.ctx.call("insb200_interpolate_w_p", _ptr(_check(wp, setup, vec)), _ptr(_check(w, setup, vec)))
    return wp


def interpolate_w_p(w, setup):
    return interpolate_w_p_(_wfield(setup), w, setup)


def velocitynorm_(out, up, setup):
    """`@. upnorm = sqrt(up1^2 + up2^2 [+ up3^2])`  src/processors.jl:141-151"""
    _sync_stream(setup)
    setup.ctx.call("insb200_velocitynorm", _ptr(_check(out, setup, False)), _ptr(_check(up, setup, True)))
    return out


def scalewithvolume_(p, setup):
    """scalewithvolume!(p, setup)  src/operators.jl:81-95"""
    _sync_stream(setup)
    setup.ctx.call("insb200_scalewithvolume", _ptr(_check(p, setup, False)))
    return p


def pressuregradient_(G, p, setup):
    """pressuregradient!(G, p, setup)  src/operators.jl:159-168"""
    _sync_stream(setup)
    setup.ctx.call("insb200_pressuregradient", _ptr(_check(G, setup, True)), _ptr(_check(p, setup, False)))
    return G


def pressuregradient(p, setup):
    return pressuregradient_(vectorfield(setup), p, setup)


def applypressure_(u, p, setup):
    """applypressure!(u, p, setup)  src/operators.jl:214-223"""
    _sync_stream(setup)
    setup.ctx.call("insb200_applypressure", _ptr(_check(u, setup, True)), _ptr(_check(p, setup, False)))
    return u


def applypressure(u, p, setup):
    return applypressure_(u.clone(), p, setup)


def psolver_spectral(setup):
    """psolver_spectral(setup)  src/pressure.jl:289-351: returns `psolve!(p)` (in place)."""

    def psolve_(p):
        _sync_stream(setup)
        setup.ctx.call("insb200_psolve_spectral", _ptr(_check(p, setup, False)))
        return p

    psolve_.setup = setup
    return psolve_


def default_psolver(setup):
    """default_psolver(setup)  src/pressure.jl:85-98 (periodic + uniform -> spectral)."""
    return psolver_spectral(setup)


def poisson_(psolver, f):
    """poisson!(psolver, f)  src/pressure.jl:22"""
    return psolver(f)


def poisson(psolver, f):
    return psolver(f.clone())


def project_(u, setup, psolver=None, p=None):
    """project!(u, setup; psolver, p)  src/pressure.jl:69-82"""
    _sync_stream(setup)
    setup.ctx.call("insb200_project", _ptr(_check(u, setup, True)), _ptr(p) if p is not None else None)
    return u


def project(u, setup, psolver=None):
    return project_(u.clone(), setup, psolver)


def pressure_(p, u, temp, t, setup, psolver=None, F=None):
    """pressure!(p, u, temp, t, setup; psolver, F)  src/pressure.jl:41-49"""
    if temp is not None:
        raise NotImplementedError("pressure! with a temperature field is not on the B200 path yet")
    _sync_stream(setup)
    setup.ctx.call("insb200_pressure", _ptr(_check(p, setup, False)), _ptr(_check(u, setup, True)), float(t))
    return p


def pressure(u, temp, t, setup, psolver=None):
    return pressure_(scalarfield(setup), u, temp, t, setup, psolver)


def smagorinsky_closure(setup):
    """smagorinsky_closure(setup)  src/operators.jl:1294-1305: m(u, theta) -> s (same buffer
    every call, like the reference)."""
    s = vectorfield(setup)

    def closure(u, theta):
        _sync_stream(setup)
        setup.ctx.call("insb200_smagorinsky", _ptr(s), _ptr(_check(u, setup, True)), float(theta))
        return s

    closure.is_smagorinsky = True
    return closure


def temperature_equation(Pr, Ra, Ge, dodissipation=True, boundary_conditions=None, gdir=2, nondim_type=1):
    """temperature_equation(; Pr, Ra, Ge, dodissipation, boundary_conditions, gdir, nondim_type)  src/setup.jl:56-86.
    The coefficients take the dtype of `Pr` (NumPy scalar arithmetic, like `eltype(Pr)` in the reference)."""
    T = np.asarray(Pr).dtype.type
    if T not in (np.float32, np.float64):
        T = np.float64
    Pr, Ra, Ge = T(Pr), T(Ra), T(Ge)
    if nondim_type == 1:      # free fall velocity
        a1, a2, a3, a4 = np.sqrt(Pr / Ra), T(1), Ge * np.sqrt(Pr / Ra), T(1) / np.sqrt(Pr * Ra)
    elif nondim_type == 2:    # heat conduction time scale
        a1, a2, a3, a4 = Pr, Pr * Ra, Ge / Ra, T(1)
    elif nondim_type == 3:
        a1, a2, a3, a4 = np.sqrt(Pr * Ge / Ra), Ge, np.sqrt(Pr * Ge / Ra), np.sqrt(Ge / (Pr * Ra))
    else:
        raise ValueError("nondim_type must be 1, 2 or 3")
    a1, a2, a3, a4 = T(a1), T(a2), T(a3), T(a4)
    return SimpleNamespace(alpha1=a1, alpha2=a2, alpha3=a3, alpha4=a4, gamma=T(a1 / a3), dodissipation=bool(dodissipation),
                           boundary_conditions=boundary_conditions, gdir=int(gdir))


def _need_temperature(setup):
    if setup.temperature is None:
        raise ValueError("setup.temperature is nothing: pass temperature=temperature_equation(...) to Setup")


def apply_bc_temp_(temp, t, setup):
    """apply_bc_temp!(temp, t, setup)  src/boundary_conditions.jl:236-246"""
    _sync_stream(setup)
    setup.ctx.call("insb200_apply_bc_temp", _ptr(_check(temp, setup, False)))
    return temp


def apply_bc_temp(temp, t, setup):
    return apply_bc_temp_(temp.clone(), t, setup)


def convection_diffusion_temp_(c, u, temp, setup):
    """convection_diffusion_temp!(c, u, temp, setup)  src/operators.jl:711-735 (adds to c)"""
    _need_temperature(setup)
    _sync_stream(setup)
    setup.ctx.call("insb200_convection_diffusion_temp", _ptr(_check(c, setup, False)), _ptr(_check(u, setup, True)),
                   _ptr(_check(temp, setup, False)))
    return c


def convection_diffusion_temp(u, temp, setup):
    return convection_diffusion_temp_(scalarfield(setup), u, temp, setup)


def dissipation_(diss, diff, u, setup):
    """dissipation!(diss, diff, u, setup)  src/operators.jl:787-808 (adds to diss; diff is scratch)"""
    _need_temperature(setup)
    _sync_stream(setup)
    setup.ctx.call("insb200_dissipation", _ptr(_check(diss, setup, False)), _ptr(_check(diff, setup, True)),
                   _ptr(_check(u, setup, True)))
    return diss


def dissipation(u, setup):
    return dissipation_(scalarfield(setup), vectorfield(setup), u, setup)


def gravity_(F, temp, setup):
    """gravity!(F, temp, setup)  src/operators.jl:916-931 (adds to F)"""
    _need_temperature(setup)
    _sync_stream(setup)
    setup.ctx.call("insb200_gravity", _ptr(_check(F, setup, True)), _ptr(_check(temp, setup, False)))
    return F


def gravity(temp, setup):
    return gravity_(vectorfield(setup), temp, setup)


def right_hand_side_(dudt, u, params_ref, t):
    """right_hand_side!(dudt, u, params_ref, t)  src/sciml.jl:37-49; params_ref = (setup, psolver)."""
    setup = params_ref[0]
    _sync_stream(setup)
    setup.ctx.call("insb200_right_hand_side", _ptr(_check(dudt, setup, True)), _ptr(_check(u, setup, True)), float(t))
    return None


def create_right_hand_side(setup, psolver=None):
    """create_right_hand_side(setup, psolver)  src/sciml.jl:13-19: returns f(u, param, t) -> dudt."""

    def right_hand_side(u, param, t):
        dudt = vectorfield(setup)
        right_hand_side_(dudt, u, (setup, psolver), t)
        return dudt

    return right_hand_side


def Qfield_(Q, u, setup):
    """Qfield!(Q, u, setup)  src/operators.jl:1441-1460"""
    _sync_stream(setup)
    setup.ctx.call("insb200_qfield", _ptr(_check(Q, setup, False)), _ptr(_check(u, setup, True)))
    return Q


def Qfield(u, setup):
    return Qfield_(scalarfield(setup), u, setup)


def dissipation_from_strain_(eps, u, setup):
    """dissipation_from_strain!(eps, u, setup)  src/operators.jl:819-836"""
    _sync_stream(setup)
    setup.ctx.call("insb200_dissipation_from_strain", _ptr(_check(eps, setup, False)), _ptr(_check(u, setup, True)))
    return eps


def dissipation_from_strain(u, setup):
    return dissipation_from_strain_(scalarfield(setup), u, setup)


def get_scale_numbers(u, setup):
    """get_scale_numbers(u, setup)  src/operators.jl:1569-1619 -> (; uavg, eps, eta, lam, Relam, L, tau, Re_int)"""
    _sync_stream(setup)
    out = (C.c_double * 8)()
    setup.ctx.call("insb200_scale_numbers", _ptr(_check(u, setup, True)), out)
    T = setup.grid.T
    names = ("uavg", "eps", "eta", "lam", "Relam", "L", "tau", "Re_int")
    return SimpleNamespace(**{k: T(v) for k, v in zip(names, out)})


def total_kinetic_energy(u, setup):
    """total_kinetic_energy(u, setup)  src/operators.jl:1551-1556"""
    _sync_stream(setup)
    out = C.c_double()
    setup.ctx.call("insb200_kinetic_energy", _ptr(_check(u, setup, True)), C.byref(out))
    return out.value


def get_cfl_timestep_(buf, u, setup):
    """get_cfl_timestep!(buf, u, setup)  src/solver.jl:101-125 (`buf` is unused: one fused reduction)."""
    _sync_stream(setup)
    out = C.c_double()
    setup.ctx.call("insb200_cfl_dt", _ptr(_check(u, setup, True)), C.byref(out))
    return setup.grid.T(out.value)


def maxabs(u, setup):
    """maximum(abs, u)  src/processors.jl:67"""
    _sync_stream(setup)
    out = C.c_double()
    setup.ctx.call("insb200_maxabs", _ptr(_check(u, setup, True)), C.byref(out))
    return out.value


# --------------------------------------------------------------------------------------------
# time stepping  (src/time_steppers)
# --------------------------------------------------------------------------------------------
class ExplicitRungeKuttaMethod:
    """ExplicitRungeKuttaMethod{T}  src/time_steppers/methods.jl:184-190 (A, c already shifted)."""

    def __init__(self, A, b, c, r, p_add_solve=True):
        self.A, self.b, self.c, self.r, self.p_add_solve = A, b, c, r, p_add_solve


def runge_kutta_method(A, b, c, r, T=np.float64, **kw):
    """runge_kutta_method  src/time_steppers/methods.jl:222-240: explicit tableaux are shifted,
    A = [A[2:end,:]; b'], c = [c[2:end]; 1]."""
    A = np.array(A, dtype=np.float64)
    b = np.array(b, dtype=np.float64)
    c = np.array(c, dtype=np.float64)
    s = A.shape[0]
    if not (A.shape == (s, s) and len(b) == s and len(c) == s):
        raise ValueError("A, b, and c must have the same sizes")
    if not np.allclose(np.triu(A), 0):
        raise NotImplementedError("implicit Runge-Kutta methods are outside the B200 hot path")
    A = np.vstack((A[1:, :], b[None, :])).astype(T)
    c = np.concatenate((c[1:], [1.0])).astype(T)
    return ExplicitRungeKuttaMethod(A, b.astype(T), c, T(r), **kw)


def _rowsum(A):
    return np.sum(np.array(A, dtype=np.float64), axis=1)


_S6 = float(np.sqrt(6.0))
# name -> (A, b, c or None (= row sums of A), r): the explicit tableaux of src/time_steppers/RKMethods.jl
# (file:line of each constructor in the comments).  Fractions are written as Python floats (Float64).
_EXPLICIT_TABLEAUX = {
    "FE11": ([[0]], [1], [0], 1),                                                                          # :45-51
    "SSP22": ([[0, 0], [1, 0]], [1 / 2, 1 / 2], None, 1),                                                  # :54-60
    "SSP42": ([[0, 0, 0, 0], [1 / 3, 0, 0, 0], [1 / 3, 1 / 3, 0, 0], [1 / 3, 1 / 3, 1 / 3, 0]], [1 / 4] * 4, None, 3),  # :63-70
    "SSP33": ([[0, 0, 0], [1, 0, 0], [1 / 4, 1 / 4, 0]], [1 / 6, 1 / 6, 2 / 3], None, 1),                  # :73-79
    "SSP43": ([[0, 0, 0, 0], [1 / 2, 0, 0, 0], [1 / 2, 1 / 2, 0, 0], [1 / 6, 1 / 6, 1 / 6, 0]], [1 / 6, 1 / 6, 1 / 6, 1 / 2], None, 2),  # :82-88
    "Wray3": ([[0, 0, 0], [8 / 15, 0, 0], [8 / 15 - 17 / 60, 5 / 12, 0]], [8 / 15 - 17 / 60, 0, 3 / 4],
              [0, 8 / 15, 8 / 15 - 17 / 60 + 5 / 12], 0),                                                  # :137-146
    "RK56": ([[0, 0, 0, 0, 0, 0], [1 / 4, 0, 0, 0, 0, 0], [1 / 8, 1 / 8, 0, 0, 0, 0], [0, 0, 1 / 2, 0, 0, 0],
              [3 / 16, -3 / 8, 3 / 8, 9 / 16, 0, 0], [-3 / 7, 8 / 7, 6 / 7, -12 / 7, 8 / 7, 0]],
             [7 / 90, 0, 16 / 45, 2 / 15, 16 / 45, 7 / 90], [0, 1 / 4, 1 / 4, 1 / 2, 3 / 4, 1], 0),         # :149-162
    "DOPRI6": ([[0, 0, 0, 0, 0, 0], [1 / 5, 0, 0, 0, 0, 0], [3 / 40, 9 / 40, 0, 0, 0, 0], [44 / 45, -56 / 15, 32 / 9, 0, 0, 0],
                [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729, 0, 0],
                [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656, 0]],
               [35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84], None, 0),                      # :165-178
    "HEM3": ([[0, 0, 0], [1 / 3, 0, 0], [-1, 2, 0]], [0, 3 / 4, 1 / 4], None, 0),                           # :235-241
    "HEM3BS": ([[0, 0, 0], [1 / 2, 0, 0], [-1, 2, 0]], [1 / 6, 2 / 3, 1 / 6], None, 0),                     # :244-250
    "HEM5": ([[0, 0, 0, 0, 0], [3 / 10, 0, 0, 0, 0], [(1 + _S6) / 30, (11 - 4 * _S6) / 30, 0, 0, 0],
              [(-79 - 31 * _S6) / 150, (-1 - 4 * _S6) / 30, (24 + 11 * _S6) / 25, 0, 0],
              [(14 + 5 * _S6) / 6, (-8 + 7 * _S6) / 6, (-9 - 7 * _S6) / 4, (9 - _S6) / 4, 0]],
             [0, 0, (16 - _S6) / 36, (16 + _S6) / 36, 1 / 9], None, 0),                                    # :253-265
    "Mid22": ([[0, 0], [1 / 2, 0]], [0, 1], [0, 1 / 2], 1 / 2),                                            # :461-467
    "MTE22": ([[0, 0], [2 / 3, 0]], [1 / 4, 3 / 4], [0, 2 / 3], 1 / 2),                                    # :470-476
    "Heun33": ([[0, 0, 0], [1 / 3, 0, 0], [0, 2 / 3, 0]], [1 / 4, 0, 3 / 4], None, 0),                      # :488-494
    "RK33C2": ([[0, 0, 0], [2 / 3, 0, 0], [1 / 3, 1 / 3, 0]], [1 / 4, 0, 3 / 4], [0, 2 / 3, 2 / 3], 0),     # :497-503
    "RK33P2": ([[0, 0, 0], [1 / 3, 0, 0], [-1, 2, 0]], [0, 3 / 4, 1 / 4], [0, 1 / 3, 1], 0),                # :506-512
    "RK44": ([[0, 0, 0, 0], [1 / 2, 0, 0, 0], [0, 1 / 2, 0, 0], [0, 0, 1, 0]], [1 / 6, 1 / 3, 1 / 3, 1 / 6], None, 0),  # :515-521
    "RK44C2": ([[0, 0, 0, 0], [1 / 4, 0, 0, 0], [0, 1 / 2, 0, 0], [1, -2, 2, 0]], [1 / 6, 0, 2 / 3, 1 / 6], [0, 1 / 4, 1 / 2, 1], 0),  # :524-530
    "RK44C23": ([[0, 0, 0, 0], [1 / 2, 0, 0, 0], [1 / 4, 1 / 4, 0, 0], [0, -1, 2, 0]], [1 / 6, 0, 2 / 3, 1 / 6], [0, 1 / 2, 1 / 2, 1], 0),  # :533-539
    "RK44P2": ([[0, 0, 0, 0], [1, 0, 0, 0], [3 / 8, 1 / 8, 0, 0], [-1 / 8, -3 / 8, 3 / 2, 0]], [1 / 6, -1 / 18, 2 / 3, 2 / 9], [0, 1, 1 / 2, 1], 0),  # :542-548
    "NSSP21": ([[0, 0], [3 / 4, 0]], [0, 1], [0, 3 / 4], 0),                                               # :589-598
    "NSSP32": ([[0, 0, 0], [1 / 3, 0, 0], [0, 1, 0]], [1 / 2, 0, 1 / 2], [0, 1 / 3, 1], 0),                 # :601-611
    "NSSP33": ([[0, 0, 0], [-4 / 9, 0, 0], [7 / 6, -1 / 2, 0]], [1 / 4, 0, 3 / 4], [0, -4 / 9, 2 / 3], 0),  # :614-624
    "NSSP53": ([[0, 0, 0, 0, 0], [1 / 7, 0, 0, 0, 0], [0, 3 / 16, 0, 0, 0], [0, 0, 1 / 3, 0, 0], [0, 0, 0, 2 / 3, 0]],
               [1 / 4, 0, 0, 0, 3 / 4], [0, 1 / 7, 3 / 16, 1 / 3, 2 / 3], 0),                               # :627-641
}


class _RKMethods:
    """RKMethods.<name>(T=...) for every EXPLICIT tableau of src/time_steppers/RKMethods.jl:45-641 (the
    implicit, DSRK and the matrix-generated SSP104 / rSSPs families are outside the explicit stepper);
    any other explicit tableau goes through `runge_kutta_method(A, b, c, r)`."""

    names = tuple(_EXPLICIT_TABLEAUX)

    def __getattr__(self, name):
        if name not in _EXPLICIT_TABLEAUX:
            raise AttributeError(f"RKMethods has no explicit tableau {name!r}")
        A, b, c, r = _EXPLICIT_TABLEAUX[name]

        def make(**kw):
            return runge_kutta_method(A, b, _rowsum(A) if c is None else c, r, **kw)

        make.__name__ = name
        return make


RKMethods = _RKMethods()


class LMWray3:
    """LMWray3{T}  src/time_steppers/methods.jl (low-storage Wray RK3) with the step of
    step_lmwray3.jl:4-104: per stage  k = f(u);  u = project(xstart + dt*a_i*k);  xstart += dt*b_i*k.
    Unrolled, stage i is  u_i = x0 + dt*(b_1 k_1 + .. + b_{i-1} k_{i-1} + a_i k_i)  added in that order, i.e.
    the explicit stage combination (step_explicit_runge_kutta.jl:35-38) with the shifted tableau below -- same
    operations in the same order, so the library runs it through `insb200_rk_step`; it keeps u, xstart and the
    k_j with a non-zero b_j (k_1 only) instead of six vector fields."""

    a = (8 / 15, 5 / 12, 3 / 4)
    b = (1 / 4, 0.0)
    c_stage = (0.0, 8 / 15, 2 / 3)

    def __init__(self, T=np.float64):
        a, b = self.a, self.b
        self.A = np.array([[a[0], 0, 0], [b[0], a[1], 0], [b[0], b[1], a[2]]], dtype=np.float64).astype(T)
        self.b = self.A[-1].copy()
        # time level reached after each stage's projection; the step ends at t + dt (step_lmwray3.jl:97)
        self.c = np.array([self.c_stage[1], self.c_stage[2], 1.0], dtype=np.float64).astype(T)
        self.r = T(0)


def ode_method_cache(method, setup):
    """ode_method_cache(method, setup)  time_stepper_caches.jl:34-49.  The stage buffers
    (ustart, ku, p) are owned by the library context, so the cache is only a token."""
    return SimpleNamespace(owner="libinsb200", nstage=len(method.b))


def create_stepper(method, setup, psolver=None, u=None, temp=None, t=0.0, n=0):
    """create_stepper(method; setup, psolver, u, temp, t, n=0)  step_explicit_runge_kutta.jl:1-2"""
    if temp is not None:
        _need_temperature(setup)
    return SimpleNamespace(setup=setup, psolver=psolver, u=u, temp=temp, t=setup.grid.T(t), n=n)


def _tableau_args(method):
    s = len(method.b)
    A = np.ascontiguousarray(np.asarray(method.A, dtype=np.float64)).ravel()
    c = np.ascontiguousarray(np.asarray(method.c, dtype=np.float64))
    return s, A.ctypes.data_as(C.POINTER(C.c_double)), c.ctypes.data_as(C.POINTER(C.c_double)), (A, c)


def _theta(setup, theta):
    m = setup.closure_model
    if m is None:
        return float("nan")
    if not getattr(m, "is_smagorinsky", False):
        raise NotImplementedError("only smagorinsky_closure(setup) is fused into the B200 stepper")
    if theta is None:
        raise ValueError("closure_model is set: pass theta")
    return float(theta)


def timestep_(method, stepper, dt, theta=None, cache=None, p=None):
    """timestep!(method, stepper, dt; theta, cache)  step_explicit_runge_kutta.jl:4-59.
    `p` (optional scalar field) receives the last stage's multiplier."""
    setup = stepper.setup
    s, A, c, keep = _tableau_args(method)
    _sync_stream(setup)
    if stepper.temp is None:
        setup.ctx.call("insb200_rk_step", _ptr(_check(stepper.u, setup, True)), float(stepper.t), float(dt), s, A, c,
                       _theta(setup, theta), _ptr(p) if p is not None else None)
    else:
        setup.ctx.call("insb200_rk_step_temp", _ptr(_check(stepper.u, setup, True)), _ptr(_check(stepper.temp, setup, False)),
                       float(stepper.t), float(dt), s, A, c, _theta(setup, theta), _ptr(p) if p is not None else None)
    T = setup.grid.T
    t = T(stepper.t + T(method.c[-1]) * T(dt))
    return create_stepper(method, setup, stepper.psolver, stepper.u, stepper.temp, t, stepper.n + 1)


def _rk_steps(method, setup, u, t, dt, nsteps, theta, ke_hist=False, temp=None):
    s, A, c, keep = _tableau_args(method)
    tt = C.c_double(float(t))
    if temp is not None:
        _sync_stream(setup)
        setup.ctx.call("insb200_rk_steps_temp", _ptr(_check(u, setup, True)), _ptr(_check(temp, setup, False)), C.byref(tt),
                       float(dt), int(nsteps), s, A, c, _theta(setup, theta))
        return setup.grid.T(tt.value), None
    ke = (C.c_double * (nsteps + 1))() if ke_hist else None
    _sync_stream(setup)
    setup.ctx.call("insb200_rk_steps", _ptr(_check(u, setup, True)), C.byref(tt), float(dt), int(nsteps), s, A, c,
                   _theta(setup, theta), ke)
    return setup.grid.T(tt.value), (np.array(ke[:]) if ke_hist else None)


def get_state(stepper):
    """get_state(stepper)  src/solver.jl:95-98"""
    return SimpleNamespace(u=stepper.u, temp=stepper.temp, t=stepper.t, n=stepper.n)


def solve_unsteady(setup, tlims, ustart, tempstart=None, method=None, psolver=None, dt=None, dt_min=None,
                   cfl=0.9, n_adapt_dt=1, docopy=True, processors=None, theta=None, cache=None):
    """solve_unsteady(; setup, tlims, ustart, method, psolver, dt, dt_min, cfl, n_adapt_dt, docopy,
    processors, theta, cache)  src/solver.jl:18-92.  Returns `(state, outputs)`.

    `processors`: dict name -> processor (see processors.py).  With a fixed dt the steps between
    two processor firings (their `nupdate`) run inside one `insb200_rk_steps` call, i.e. without
    touching the reference layout or the host."""
    T = setup.grid.T
    method = method or RKMethods.RK44(T=T)
    psolver = psolver or default_psolver(setup)
    u = ustart.clone() if docopy else ustart
    temp = None if tempstart is None else (tempstart.clone() if docopy else tempstart)
    tstart, tend = T(tlims[0]), T(tlims[1])
    stepper = create_stepper(method, setup, psolver, u, temp, tstart)
    processors = processors or {}
    state = Observable(get_state(stepper))
    initialized = {k: v.initialize(state) for k, v in processors.items()}
    nups = [getattr(v, "nupdate", 1) for v in processors.values()]
    if dt is None:
        cfl = T(cfl)
        d = None
        while stepper.t < tend:
            if stepper.n % n_adapt_dt == 0:
                d = T(cfl * get_cfl_timestep_(None, stepper.u, setup))
                d = d if dt_min is None else max(d, T(dt_min))
            d = min(d, T(tend - stepper.t))
            stepper = timestep_(method, stepper, d, theta=theta, cache=cache)
            state.set(get_state(stepper))
    else:
        nstep = int(round(float(tend - tstart) / float(dt)))
        d = T((tend - tstart) / T(nstep))
        chunk = int(np.gcd.reduce(nups)) if nups else nstep
        chunk = max(1, min(chunk, nstep)) if nstep > 0 else 1
        it = 0
        while it < nstep:
            m = min(chunk, nstep - it)
            t, _ = _rk_steps(method, setup, stepper.u, stepper.t, d, m, theta, temp=stepper.temp)
            it += m
            stepper = create_stepper(method, setup, psolver, stepper.u, stepper.temp, t, stepper.n + m)
            state.set(get_state(stepper))
    outputs = {k: processors[k].finalize(initialized[k], state) for k in processors}
    return SimpleNamespace(u=stepper.u, temp=stepper.temp, t=stepper.t, n=stepper.n), outputs


def solve_unsteady_host(setup, tlims, ustart_host, method=None, dt=None, cfl=0.9, n_adapt_dt=1, theta=None,
                        max_steps=0, ke_hist=0):
    """solve_unsteady with HOST buffers through the single C-ABI call
    `insb200_solve_unsteady_host`: `ustart_host` is a Julia-shaped NumPy array or a pinned CPU
    tensor in device-field shape; it is advanced IN PLACE (H2D, all steps on the device, D2H).
    Returns (nsteps, t_final, ke_history or None)."""
    T = setup.grid.T
    method = method or RKMethods.RK44(T=T)
    s, A, c, keep = _tableau_args(method)
    if isinstance(ustart_host, torch.Tensor):
        if ustart_host.is_cuda or not ustart_host.is_contiguous():
            raise TypeError("host entry point takes contiguous CPU memory")
        ptr = C.c_void_p(ustart_host.data_ptr())
    else:
        if not (ustart_host.flags.f_contiguous and ustart_host.dtype == T):
            raise TypeError("host array must be Fortran-contiguous with the setup's dtype")
        ptr = C.c_void_p(ustart_host.ctypes.data)
    ke = (C.c_double * ke_hist)() if ke_hist else None
    ns, tout = C.c_int(), C.c_double()
    _sync_stream(setup)
    setup.ctx.call("insb200_solve_unsteady_host", ptr, float(tlims[0]), float(tlims[1]), -1.0 if dt is None else float(dt),
                   float(cfl), int(n_adapt_dt), int(max_steps), s, A, c, _theta(setup, theta), ke, int(ke_hist),
                   C.byref(ns), C.byref(tout))
    return ns.value, T(tout.value), (np.array(ke[: min(ke_hist, ns.value + 1)]) if ke_hist else None)


class Observable:
    """Minimal stand-in for Observables.jl (`Observable(state)`, `on(f, obs)`, `obs[] = x`),
    the hook processors listen on (src/processors.jl:39-40, src/solver.jl:49-50,68,78)."""

    def __init__(self, value):
        self.value = value
        self.listeners = []

    def on(self, f):
        self.listeners.append(f)
        return f

    def set(self, value):
        self.value = value
        for f in self.listeners:
            f(value)

    def get(self):
        return self.value


# --------------------------------------------------------------------------------------------
# slab helpers (host side of the multi-GPU decomposition: SURVEY 8e)
# --------------------------------------------------------------------------------------------
def decomposition(n_first, n_last, rank, nranks):
    """(z0, nzl, kx0, kxl) of `rank`: its planes of the last axis and its kx range of the transposed spectrum of the
    distributed pressure solve -- `insb200_decomposition`, the host arithmetic `insb200_create` itself uses."""
    lib = _lib.load()
    z0, nzl, kx0, kxl = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    if lib.insb200_decomposition(int(n_first), int(n_last), int(rank), int(nranks), C.byref(z0), C.byref(nzl), C.byref(kx0), C.byref(kxl)):
        raise ValueError("the last axis must be divisible by the number of ranks")
    return z0.value, nzl.value, kx0.value, kxl.value


def slab_bounds(n_last, rank, nranks):
    """[first, last) interior planes of the last axis owned by `rank`."""
    z0, nzl, _, _ = decomposition(2, n_last, rank, nranks)
    return z0, z0 + nzl


def local_slab(a_global, D, rank, nranks):
    """Cut the local ghosted slab (reference layout) of `rank` out of a GLOBAL ghosted Julia-shaped
    array: interior planes [lo, hi) of axis D-1 plus one ghost plane each side."""
    n_last = a_global.shape[D - 1] - 2
    lo, hi = slab_bounds(n_last, rank, nranks)
    idx = [slice(None)] * a_global.ndim
    idx[D - 1] = slice(lo, hi + 2)
    return np.asfortranarray(a_global[tuple(idx)])


def gather_slabs(slabs, D):
    """Inverse of local_slab for the interior planes; ghosts of the last axis are rebuilt
    periodically.  `slabs`: list of local arrays ordered by rank."""
    parts = []
    for s in slabs:
        idx = [slice(None)] * s.ndim
        idx[D - 1] = slice(1, s.shape[D - 1] - 1)
        parts.append(s[tuple(idx)])
    core = np.concatenate(parts, axis=D - 1)
    first = [slice(None)] * core.ndim
    last = [slice(None)] * core.ndim
    first[D - 1] = slice(0, 1)
    last[D - 1] = slice(core.shape[D - 1] - 1, core.shape[D - 1])
    return np.asfortranarray(np.concatenate((core[tuple(last)], core, core[tuple(first)]), axis=D - 1))
