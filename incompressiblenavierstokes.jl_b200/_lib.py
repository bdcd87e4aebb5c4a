"""ctypes binding of libinsb200.so -- one prototype per entry point of include/insb200.h.

The library is REQUIRED: there is no Python / CPU fallback for any operator.  Importing this
module without the built library raises, and creating a context without a CUDA device fails in
`insb200_create` with "no CUDA device available".
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libinsb200.so")

c_void_p, c_int, c_double, c_ll = C.c_void_p, C.c_int, C.c_double, C.c_longlong
P = C.POINTER

# insb200_force_fn: int (*)(void *user, double t, const void **field)
FORCE_FN = C.CFUNCTYPE(c_int, c_void_p, c_double, P(c_void_p))

# name -> (restype, argtypes); mirrors include/insb200.h declaration by declaration
PROTOTYPES = {
    "insb200_version": (c_int, []),
    "insb200_comm_unique_id": (c_int, [c_void_p]),
    "insb200_create": (c_int, [P(c_void_p), c_int, P(c_int), c_int, P(c_double), c_double, c_int, c_int, c_int, c_void_p]),
    "insb200_destroy": (c_int, [c_void_p]),
    "insb200_last_error": (C.c_char_p, [c_void_p]),
    "insb200_set_stream": (c_int, [c_void_p, c_void_p]),
    "insb200_synchronize": (c_int, [c_void_p]),
    "insb200_local_geometry": (c_int, [c_void_p, P(c_int), P(c_int)]),
    "insb200_decomposition": (c_int, [c_int, c_int, c_int, c_int, P(c_int), P(c_int), P(c_int), P(c_int)]),
    "insb200_apply_bc_u": (c_int, [c_void_p, c_void_p]),
    "insb200_apply_bc_p": (c_int, [c_void_p, c_void_p]),
    "insb200_set_bodyforce": (c_int, [c_void_p, c_void_p]),
    "insb200_set_bodyforce_callback": (c_int, [c_void_p, FORCE_FN, c_void_p]),
    "insb200_momentum": (c_int, [c_void_p, c_void_p, c_void_p, c_double]),
    "insb200_convectiondiffusion": (c_int, [c_void_p, c_void_p, c_void_p, c_int]),
    "insb200_smagorinsky": (c_int, [c_void_p, c_void_p, c_void_p, c_double]),
    "insb200_divergence": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_scalewithvolume": (c_int, [c_void_p, c_void_p]),
    "insb200_pressuregradient": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_applypressure": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_psolve_spectral": (c_int, [c_void_p, c_void_p]),
    "insb200_project": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_pressure": (c_int, [c_void_p, c_void_p, c_void_p, c_double]),
    "insb200_cfl_dt": (c_int, [c_void_p, c_void_p, P(c_double)]),
    "insb200_kinetic_energy": (c_int, [c_void_p, c_void_p, P(c_double)]),
    "insb200_maxabs": (c_int, [c_void_p, c_void_p, P(c_double)]),
    "insb200_reduce_noalloc": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_rk_step": (c_int, [c_void_p, c_void_p, c_double, c_double, c_int, P(c_double), P(c_double), c_double, c_void_p]),
    "insb200_rk_steps": (c_int, [c_void_p, c_void_p, P(c_double), c_double, c_int, c_int, P(c_double), P(c_double), c_double, P(c_double)]),
    "insb200_solve_unsteady_host": (c_int, [c_void_p, c_void_p, c_double, c_double, c_double, c_double, c_int, c_int, c_int,
                                            P(c_double), P(c_double), c_double, P(c_double), c_int, P(c_int), P(c_double)]),
    "insb200_model_bytes_per_cell_step": (c_double, [c_void_p, c_int]),
    "insb200_launch_count": (c_ll, [c_void_p]),
    "insb200_fft_exec_count": (c_ll, [c_void_p]),
    "insb200_set_profiling": (c_int, [c_void_p, c_int]),
    "insb200_timers": (c_int, [c_void_p, P(C.c_char_p), P(c_double), P(c_ll), c_int, P(c_int)]),
    "insb200_reset_timers": (c_int, [c_void_p]),
    "insb200_set_variant": (c_int, [c_void_p, c_int]),
    "insb200_set_param": (c_int, [c_void_p, C.c_char_p, c_double]),
    "insb200_vorticity": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_interpolate_u_p": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_interpolate_w_p": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_velocitynorm": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_workspace_bytes": (c_ll, [c_void_p]),
    "insb200_graph_replays": (c_ll, [c_void_p]),
    "insb200_set_temperature": (c_int, [c_void_p, c_double, c_double, c_double, c_double, c_double, c_int, c_int]),
    "insb200_apply_bc_temp": (c_int, [c_void_p, c_void_p]),
    "insb200_convection_diffusion_temp": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "insb200_dissipation": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "insb200_gravity": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_momentum_temp": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_double]),
    "insb200_rk_step_temp": (c_int, [c_void_p, c_void_p, c_void_p, c_double, c_double, c_int, P(c_double), P(c_double), c_double, c_void_p]),
    "insb200_rk_steps_temp": (c_int, [c_void_p, c_void_p, c_void_p, P(c_double), c_double, c_int, c_int, P(c_double), P(c_double), c_double]),
    "insb200_right_hand_side": (c_int, [c_void_p, c_void_p, c_void_p, c_double]),
    "insb200_qfield": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_dissipation_from_strain": (c_int, [c_void_p, c_void_p, c_void_p]),
    "insb200_scale_numbers": (c_int, [c_void_p, c_void_p, P(c_double)]),
}

_lib = None


def load():
    """dlopen libinsb200.so (RTLD_GLOBAL not needed) and attach prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python incompressiblenavierstokes.jl_b200/build.py` "
            "(nvcc, sm_100a).  There is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class INSB200Error(RuntimeError):
    pass


def check(lib, ctx, rc):
    if rc != 0:
        msg = lib.insb200_last_error(ctx)
        raise INSB200Error(msg.decode() if msg else f"libinsb200 call failed with status {rc}")
