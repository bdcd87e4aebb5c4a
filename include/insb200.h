/* insb200.h -- C ABI of libinsb200.so: the B200-native (sm_100a) time-stepping core that sits
 * behind the operator API of IncompressibleNavierStokes.jl (reference, Julia, v2.2.0).
 *
 * The reference has no FFI today: its operators are Julia generic functions `op!(out, in..., setup)`.
 * Every entry point below replaces the body of one of them for the periodic, uniform-grid hot path
 * (src/utils.jl:1-13 defines that domain); the reference function it replaces is cited as
 * `file:line` (relative to the reference tree).  INTEGRATION.md shows the `ccall` stubs.
 *
 * Conventions
 *  - every function returns 0 on success, non-zero on failure; the message is available from
 *    insb200_last_error(ctx) (ctx == NULL: message of the last failed insb200_create).
 *  - "device field" arguments are raw CUDA device pointers to arrays in the REFERENCE LAYOUT:
 *    column-major, one ghost volume each side: scalar (N1,..,ND), vector (N1,..,ND,D) with
 *    N_a = nlocal_a + 2 (src/grid.jl:121, src/initializers.jl:1-6).  The caller owns them.
 *  - dtype: 0 = Float32, 1 = Float64.  All arithmetic is done in that type.
 *  - calls are asynchronous on the stream given by insb200_set_stream (default: a stream owned
 *    by the ctx), except the reductions and *_host calls, which synchronise before returning.
 *  - one ctx per GPU / per process rank.  With nranks > 1 the domain is slab-decomposed along the
 *    LAST axis (z in 3-D, y in 2-D): rank r owns n_last/nranks consecutive planes, device fields
 *    are the local ghosted slabs and the ghost planes of the last axis are filled over NCCL.
 *  - a ctx is not thread-safe; distinct ctx are independent.
 */
#ifndef INSB200_H
#define INSB200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct insb200_ctx insb200_ctx;

#define INSB200_F32 0
#define INSB200_F64 1
#define INSB200_NCCL_ID_BYTES 128
#define INSB200_MAX_STAGES 16

/* ---- life cycle ---------------------------------------------------------------------------- */

/* Library version (major*10000 + minor*100 + patch). */
int insb200_version(void);

/* Fill a 128-byte NCCL unique id (call on rank 0, broadcast it out of band, pass to create). */
int insb200_comm_unique_id(void *id128);

/* Replaces Setup(...)'s device side + psolver_spectral(setup) construction + ode_method_cache:
 * src/setup.jl:2-46, src/pressure.jl:289-316, src/time_steppers/time_stepper_caches.jl:34-49.
 *   D        2 or 3
 *   n[D]     GLOBAL number of cells per axis (all even, utils.jl:10-12); n[D-1] % nranks == 0
 *   dtype    INSB200_F32 / INSB200_F64
 *   h[D]     uniform spacing per axis (Delta[a][1] of the reference grid)
 *   Re       Reynolds number
 *   device   CUDA device ordinal
 *   rank, nranks, nccl_id   slab decomposition (nccl_id may be NULL when nranks == 1) */
int insb200_create(insb200_ctx **ctx, int D, const int *n, int dtype, const double *h, double Re,
                   int device, int rank, int nranks, const void *nccl_id);
int insb200_destroy(insb200_ctx *ctx);
const char *insb200_last_error(const insb200_ctx *ctx);
/* All later calls enqueue on `cuda_stream` (a cudaStream_t; NULL = the CUDA legacy default stream).
 * Until the first call a non-blocking stream owned by the ctx is used. */
int insb200_set_stream(insb200_ctx *ctx, void *cuda_stream);
int insb200_synchronize(insb200_ctx *ctx);

/* Local slab geometry: nloc[D] interior cells of this rank, offset of its first plane on the last axis. */
int insb200_local_geometry(const insb200_ctx *ctx, int *nloc, int *last_axis_offset);

/* The slab decomposition itself, as pure host arithmetic (no device, no ctx; insb200_create uses the same function):
 * rank owns the planes [z0, z0 + nzl) of the last axis (n_last cells) and, in the transposed spectrum of the
 * distributed pressure solve, the wavenumbers kx in [kx0, kx0 + kxl) of the halved first axis (n_first / 2 + 1 values:
 * n_first / 2 + 1 over nranks columns per rank, the last rank also takes the remainder).  The reference is
 * single-device; this is the contract the multi-GPU tests pin.  Returns 0, or 1 on bad arguments. */
int insb200_decomposition(int n_first, int n_last, int rank, int nranks, int *z0, int *nzl, int *kx0, int *kxl);

/* ---- operators on device fields (reference layout) ----------------------------------------- */

/* apply_bc_u!(u, t, setup), PeriodicBC: src/boundary_conditions.jl:159-167, 276-288 */
int insb200_apply_bc_u(insb200_ctx *ctx, void *u);
/* apply_bc_p!(p, t, setup), PeriodicBC: src/boundary_conditions.jl:197-206, 306-318 */
int insb200_apply_bc_p(insb200_ctx *ctx, void *p);

/* Steady body force field stored by Setup (src/setup.jl:25-32), consumed by applybodyforce!
 * (src/operators.jl:861-879).  `f` = device vector field in reference layout, copied; NULL clears. */
int insb200_set_bodyforce(insb200_ctx *ctx, const void *f);

/* Unsteady body force, the `issteadybodyforce = false` branch of applybodyforce! (src/operators.jl:870-877:
 * `Fα .+= bodyforce.(α, xin..., t)` at every call).  The library cannot evaluate the caller's function, so the caller
 * does -- with the reference's own broadcast -- whenever the path needs the force at a new time: `fn(user, t, &field)`
 * must set `field` to a DEVICE vector field in the reference layout that holds bodyforce.(α, xu[α]..., t) over the whole
 * array (ghosts included, as the broadcast of the reference), valid until the next invocation, with the work that fills
 * it either complete or enqueued on the context's stream; return 0, anything else aborts the calling entry point.
 * It is invoked once per RK stage at the stage's time (momentum!(ku[i], u, temp, t, setup),
 * src/time_steppers/step_explicit_runge_kutta.jl:21,32: tstart for the first stage, tstart + c[i-1] dt afterwards) and
 * by insb200_momentum / insb200_pressure at their `t`.  Steps with a callback are not replayed as CUDA graphs.
 * fn = NULL removes the callback (and the force). */
typedef int (*insb200_force_fn)(void *user, double t, const void **field);
int insb200_set_bodyforce_callback(insb200_ctx *ctx, insb200_force_fn fn, void *user);

/* momentum!(F, u, temp=nothing, t, setup): src/operators.jl:967-976 (fill! + convectiondiffusion!
 * :634-690 + applybodyforce! :861-879).  Ghosts of u must be valid; ghosts of F are 0, or the
 * ghost values of the stored force field (`F .+= bodyforce` covers the whole array). */
int insb200_momentum(insb200_ctx *ctx, void *F, const void *u, double t);

/* convection! / diffusion! / convectiondiffusion! applied to a zero field:
 * src/operators.jl:378-415, 537-573, 634-690.  terms: 1 = convection, 2 = diffusion, 3 = both. */
int insb200_convectiondiffusion(insb200_ctx *ctx, void *F, const void *u, int terms);

/* smagorinsky_closure(setup)(u, theta): src/operators.jl:1294-1305 (smagtensor! :1135-1151,
 * apply_bc_p!(sigma), divoftensor! :1166-1238).  Writes s; ghosts of s are set to 0. */
int insb200_smagorinsky(insb200_ctx *ctx, void *s, const void *u, double theta);

/* divergence!(div, u, setup): src/operators.jl:106-125 (interior written, ghosts untouched) */
int insb200_divergence(insb200_ctx *ctx, void *div, const void *u);
/* Observer fields (observefield, src/processors.jl:77-198), single rank, reference layout:
 * vorticity!(w, u, setup): src/operators.jl:989-1021 -- w is a scalar field in 2-D, a vector field in 3-D; written on
 *   indices 1..N-1 of every axis (Julia, 1-based), the ghosts of u must be valid (apply_bc_u! first, processors.jl:153);
 * interpolate_u_p!(up, u, setup): src/operators.jl:1311-1326 (Ip written);
 * interpolate_w_p!(wp, w, setup): src/operators.jl:1338-1372 (Ip written; wp scalar in 2-D, vector in 3-D);
 * velocitynorm: `@. upnorm = sqrt(up1^2 + up2^2 [+ up3^2])` over the whole array, src/processors.jl:141-151. */
int insb200_vorticity(insb200_ctx *ctx, void *w, const void *u);
int insb200_interpolate_u_p(insb200_ctx *ctx, void *up, const void *u);
int insb200_interpolate_w_p(insb200_ctx *ctx, void *wp, const void *w);
int insb200_velocitynorm(insb200_ctx *ctx, void *out, const void *up);
/* scalewithvolume!(p, setup): src/operators.jl:81-95 (whole array incl. ghosts) */
int insb200_scalewithvolume(insb200_ctx *ctx, void *p);
/* pressuregradient!(G, p, setup): src/operators.jl:159-178 (interior written) */
int insb200_pressuregradient(insb200_ctx *ctx, void *G, const void *p);
/* applypressure!(u, p, setup): src/operators.jl:214-233 */
int insb200_applypressure(insb200_ctx *ctx, void *u, const void *p);

/* psolver_spectral(setup)(p) == poisson!(psolver, p): src/pressure.jl:22, 318-350.
 * In: volume-scaled right-hand side in p[Ip]; out: solution in p[Ip] (ghosts untouched). */
int insb200_psolve_spectral(insb200_ctx *ctx, void *p);

/* project!(u, setup; psolver, p): src/pressure.jl:69-82.  p (nullable) receives the multiplier
 * with its ghosts filled. Ghosts of u are left as the reference leaves them (not refreshed). */
int insb200_project(insb200_ctx *ctx, void *u, void *p);

/* pressure!(p, u, temp=nothing, t, setup; psolver, F): src/pressure.jl:41-49. */
int insb200_pressure(insb200_ctx *ctx, void *p, const void *u, double t);

/* ---- reductions (synchronise; results on the host) ----------------------------------------- */

/* get_cfl_timestep!(buf, u, setup): src/solver.jl:101-125 (the caller multiplies by cfl) */
int insb200_cfl_dt(insb200_ctx *ctx, const void *u, double *dt);
/* total_kinetic_energy(u, setup) / energy_history_plot: src/operators.jl:1516-1556,
 * ext/IncompressibleNavierStokesMakieExt.jl:281-295 */
int insb200_kinetic_energy(insb200_ctx *ctx, const void *u, double *E);
/* maximum(abs, u) of timelogger: src/processors.jl:67 */
int insb200_maxabs(insb200_ctx *ctx, const void *u, double *m);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The fused reduction pass of the three
 * entries above without any allocation, host copy or synchronisation: out4_dev is a caller-owned DEVICE buffer of four
 * doubles that receives { total kinetic energy, max|u_1|, max|u_2|, max|u_3| } (the inputs of insb200_kinetic_energy,
 * insb200_maxabs and the CFL bound), enqueued on the context's stream.  The per-block partial sums live in a static
 * 2 MiB device array inside the library (held by every GPU that loads it until the process exits) that ALL calls and
 * all contexts on a device share: calls that overlap on the GPU (different streams) race; calls on one stream are safe.
 * Limit: 65536 blocks of 64 x 4 cells, i.e. grids up to 256^3 (3-D) or 4096^2 (2-D); larger grids and nranks > 1 are
 * rejected with an error message.  The first call on a device looks the array's address up once
 * (cudaGetSymbolAddress): make one warm-up call before capturing a graph. */
int insb200_reduce_noalloc(insb200_ctx *ctx, const void *u, double *out4_dev);

/* ---- time stepping -------------------------------------------------------------------------- */

/* timestep!(::ExplicitRungeKuttaMethod, stepper, dt; theta, cache):
 * src/time_steppers/step_explicit_runge_kutta.jl:4-59.
 *   u      device vector field (reference layout), advanced in place, all ghosts valid on return
 *   s,A,c  SHIFTED tableau of runge_kutta_method (src/time_steppers/methods.jl:231-236):
 *          A is s*s row-major, c has s entries (RK44: RKMethods.jl:515-521)
 *   theta  Smagorinsky constant; NaN = no closure model (setup.closure_model === nothing)
 *   p_out  nullable device scalar field receiving the last stage's multiplier p (ghosts filled) */
int insb200_rk_step(insb200_ctx *ctx, void *u, double t, double dt, int s, const double *A,
                    const double *c, double theta, void *p_out);

/* nsteps calls of timestep! with fixed dt (the fixed-step loop of solve_unsteady,
 * src/solver.jl:71-79) without leaving the device.  *t is advanced.  ke_hist (nullable, host,
 * nsteps+1 doubles) receives the total kinetic energy before the first and after every step. */
int insb200_rk_steps(insb200_ctx *ctx, void *u, double *t, double dt, int nsteps, int s,
                     const double *A, const double *c, double theta, double *ke_hist);

/* solve_unsteady(; setup, tlims, ustart, method, dt, cfl, n_adapt_dt, theta) with HOST buffers:
 * src/solver.jl:18-92.  u_host: reference-layout vector field in host memory (pinned or not),
 * copied to the device, advanced from t0 to tend, copied back.  dt <= 0 selects the adaptive
 * loop (:53-69) with `cfl` and `n_adapt_dt`; dt > 0 the fixed loop (:71-79, dt re-derived from
 * round((tend-t0)/dt)).  max_steps > 0 bounds the number of steps.  ke_hist (nullable) holds
 * ke_cap doubles: energy at start and after every step while it fits.  Outputs: number of steps
 * taken and the final time. */
int insb200_solve_unsteady_host(insb200_ctx *ctx, void *u_host, double t0, double tend, double dt,
                                double cfl, int n_adapt_dt, int max_steps, int s, const double *A,
                                const double *c, double theta, double *ke_hist, int ke_cap,
                                int *nsteps_out, double *t_out);

/* ---- temperature equation and the callers next to the path (single rank) ------------- */

/* temperature_equation(; Pr, Ra, Ge, dodissipation, gdir, nondim_type): src/setup.jl:56-86.  The caller forms the
 * coefficients alpha1..alpha4 and gamma = alpha1/alpha3 in the field type (exactly representable as double);
 * gdir is the 1-based gravity axis.  Temperature boundary conditions: periodic only. */
int insb200_set_temperature(insb200_ctx *ctx, double alpha1, double alpha2, double alpha3, double alpha4,
                            double gamma, int dodissipation, int gdir);
/* apply_bc_temp!(temp, t, setup), PeriodicBC: src/boundary_conditions.jl:236-246, 338-339 */
int insb200_apply_bc_temp(insb200_ctx *ctx, void *temp);
/* convection_diffusion_temp!(c, u, temp, setup): src/operators.jl:711-735 (ADDS to c on Ip) */
int insb200_convection_diffusion_temp(insb200_ctx *ctx, void *c, const void *u, const void *temp);
/* dissipation!(diss, diff, u, setup): src/operators.jl:787-808 (diff is overwritten by diffusion!(0, u); ADDS to diss) */
int insb200_dissipation(insb200_ctx *ctx, void *diss, void *diff, const void *u);
/* gravity!(F, temp, setup): src/operators.jl:916-931 (ADDS alpha2 * avg(temp) to F[:, gdir] on Iu[gdir]) */
int insb200_gravity(insb200_ctx *ctx, void *F, const void *temp);
/* momentum!(F, u, temp, t, setup) with temp !== nothing: src/operators.jl:967-976 (temp may be NULL) */
int insb200_momentum_temp(insb200_ctx *ctx, void *F, const void *u, const void *temp, double t);
/* timestep!(::ExplicitRungeKuttaMethod, ...) with the temperature hooks: step_explicit_runge_kutta.jl:4-59
 * (cache: time_stepper_caches.jl:34-49).  u and temp are advanced in place with valid ghosts; other arguments as
 * insb200_rk_step.  First version: one kernel per reference kernel on the reference layout (not fused). */
int insb200_rk_step_temp(insb200_ctx *ctx, void *u, void *temp, double t, double dt, int s, const double *A,
                         const double *c, double theta, void *p_out);
int insb200_rk_steps_temp(insb200_ctx *ctx, void *u, void *temp, double *t, double dt, int nsteps, int s,
                          const double *A, const double *c, double theta);
/* right_hand_side!(dudt, u, params_ref, t) of the SciML glue: src/sciml.jl:37-49 (u is not modified) */
int insb200_right_hand_side(insb200_ctx *ctx, void *dudt, const void *u, double t);
/* Qfield!(Q, u, setup): src/operators.jl:1441-1460 (Ip written) */
int insb200_qfield(insb200_ctx *ctx, void *Q, const void *u);
/* dissipation_from_strain!(eps, u, setup): src/operators.jl:819-836 (Ip written; ghosts of u must be valid) */
int insb200_dissipation_from_strain(insb200_ctx *ctx, void *eps, const void *u);
/* get_scale_numbers(u, setup): src/operators.jl:1569-1619.  out8 (host) = uavg, epsilon, eta, lambda, Re_lambda,
 * L, tau, Re_int; synchronises.  Reductions accumulate in double (the reference sums in T). */
int insb200_scale_numbers(insb200_ctx *ctx, const void *u, double *out8);

/* ---- introspection --------------------------------------------------------------------------- */

/* Algorithmic bytes per cell per RK step of the current configuration (SURVEY.md 8d model:
 * (46 D + 8) words + D words/stage with a stored body force). */
double insb200_model_bytes_per_cell_step(const insb200_ctx *ctx, int nstages);

/* Number of kernels of this library launched so far on this ctx, and of cuFFT executions. */
long long insb200_launch_count(const insb200_ctx *ctx);
long long insb200_fft_exec_count(const insb200_ctx *ctx);

/* Per-kernel device timers (CUDA events on the launching stream).  enable=1 starts recording
 * every launch; insb200_timers synchronises and returns up to `cap` entries: name (static
 * string), accumulated milliseconds, launch count.  Returns the number of entries via *n. */
int insb200_set_profiling(insb200_ctx *ctx, int enable);
int insb200_timers(insb200_ctx *ctx, const char **names, double *ms, long long *counts, int cap, int *n);
int insb200_reset_timers(insb200_ctx *ctx);

/* Kernel variant switch for A/B runs: 0 = reference-structure kernels (one thread per cell,
 * direct global loads), 1 = LDG-tiled shared-memory kernels, 2 = TMA plane-pipeline kernels (3-D;
 * default).  Results are bit-identical across variants. */
int insb200_set_variant(insb200_ctx *ctx, int variant);

/* Tuning knobs for A/B runs.  Unless noted, results are bit-identical for every setting.
 *   stage kernels:  "variant" 0|1|2, "tile_y" 8|16, "zchunk" planes marched per block (0 = automatic),
 *                   "fuse_gradient" 0|1 (projection of stage i applied by the loader of stage i+1),
 *                   "stage_x2" 0|1|2 (packed-FP32 stage kernel: 1 exact, 2 contracted -- 2 differs by round-off),
 *                   2-D: "fuse2d" 0|1 (warp-marching fused stage kernel), "rows2d" 0|1, "rows_chunk" rows per block
 *   closure:        "smag_v" 1|2, "smag_ty" 8|16, "smag_zchunk"
 *   Poisson solve:  "zsolve" 0|1 (fused last-axis kernel instead of cuFFT's z pass; differs from cuFFT by round-off),
 *                   "zcols" 0|4|8|16, "zthreads", "zradix16" 0|1 (Float32 radix-16 passes; round-off),
 *                   "fft_chunk" planes per cufftExec (0 = whole slab), "fft_pipe" / "fft_pipe_min_ranks"
 *   arithmetic:     "exact" 0|1 (1: the kernels built without FMA contraction, bit-identical to the oracle; 0, default: the
 *                   contracted kernels of fast.cu, north_star's tolerance), "fast_gen" 0|1
 *   multi-GPU:      "p2p" 0|1 (exchange over peer memory mapped with CUDA IPC; 0: NCCL send/recv; before the first solve),
 *                   "p2p_halo" 0|1, "fast_barrier" 0|1 (flag words in peer memory | 4-byte ncclAllReduce),
 *                   "zline" -1|0|1 (last axis of the pressure solve without transposes, csrc/zline.cuh: automatic = on when
 *                   the peers are mapped | distributed FFT | on, 3-D also on one rank; differs from the FFT by round-off),
 *                   "zline_chunks" sub-chunks per slab (0 = automatic),
 *                   distributed FFT: "transpose_mode" -1|0|1|2 (automatic | scatter kernels | pitched peer copies | pack +
 *                   contiguous peer copies), "scatter_blocks"; "stage_x2" / "p2p_fused" need -DINSB200_EXPERIMENTAL
 *   "graphs" 1|0|-1 (CUDA-graph replay of whole steps: on, off, automatic for slabs of <= 2^22 cells). */
int insb200_set_param(insb200_ctx *ctx, const char *name, double value);

/* Whole RK steps are replayed as CUDA graphs where that is possible (one rank, fixed dt, a non-default stream, timers
 * off; set_param("graphs", 0) disables it): number of steps that ran as a graph replay so far. */
long long insb200_graph_replays(const insb200_ctx *ctx);

/* Device memory owned by the ctx, in bytes. */
long long insb200_workspace_bytes(const insb200_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* INSB200_H */
