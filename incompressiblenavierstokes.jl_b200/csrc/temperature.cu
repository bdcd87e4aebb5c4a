// temperature.cu -- the callers either side of the velocity hot path, sm_100a:
//   * temperature equation on the periodic path: apply_bc_temp!, convection_diffusion_temp!, dissipation!,
//     gravity!, momentum!(.., temp, ..) and the explicit RK step with its temperature hooks
//     (src/operators.jl:711-735, 787-808, 916-931, 967-976; step_explicit_runge_kutta.jl:4-59),
//   * right_hand_side! of the SciML glue (src/sciml.jl:37-49),
//   * Qfield!, dissipation_from_strain! and get_scale_numbers (src/operators.jl:1441-1460, 819-836, 1569-1619).
// All fields are raw device pointers in the REFERENCE layout (N_a = n_a + 2, component slowest).  This first
// version of the temperature step is UNFUSED -- one kernel per reference kernel, composed from the public
// operators of include/insb200.h on the reference layout -- and single-rank; the velocity-only path keeps its
// fused kernels (tma_stage.cuh).  Arithmetic: the reference's operation order, true divisions (these kernels
// are not on the measured path), no FMA contraction (-fmad=false): bit-identical to the oracle for any h.
#include <algorithm>
#include <type_traits>

#include "ctx.cuh"

namespace insb200 {

// strides of the reference layout; o(i,j,k) = offset of interior cell (i,j,k), 0-based
struct RefGeo {
  int n[3];
  i64 e[3];  // element strides of the three axes (e[2] = 0 in 2-D)
  i64 sc;    // component stride
  i64 org;   // offset of interior cell (0,0,0)
};

static RefGeo ref_geo(const insb200_ctx *c) {
  RefGeo g;
  const i64 N0 = c->nloc[0] + 2, N1 = c->nloc[1] + 2, N2 = c->D == 3 ? c->nloc[2] + 2 : 1;
  for (int a = 0; a < 3; a++) g.n[a] = c->nloc[a];
  g.e[0] = 1; g.e[1] = N0; g.e[2] = c->D == 3 ? N0 * N1 : 0;
  g.sc = N0 * N1 * N2;
  g.org = g.e[0] + g.e[1] + g.e[2];
  return g;
}

template <typename T>
struct H3 {
  T v[3];
};

#define TQ_CELL(g)                                                   \
  const int i = blockIdx.x * blockDim.x + threadIdx.x;               \
  const int j = blockIdx.y * blockDim.y + threadIdx.y;               \
  const int k = blockIdx.z;                                          \
  if (i >= (g).n[0] || j >= (g).n[1]) return;                        \
  const i64 o = (g).org + i + j * (g).e[1] + k * (g).e[2];

// avg(phi, Delta, I, b)  src/operators.jl:59-62 on a uniform grid: (h*phi[I] + h*phi[I+e_b]) / (h + h)
template <typename T>
__device__ __forceinline__ T avg_b(const T *__restrict__ phi, i64 o, i64 eb, T h) {
  return (h * phi[o] + h * phi[o + eb]) / (h + h);
}

// convection_diffusion_temp_kernel!  src/operators.jl:723-735:  c[I] += sum_b (-(uT2-uT1) + a4 (dT2-dT1)) / D[b]
template <typename T, int D>
__global__ void k_convdiff_temp(T *__restrict__ cc, const T *__restrict__ u, const T *__restrict__ temp, RefGeo g, T a4, H3<T> h) {
  TQ_CELL(g)
  T cI = T(0);
#pragma unroll
  for (int b = 0; b < D; b++) {
    const i64 eb = g.e[b];
    const T hb = h.v[b];
    const T dT1 = (temp[o] - temp[o - eb]) / hb;
    const T dT2 = (temp[o + eb] - temp[o]) / hb;
    const T uT1 = u[b * g.sc + o - eb] * avg_b(temp, o - eb, eb, hb);
    const T uT2 = u[b * g.sc + o] * avg_b(temp, o, eb, hb);
    cI += (-(uT2 - uT1) + a4 * (dT2 - dT1)) / hb;
  }
  cc[o] += cI;
}

// interpolate! of dissipation!  src/operators.jl:795-803:  diss[I] += sum_b coef (u[I-e_b,b] diff[I-e_b,b] + u[I,b] diff[I,b]) / 2
// with coef = Re*a1/gamma; the ghosts of diff are 0 (fill!(diff, 0) + diffusion! writes Iu only), as in the reference
template <typename T, int D>
__global__ void k_dissipation(T *__restrict__ diss, const T *__restrict__ diff, const T *__restrict__ u, RefGeo g, T coef) {
  TQ_CELL(g)
  T d = T(0);
#pragma unroll
  for (int b = 0; b < D; b++) {
    const i64 ob = b * g.sc + o;
    d += coef * (u[ob - g.e[b]] * diff[ob - g.e[b]] + u[ob] * diff[ob]) * T(0.5);
  }
  diss[o] += d;
}

// g! of gravity!  src/operators.jl:922-926:  F[I, gdir] += a2 * avg(temp, Delta, I, gdir)
template <typename T>
__global__ void k_gravity(T *__restrict__ F, const T *__restrict__ temp, RefGeo g, int gd, T a2, T hg) {
  TQ_CELL(g)
  F[gd * g.sc + o] += a2 * avg_b(temp, o, g.e[gd], hg);
}

// out = base + sum_j coef[j] * kin[j] over a whole array (step_explicit_runge_kutta.jl:35-38, 40-43); j ascending
template <typename T>
struct CombineArgs {
  const T *kin[INSB200_MAX_STAGES];
  T coef[INSB200_MAX_STAGES];
  int nk;
};
template <typename T>
__global__ void k_combine(T *__restrict__ out, const T *__restrict__ base, CombineArgs<T> a, i64 n) {
  const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  T v = base[t];
  for (int q = 0; q < a.nk; q++) v = v + a.coef[q] * a.kin[q][t];
  out[t] = v;
}

// a .+= b over a whole array (`ku[i] .+= m(u, theta)`, step_explicit_runge_kutta.jl:29)
template <typename T>
__global__ void k_add(T *__restrict__ a, const T *__restrict__ b, i64 n) {
  const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) a[t] = a[t] + b[t];
}

// Q!  src/operators.jl:1446-1456:  q -= (u[I,a]-u[I-e_b,a]) / D[b] * (u[I,b]-u[I-e_a,b]) / D[a] / 2, a outer, b inner
template <typename T, int D>
__global__ void k_qfield(T *__restrict__ Q, const T *__restrict__ u, RefGeo g, H3<T> h) {
  TQ_CELL(g)
  T q = T(0);
#pragma unroll
  for (int a = 0; a < D; a++)
#pragma unroll
    for (int b = 0; b < D; b++)
      q -= (u[a * g.sc + o] - u[a * g.sc + o - g.e[b]]) / h.v[b] * (u[b * g.sc + o] - u[b * g.sc + o - g.e[a]]) / h.v[a] * T(0.5);
  Q[o] = q;
}

// dx(u, I, a, b, D[b], Du[b])  src/operators.jl:1023-1033 with true divisions
template <typename T>
__device__ __forceinline__ T ddx_ref(const T *__restrict__ u, const RefGeo &g, i64 o, int a, int b, T hb) {
  const T *ua = u + a * g.sc;
  const i64 ea = g.e[a], eb = g.e[b];
  if (a == b) return (ua[o] - ua[o - eb]) / hb;
  return ((ua[o + eb] - ua[o]) / hb + (ua[o - ea + eb] - ua[o - ea]) / hb + (ua[o] - ua[o - eb]) / hb +
          (ua[o - ea] - ua[o - ea - eb]) / hb) * T(0.25);
}

// sum(S .* S) with S = (G + G')/2  (column-major order of the SMatrix, left fold; operators.jl:833-835)
template <typename T, int D>
__device__ __forceinline__ T strain_ss(const T *__restrict__ u, const RefGeo &g, i64 o, const H3<T> &h) {
  T G[3][3];
#pragma unroll
  for (int a = 0; a < D; a++)
#pragma unroll
    for (int b = 0; b < D; b++) G[a][b] = ddx_ref(u, g, o, a, b, h.v[b]);
  T ss = T(0);
  bool first = true;
#pragma unroll
  for (int b = 0; b < D; b++)
#pragma unroll
    for (int a = 0; a < D; a++) {
      const T S = (G[a][b] + G[b][a]) * T(0.5);
      const T t = S * S;
      ss = first ? t : ss + t;
      first = false;
    }
  return ss;
}

// dissipation_from_strain_kernel!  src/operators.jl:830-836:  eps[I] = 2 * visc * sum(S .* S)
template <typename T, int D>
__global__ void k_strain_dissipation(T *__restrict__ eps, const T *__restrict__ u, RefGeo g, H3<T> h, T visc) {
  TQ_CELL(g)
  eps[o] = T(2) * visc * strain_ss<T, D>(u, g, o, h);
}

}  // namespace insb200

// =================================================================================================
// get_scale_numbers (src/operators.jl:1569-1619): two fused reductions + one spectrum reduction
// =================================================================================================
namespace insb200 {

__device__ __forceinline__ double wsum(double v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

// block partials of (a) sum_c u[I,c]^2 and (b) eps[I] = 2 visc S:S over the interior; per-cell values in T, sums in double
template <typename T, int D>
__global__ void k_scale_partials(const T *__restrict__ u, RefGeo g, H3<T> h, T visc, double *__restrict__ part) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z;
  double s0 = 0.0, s1 = 0.0;
  if (i < g.n[0] && j < g.n[1]) {
    const i64 o = g.org + i + j * g.e[1] + k * g.e[2];
    T q = T(0);
#pragma unroll
    for (int c = 0; c < D; c++) q += u[c * g.sc + o] * u[c * g.sc + o];
    s0 = (double)q;
    s1 = (double)(T(2) * visc * strain_ss<T, D>(u, g, o, h));
  }
  __shared__ double sh[2][32];
  const int tid = threadIdx.y * blockDim.x + threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int nw = (blockDim.x * blockDim.y + 31) >> 5;
  s0 = wsum(s0); s1 = wsum(s1);
  if (lane == 0) { sh[0][w] = s0; sh[1][w] = s1; }
  __syncthreads();
  if (w == 0) {
    s0 = lane < nw ? sh[0][lane] : 0.0;
    s1 = lane < nw ? sh[1][lane] : 0.0;
    s0 = wsum(s0); s1 = wsum(s1);
    if (lane == 0) {
      const i64 b = ((i64)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
      part[2 * b] = s0;
      part[2 * b + 1] = s1;
    }
  }
}

// copy the interior of one component into a dense (n0, n1, n2) array (the `view(u, Ip, :)` of :1597)
template <typename T>
__global__ void k_strip(T *__restrict__ dst, const T *__restrict__ src, RefGeo g) {
  TQ_CELL(g)
  dst[((i64)k * g.n[1] + j) * g.n[0] + i] = src[o];
}

// block partials of e = |uhat|^2 / (2 prod(Np)^2) / |k| over the first K = n/2 modes of every axis, k = 0 excluded
// (:1599-1615); uhat is the R2C spectrum (n0/2+1 fastest), the first K modes of the full transform are a subset of it
template <typename T, typename C>
__global__ void k_spectrum_partials(const C *__restrict__ uh, int nxh, int n1, int K0, int K1, int K2, T scale, double *__restrict__ part) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z;
  double s = 0.0;
  if (i < K0 && j < K1 && k < K2 && (i | j | k) != 0) {
    const C v = uh[((i64)k * n1 + j) * nxh + i];
    const T e = (v.x * v.x + v.y * v.y) / scale;
    const T kk = sqrt((T)(i * i + j * j + k * k));
    s = (double)(e / kk);
  }
  __shared__ double sh[32];
  const int tid = threadIdx.y * blockDim.x + threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int nw = (blockDim.x * blockDim.y + 31) >> 5;
  s = wsum(s);
  if (lane == 0) sh[w] = s;
  __syncthreads();
  if (w == 0) {
    s = lane < nw ? sh[lane] : 0.0;
    s = wsum(s);
    if (lane == 0) part[((i64)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = s;
  }
}

// out[q] (+)= sum_b part[stride*b + q], one block, fixed order -> deterministic
__global__ void k_sum_partials(const double *__restrict__ part, i64 nb, int stride, int q, double *__restrict__ out, int accumulate) {
  double s = 0.0;
  for (i64 b = threadIdx.x; b < nb; b += blockDim.x) s += part[stride * b + q];
  __shared__ double sh[32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  s = wsum(s);
  if (lane == 0) sh[w] = s;
  __syncthreads();
  if (w == 0) {
    s = lane < nw ? sh[lane] : 0.0;
    s = wsum(s);
    if (lane == 0) *out = accumulate ? *out + s : s;
  }
}

#define TQ_LAUNCHED(c)                 \
  do {                                 \
    (c)->launches++;                   \
    INS_CUDA((c), cudaGetLastError()); \
  } while (0)

static inline dim3 tq_block() { return dim3(64, 4, 1); }
static inline dim3 tq_grid(const insb200_ctx *c) { return dim3((c->nloc[0] + 63) / 64, (c->nloc[1] + 3) / 4, c->nloc[2]); }
static inline i64 ref_elems(const insb200_ctx *c) {
  i64 e = 1;
  for (int a = 0; a < c->D; a++) e *= c->nloc[a] + 2;
  return e;
}
template <typename T>
static H3<T> make_h(const insb200_ctx *c) {
  H3<T> h;
  for (int a = 0; a < 3; a++) h.v[a] = T(c->h[a]);
  return h;
}

static int single_rank(insb200_ctx *c, const char *what) {
  if (c->nranks > 1) INS_FAIL(c, std::string(what) + " runs on one rank only (nranks > 1 not supported yet)");
  return 0;
}

template <typename T, int D>
static int op_scale_numbers(insb200_ctx *c, const void *u_, double *out8) {
  typedef typename std::conditional<sizeof(T) == 4, cufftComplex, cufftDoubleComplex>::type C;
  const T *u = (const T *)u_;
  const RefGeo g = ref_geo(c);
  const H3<T> h = make_h<T>(c);
  const T visc = T(1) / T(c->Re);
  const dim3 grd = tq_grid(c), blk = tq_block();
  const i64 nb = (i64)grd.x * grd.y * grd.z;
  const i64 ncell = (i64)c->nloc[0] * c->nloc[1] * c->nloc[2];
  const int nxh = c->nloc[0] / 2 + 1;
  const i64 nspec = (i64)nxh * c->nloc[1] * c->nloc[2];
  // scratch of this observer: partials (2 nb), 3 results, dense real copy, half spectrum
  double *part = nullptr;
  T *dense = nullptr;
  C *spec = nullptr;
  cufftHandle plan = 0;
  auto cleanup = [&]() {
    if (plan) cufftDestroy(plan);
    cudaFree(part); cudaFree(dense); cudaFree(spec);
  };
#define SN_CUDA(call)                                                                     \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) { cleanup(); INS_FAIL(c, std::string("CUDA: ") + cudaGetErrorString(e_) + " in " #call); } \
  } while (0)
  SN_CUDA(cudaMalloc((void **)&part, (size_t)(2 * nb + 4) * sizeof(double)));
  SN_CUDA(cudaMalloc((void **)&dense, (size_t)ncell * sizeof(T)));
  SN_CUDA(cudaMalloc((void **)&spec, (size_t)nspec * sizeof(C)));
  double *res = part + 2 * nb;
  k_scale_partials<T, D><<<grd, blk, 0, c->stream>>>(u, g, h, visc, part);
  c->launches++;
  k_sum_partials<<<1, 1024, 0, c->stream>>>(part, nb, 2, 0, res + 0, 0);
  k_sum_partials<<<1, 1024, 0, c->stream>>>(part, nb, 2, 1, res + 1, 0);
  c->launches += 2;
  // integral length scale: per component R2C transform of u[Ip, a], shell-free sum of e / |k| over the first K modes
  int dims[3];
  for (int a = 0; a < D; a++) dims[a] = c->nloc[D - 1 - a];
  if (cufftPlanMany(&plan, D, dims, nullptr, 1, 0, nullptr, 1, 0, sizeof(T) == 4 ? CUFFT_R2C : CUFFT_D2Z, 1) != CUFFT_SUCCESS ||
      cufftSetStream(plan, c->stream) != CUFFT_SUCCESS) {
    cleanup();
    INS_FAIL(c, "cuFFT plan of get_scale_numbers failed");
  }
  const int K0 = c->nloc[0] / 2, K1 = c->nloc[1] / 2, K2 = D == 3 ? c->nloc[2] / 2 : 1;
  const T pn = (T)(double)ncell;
  const T scale = T(2) * pn * pn;  // 2 * prod(Np)^2
  dim3 sgrd((K0 + 63) / 64, (K1 + 3) / 4, K2);
  const i64 snb = (i64)sgrd.x * sgrd.y * sgrd.z;  // <= nb
  for (int a = 0; a < D; a++) {
    k_strip<T><<<grd, blk, 0, c->stream>>>(dense, u + a * g.sc, g);
    c->launches++;
    cufftResult r;
    if constexpr (sizeof(T) == 4) r = cufftExecR2C(plan, (cufftReal *)dense, spec);
    else r = cufftExecD2Z(plan, (cufftDoubleReal *)dense, spec);
    if (r != CUFFT_SUCCESS) { cleanup(); INS_FAIL(c, "cuFFT exec of get_scale_numbers failed"); }
    c->fft_execs++;
    k_spectrum_partials<T, C><<<sgrd, blk, 0, c->stream>>>(spec, nxh, c->nloc[1], K0, K1, K2, scale, part);
    k_sum_partials<<<1, 1024, 0, c->stream>>>(part, snb, 1, 0, res + 2, a > 0);
    c->launches += 2;
  }
  double hres[3];
  SN_CUDA(cudaGetLastError());
  SN_CUDA(cudaMemcpyAsync(hres, res, 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SN_CUDA(cudaStreamSynchronize(c->stream));
  cleanup();
#undef SN_CUDA
  // :1577-1618 in T.  `field = @. u^2 * Omega_u` runs over ALL components for every alpha (:1582-1584), so
  // uavg^2 = D * sum_{I,c} u^2 / ncells on the uniform grid (Omega_u cancels)
  const T uavg = std::sqrt(T((double)D * hres[0] / (double)ncell));
  const T eps = T(hres[1] / (double)ncell);
  const T eta = std::pow(visc * visc * visc / eps, T(0.25));
  const T lam = std::sqrt(T(5) * visc / eps) * uavg;
  const T Relam = lam * uavg / std::sqrt(T(3)) / visc;
  const T L = T(3 * M_PI) / T(2) / (uavg * uavg) * T(hres[2]);
  const T tau = L / uavg;
  const T Re_int = L * uavg / visc;
  const T o[8] = {uavg, eps, eta, lam, Relam, L, tau, Re_int};
  for (int q = 0; q < 8; q++) out8[q] = (double)o[q];
  return 0;
}

}  // namespace insb200

// =================================================================================================
// temperature operators and the RK step with temperature (reference layout, single rank)
// =================================================================================================
namespace insb200 {

static int need_temperature(insb200_ctx *c) {
  if (!c->tq.set) INS_FAIL(c, "no temperature equation: call insb200_set_temperature first (setup.temperature === nothing)");
  return single_rank(c, "the temperature equation");
}

template <typename T, int D>
static int op_convdiff_temp(insb200_ctx *c, void *cc, const void *u, const void *temp) {
  k_convdiff_temp<T, D><<<tq_grid(c), tq_block(), 0, c->stream>>>((T *)cc, (const T *)u, (const T *)temp, ref_geo(c), T(c->tq.a4), make_h<T>(c));
  TQ_LAUNCHED(c);
  return 0;
}

template <typename T, int D>
static int op_dissipation(insb200_ctx *c, void *diss, void *diff, const void *u) {
  // fill!(diff, 0); diffusion!(diff, u, setup)  (:792-793) -- the public operator zero-fills and writes Iu only
  INS_TRY(insb200_convectiondiffusion(c, diff, u, 2));
  const T coef = T(c->Re) * T(c->tq.a1) / T(c->tq.gamma);
  k_dissipation<T, D><<<tq_grid(c), tq_block(), 0, c->stream>>>((T *)diss, (const T *)diff, (const T *)u, ref_geo(c), coef);
  TQ_LAUNCHED(c);
  return 0;
}

template <typename T, int D>
static int op_gravity(insb200_ctx *c, void *F, const void *temp) {
  const int gd = c->tq.gdir - 1;
  k_gravity<T><<<tq_grid(c), tq_block(), 0, c->stream>>>((T *)F, (const T *)temp, ref_geo(c), gd, T(c->tq.a2), T(c->h[gd]));
  TQ_LAUNCHED(c);
  return 0;
}

template <typename T, int D>
static int op_qfield(insb200_ctx *c, void *Q, const void *u) {
  k_qfield<T, D><<<tq_grid(c), tq_block(), 0, c->stream>>>((T *)Q, (const T *)u, ref_geo(c), make_h<T>(c));
  TQ_LAUNCHED(c);
  return 0;
}

template <typename T, int D>
static int op_strain_dissipation(insb200_ctx *c, void *eps, const void *u) {
  k_strain_dissipation<T, D><<<tq_grid(c), tq_block(), 0, c->stream>>>((T *)eps, (const T *)u, ref_geo(c), make_h<T>(c), T(1) / T(c->Re));
  TQ_LAUNCHED(c);
  return 0;
}

template <typename T>
static int combine(insb200_ctx *c, T *out, const T *base, const CombineArgs<T> &a, i64 n) {
  k_combine<T><<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(out, base, a, n);
  TQ_LAUNCHED(c);
  return 0;
}

// timestep!(::ExplicitRungeKuttaMethod, ...) with temp !== nothing, statement by statement
// (step_explicit_runge_kutta.jl:4-59).  Cache = ode_method_cache (time_stepper_caches.jl:34-49) in c->tws.
template <typename T, int D>
static int op_rk_steps_temp(insb200_ctx *c, void *u_, void *temp_, double *t, double dt, int nsteps, int s, const double *A,
                            const double *cc, double theta, void *p_out) {
  const bool smag = !std::isnan(theta);
  const i64 N = ref_elems(c);
  const size_t vb = (size_t)N * D * sizeof(T), sb = (size_t)N * sizeof(T);
  // ustart, ku[s], diff, [closure term], tempstart, ktemp[s], p
  const size_t need = vb * (size_t)(s + 2 + (smag ? 1 : 0)) + sb * (size_t)(s + 2) + 256;
  if (c->tws_bytes < need) {
    INS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->tws) { cudaFree(c->tws); c->ws_bytes -= (i64)c->tws_bytes; c->tws = nullptr; c->tws_bytes = 0; }
    INS_CUDA(c, cudaMalloc(&c->tws, need));
    c->tws_bytes = need;
    c->ws_bytes += (i64)need;
  }
  T *w = (T *)c->tws;
  T *ustart = w; w += N * D;
  T *ku[INSB200_MAX_STAGES];
  for (int i = 0; i < s; i++) { ku[i] = w; w += N * D; }
  T *diff = w; w += N * D;
  T *sclo = nullptr;
  if (smag) { sclo = w; w += N * D; }
  T *tempstart = w; w += N;
  T *ktemp[INSB200_MAX_STAGES];
  for (int i = 0; i < s; i++) { ktemp[i] = w; w += N; }
  T *p = w;
  T *u = (T *)u_, *temp = (T *)temp_;
  T tt = T(*t);
  for (int n = 0; n < nsteps; n++) {
    const T tstart = tt;
    INS_CUDA(c, cudaMemcpyAsync(ustart, u, vb, cudaMemcpyDeviceToDevice, c->stream));        // :14
    INS_CUDA(c, cudaMemcpyAsync(tempstart, temp, sb, cudaMemcpyDeviceToDevice, c->stream));  // :15
    for (int i = 0; i < s; i++) {
      INS_TRY(insb200_apply_bc_u(c, u));                                                      // :19
      INS_TRY(insb200_apply_bc_p(c, temp));                                                   // :20 (PeriodicBC -> apply_bc_p!)
      INS_TRY(insb200_momentum(c, ku[i], u, (double)tt));                                     // :21
      INS_TRY((op_gravity<T, D>(c, ku[i], temp)));                                            //      gravity! (operators.jl:974)
      INS_CUDA(c, cudaMemsetAsync(ktemp[i], 0, sb, c->stream));                               // :23
      INS_TRY((op_convdiff_temp<T, D>(c, ktemp[i], u, temp)));                                // :24
      if (c->tq.dodissipation) INS_TRY((op_dissipation<T, D>(c, ktemp[i], diff, u)));         // :25
      if (smag) {                                                                             // :29
        INS_TRY(insb200_smagorinsky(c, sclo, u, theta));
        k_add<T><<<(unsigned)((N * D + 255) / 256), 256, 0, c->stream>>>(ku[i], sclo, N * D);
        TQ_LAUNCHED(c);
      }
      tt = tstart + T(cc[i]) * T(dt);                                                         // :32
      CombineArgs<T> au, at;
      au.nk = at.nk = 0;
      for (int j = 0; j <= i; j++) {
        if (A[i * s + j] == 0.0) continue;  // adds +-0
        au.kin[au.nk] = ku[j]; au.coef[au.nk++] = T(dt) * T(A[i * s + j]);                    // `dt * A[i,j]` formed first (:37)
        at.kin[at.nk] = ktemp[j]; at.coef[at.nk++] = T(dt) * T(A[i * s + j]);
      }
      INS_TRY((combine<T>(c, u, ustart, au, N * D)));                                         // :35-38
      INS_TRY((combine<T>(c, temp, tempstart, at, N)));                                       // :40-43
      INS_TRY(insb200_apply_bc_u(c, u));                                                      // :48
      INS_TRY(insb200_project(c, u, (i == s - 1 && n == nsteps - 1 && p_out) ? p : nullptr)); // :49
    }
    INS_TRY(insb200_apply_bc_u(c, u));                                                        // :55
    INS_TRY(insb200_apply_bc_p(c, temp));                                                     // :56
  }
  if (p_out && nsteps > 0) INS_CUDA(c, cudaMemcpyAsync(p_out, p, sb, cudaMemcpyDeviceToDevice, c->stream));
  *t = (double)tt;
  return 0;
}

// right_hand_side!(dudt, u, params, t)  src/sciml.jl:37-49 (u is not touched: the ghost fill works on a copy)
template <typename T, int D>
static int op_right_hand_side(insb200_ctx *c, void *dudt, const void *u, double t) {
  const size_t vb = (size_t)ref_elems(c) * D * sizeof(T);
  void *tmp = nullptr;
  INS_CUDA(c, cudaMallocAsync(&tmp, vb, c->stream));
  int r = 0;
  if (cudaMemcpyAsync(tmp, u, vb, cudaMemcpyDeviceToDevice, c->stream) != cudaSuccess) r = fail(c, __FILE__, __LINE__, "CUDA: copy of u failed");
  if (!r) r = insb200_apply_bc_u(c, tmp);             // :44
  if (!r) r = insb200_momentum(c, dudt, tmp, t);      // :45
  if (!r) r = insb200_apply_bc_u(c, dudt);            // :46 (dudt = true is a plain periodic fill)
  if (!r) r = insb200_project(c, dudt, nullptr);      // :47
  cudaFreeAsync(tmp, c->stream);
  return r;
}

}  // namespace insb200

using namespace insb200;

#define TQ_DISPATCH(c, FN, ...)                                                                        \
  ((c)->dtype == INSB200_F32 ? ((c)->D == 2 ? FN<float, 2>(__VA_ARGS__) : FN<float, 3>(__VA_ARGS__))  \
                             : ((c)->D == 2 ? FN<double, 2>(__VA_ARGS__) : FN<double, 3>(__VA_ARGS__)))
#define TQ_PROLOG(c)  \
  if (!(c)) return 1; \
  if (cudaSetDevice((c)->device) != cudaSuccess) INS_FAIL((c), "cudaSetDevice failed")

extern "C" int insb200_set_temperature(insb200_ctx *c, double a1, double a2, double a3, double a4, double gamma, int dodissipation,
                                       int gdir) {
  TQ_PROLOG(c);
  if (gdir < 1 || gdir > c->D) INS_FAIL(c, "gdir must be in 1..D");
  if (!(gamma != 0.0)) INS_FAIL(c, "gamma must be non-zero");
  c->tq.set = true;
  c->tq.a1 = a1; c->tq.a2 = a2; c->tq.a3 = a3; c->tq.a4 = a4; c->tq.gamma = gamma;
  c->tq.dodissipation = dodissipation != 0;
  c->tq.gdir = gdir;
  return 0;
}
extern "C" int insb200_apply_bc_temp(insb200_ctx *c, void *temp) {
  TQ_PROLOG(c);
  if (!temp) INS_FAIL(c, "temp is NULL");
  return insb200_apply_bc_p(c, temp);
}
extern "C" int insb200_convection_diffusion_temp(insb200_ctx *c, void *cc, const void *u, const void *temp) {
  TQ_PROLOG(c);
  if (!cc || !u || !temp) INS_FAIL(c, "NULL field");
  INS_TRY(need_temperature(c));
  return TQ_DISPATCH(c, op_convdiff_temp, c, cc, u, temp);
}
extern "C" int insb200_dissipation(insb200_ctx *c, void *diss, void *diff, const void *u) {
  TQ_PROLOG(c);
  if (!diss || !diff || !u) INS_FAIL(c, "NULL field");
  INS_TRY(need_temperature(c));
  return TQ_DISPATCH(c, op_dissipation, c, diss, diff, u);
}
extern "C" int insb200_gravity(insb200_ctx *c, void *F, const void *temp) {
  TQ_PROLOG(c);
  if (!F || !temp) INS_FAIL(c, "NULL field");
  INS_TRY(need_temperature(c));
  return TQ_DISPATCH(c, op_gravity, c, F, temp);
}
extern "C" int insb200_momentum_temp(insb200_ctx *c, void *F, const void *u, const void *temp, double t) {
  TQ_PROLOG(c);
  if (!F || !u) INS_FAIL(c, "NULL field");
  INS_TRY(insb200_momentum(c, F, u, t));
  if (!temp) return 0;
  INS_TRY(need_temperature(c));
  return TQ_DISPATCH(c, op_gravity, c, F, temp);
}
static int check_tableau(insb200_ctx *c, int s, const double *A, const double *cc) {
  if (s < 1 || s > INSB200_MAX_STAGES) INS_FAIL(c, "number of stages out of range");
  if (!A || !cc) INS_FAIL(c, "tableau is NULL");
  for (int i = 0; i < s; i++)
    for (int j = i + 1; j < s; j++)
      if (A[i * s + j] != 0.0) INS_FAIL(c, "tableau is not explicit (shifted A must be lower triangular)");
  return 0;
}
extern "C" int insb200_rk_step_temp(insb200_ctx *c, void *u, void *temp, double t, double dt, int s, const double *A, const double *cc,
                                    double theta, void *p_out) {
  TQ_PROLOG(c);
  if (!u || !temp) INS_FAIL(c, "NULL field");
  INS_TRY(need_temperature(c));
  if (c->force_cb) INS_FAIL(c, "the temperature stepper takes steady body forces only (no unsteady-force callback)");
  INS_TRY(check_tableau(c, s, A, cc));
  double tt = t;
  return TQ_DISPATCH(c, op_rk_steps_temp, c, u, temp, &tt, dt, 1, s, A, cc, theta, p_out);
}
extern "C" int insb200_rk_steps_temp(insb200_ctx *c, void *u, void *temp, double *t, double dt, int nsteps, int s, const double *A,
                                     const double *cc, double theta) {
  TQ_PROLOG(c);
  if (!u || !temp || !t) INS_FAIL(c, "NULL argument");
  if (nsteps < 0) INS_FAIL(c, "nsteps < 0");
  INS_TRY(need_temperature(c));
  if (c->force_cb) INS_FAIL(c, "the temperature stepper takes steady body forces only (no unsteady-force callback)");
  INS_TRY(check_tableau(c, s, A, cc));
  return TQ_DISPATCH(c, op_rk_steps_temp, c, u, temp, t, dt, nsteps, s, A, cc, theta, nullptr);
}
extern "C" int insb200_right_hand_side(insb200_ctx *c, void *dudt, const void *u, double t) {
  TQ_PROLOG(c);
  if (!dudt || !u) INS_FAIL(c, "NULL field");
  if (c->force_cb) INS_FAIL(c, "right_hand_side! takes steady body forces only (no unsteady-force callback)");
  return TQ_DISPATCH(c, op_right_hand_side, c, dudt, u, t);
}
extern "C" int insb200_qfield(insb200_ctx *c, void *Q, const void *u) {
  TQ_PROLOG(c);
  if (!Q || !u) INS_FAIL(c, "NULL field");
  INS_TRY(single_rank(c, "Qfield!"));
  return TQ_DISPATCH(c, op_qfield, c, Q, u);
}
extern "C" int insb200_dissipation_from_strain(insb200_ctx *c, void *eps, const void *u) {
  TQ_PROLOG(c);
  if (!eps || !u) INS_FAIL(c, "NULL field");
  INS_TRY(single_rank(c, "dissipation_from_strain!"));
  return TQ_DISPATCH(c, op_strain_dissipation, c, eps, u);
}
extern "C" int insb200_scale_numbers(insb200_ctx *c, const void *u, double *out8) {
  TQ_PROLOG(c);
  if (!u || !out8) INS_FAIL(c, "NULL argument");
  INS_TRY(single_rank(c, "get_scale_numbers"));
  return TQ_DISPATCH(c, op_scale_numbers, c, u, out8);
}
