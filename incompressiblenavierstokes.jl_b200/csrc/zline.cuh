// zline.cuh -- last-axis part of the spectral pressure solve on a SLAB-DECOMPOSED grid WITHOUT transposes.
//
// The reference divides the 3-D spectrum by the symbol ax+ay+az (src/pressure.jl:323-344).  az[k] = 4 W sin^2(pi k/n)/dz^2
// is exactly the eigenvalue of the periodic second difference, so for one in-plane mode (kx, ky) "transform along z,
// divide, transform back" IS the solution of the periodic constant-coefficient tridiagonal system
//
//        p[j-1] - (2 + lam) p[j] + p[j+1] = h[j],      lam = (ax+ay) dz^2 / W,   h = (dz^2 / W) f^ / (nx ny)
//
// (W = cell volume, which the reference folds into its symbol).  With rho the root of rho^2 - (2+lam) rho + 1 = 0 inside
// the unit circle the operator factors as -(1/rho)(1 - rho S)(1 - rho S^-1), S the shift, i.e.
//
//        g[j] = h[j] + rho g[j-1]   (causal),      q[j] = g[j] + rho q[j+1]   (anticausal),      p = -rho q:
//
// two first-order recursions that MARCH through the planes of the in-slab spectrum [z local][ky][kx] -- the layout cuFFT's
// in-plane transforms leave, with no transpose.  A chunk of the line (a rank's slab, or a sub-chunk of it) needs from the
// others only the state each recursion enters it with, and both follow from two weighted sums per chunk and mode:
//
//   pass A (forward, in place): g0[j] = h[j] + rho g0[j-1] from g0[-1] = 0, and on the way
//        A_v = g0[mv-1] = sum_i rho^(mv-1-i) h[i],      B_v = sum_i rho^i h[i]          -> the chunk's exchange area
//   (peer barrier)
//   pass B (backward, in place): the true causal sequence is g[j] = g0[j] + rho^(j+1) G_(v-1) with
//        G_(v-1) = 1/(1-rho^n) sum_k rho^(mv k) A_(v-1-k)                                 (state entering from below)
//        Q_(v+1) = (rho G_v + 1/(1-rho^n) sum_k rho^(mv k) B_(v+1+k)) / (1 - rho^2)       (state entering from above)
//   and q[j] = g[j] + rho q[j+1] from q[mv] = Q_(v+1); p[j] = -rho * scale * q[j].
//
// The chain sums decay like rho^(mv k): all but a few hundred of the lowest modes need their nearest neighbour only, and
// the kernel reads exactly those words from the peers' exchange areas over NVLink (peer loads inside the solve kernel:
// the collective is fused into it).  Per solve a GPU reads 2 words per mode and chunk from its neighbours instead of
// sending 2 x 7/8 of its spectrum (8 GPUs, 1024^3, Float32: 8 MB instead of 824 MB), and the two passes run at HBM
// speed: 4 sweeps over the slab's spectrum.
//
// The mode kx = ky = 0 has lam = 0 (rho = 1, singular: the reference zeroes the constant mode): its n values are
// gathered and solved with the periodic Green's function of the 1-D second difference, G(d) = d (n-d) / (2n).
//
// Arithmetic is Float64 whatever the data type (lam -> rho loses half the digits for the lowest modes: 1 - rho ~
// sqrt(lam)); storage is the data type.  All data routines are __host__ __device__: tools/zline_emul.cu runs them on the
// CPU for P emulated ranks against a direct solve (tests/test_zline_emul.py, no GPU needed).
#pragma once
#include "comm.cuh"
#include "tiled.cuh"
#include "zsolve.cuh"

namespace insb200 {

struct ZlGeo {
  int nkx, pitch, nky;  // valid kx per row, row pitch and rows per plane (ny; 1 in 2-D), all in complex words
  long long pstride;    // complex words per plane (pitch * nky)
  int mloc, C, mv;      // planes of this rank, sub-chunks per rank, planes per sub-chunk
  int P, rank, n;       // ranks, this rank, points of the whole line (P * mloc)
  double kappa;         // dz^2 / W
  double scale;         // kappa / (points of the in-plane transforms)
  long long xoff;       // this solve's exchange area inside every rank's exchange buffer (complex words)
};
// exchange area of one rank: [C][A | B][pstride] sums, then the mloc values of the zero mode's column
__host__ __device__ inline long long zl_area_words(const ZlGeo &g) { return (long long)g.C * 2 * g.pstride + ((g.mloc + 15) & ~15); }

struct ZlMode {
  double rho, lr;  // root inside the unit circle and its logarithm
};
template <typename T>
__host__ __device__ inline ZlMode zl_mode(T axy, double kappa) {
  const double lam = (double)axy * kappa;
  ZlMode m;
  m.lr = -log1p(0.5 * lam + sqrt(lam * (1.0 + 0.25 * lam)));  // rho = 1 / (b + sqrt(b^2 - 1)), b = 1 + lam/2
  m.rho = exp(m.lr);
  return m;
}

constexpr int kZlU = 8;            // planes per batch of loads
constexpr double kZlTiny = 92.0;   // exp(-92) = 1e-40: below this weight a term cannot change a Float64 sum

// a word of an exchange area (possibly a peer's, over NVLink): a STRONG load at system scope -- coherent with the owner's
// stores whatever caches a weak load could be served from (the reading thread takes no part in the barrier itself)
template <typename T>
__host__ __device__ inline Cx<T> zl_ldx(const Cx<T> *p) {
#ifdef __CUDA_ARCH__
  if constexpr (sizeof(T) == 4) {
    float x, y;
    asm volatile("ld.relaxed.sys.global.v2.f32 {%0, %1}, [%2];" : "=f"(x), "=f"(y) : "l"(p) : "memory");
    return Cx<T>{x, y};
  } else {
    double x, y;
    asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "l"(p) : "memory");
    return Cx<T>{x, y};
  }
#else
  return *p;
#endif
}
// 8- / 16-byte accesses to the spectrum (Cx<T> alone only promises the alignment of T)
template <typename T>
__host__ __device__ inline Cx<T> zl_ld(const Cx<T> *p) {
#ifdef __CUDA_ARCH__
  if constexpr (sizeof(T) == 4) { const float2 v = *reinterpret_cast<const float2 *>(p); return Cx<T>{v.x, v.y}; }
  else { const double2 v = *reinterpret_cast<const double2 *>(p); return Cx<T>{v.x, v.y}; }
#else
  return *p;
#endif
}
template <typename T>
__host__ __device__ inline void zl_st(Cx<T> *p, double x, double y) {
#ifdef __CUDA_ARCH__
  if constexpr (sizeof(T) == 4) *reinterpret_cast<float2 *>(p) = make_float2((float)x, (float)y);
  else *reinterpret_cast<double2 *>(p) = make_double2(x, y);
#else
  *p = Cx<T>{(T)x, (T)y};
#endif
}

// ---- pass A: causal recursion from a zero state over a chunk, in place; returns A (the end state) and B -------------
template <typename T>
__host__ __device__ inline void zl_forward(Cx<T> *s, long long ps, int mv, double rho, double (&A)[2], double (&B)[2]) {
  double gx = 0, gy = 0, bx = 0, by = 0, pw = 1;
  int j = 0;
  for (; j + kZlU <= mv; j += kZlU) {
    Cx<T> v[kZlU];
#pragma unroll
    for (int u = 0; u < kZlU; u++) v[u] = zl_ld<T>(s + (long long)(j + u) * ps);
#pragma unroll
    for (int u = 0; u < kZlU; u++) {
      const double hx = (double)v[u].x, hy = (double)v[u].y;
      gx = hx + rho * gx; gy = hy + rho * gy;
      bx += pw * hx; by += pw * hy;
      pw *= rho;
      zl_st<T>(s + (long long)(j + u) * ps, gx, gy);
    }
  }
  for (; j < mv; j++) {
    const Cx<T> v = zl_ld<T>(s + (long long)j * ps);
    const double hx = (double)v.x, hy = (double)v.y;
    gx = hx + rho * gx; gy = hy + rho * gy;
    bx += pw * hx; by += pw * hy;
    pw *= rho;
    zl_st<T>(s + (long long)j * ps, gx, gy);
  }
  A[0] = gx; A[1] = gy; B[0] = bx; B[1] = by;
}

// ---- the chain over the chunks: 1/(1-rho^n) sum_k rho^(mv k) X_(v0 + dir k), X = A (which 0) or B (which 1) ------------
template <typename T>
__host__ __device__ inline void zl_chain(const PeerPtrs &X, const ZlGeo &g, long long moff, int which, int v0, int dir, double rm,
                                         double inv1n, double (&out)[2]) {
  const int V = g.P * g.C;
  double w = 1, sx = 0, sy = 0;
  int v = v0;
  for (int k = 0; k < V; k++) {
    const int q = v / g.C, cc = v - q * g.C;
    const Cx<T> a = zl_ldx<T>(reinterpret_cast<const Cx<T> *>(X.p[q]) + g.xoff + (long long)(cc * 2 + which) * g.pstride + moff);
    sx += w * (double)a.x; sy += w * (double)a.y;
    w *= rm;
    if (w < 1e-20) break;
    v += dir;
    if (v < 0) v += V;
    if (v >= V) v -= V;
  }
  out[0] = sx * inv1n; out[1] = sy * inv1n;
}

// ---- pass B: anticausal recursion over a chunk, in place: g[j] = g0[j] + rho^(j+1) Gin, q[j] = g[j] + rho q[j+1], writes
// cs * q.  rho^(j+1) is carried backwards by multiplying with 1/rho, started at the first plane (coming from above) whose
// weight is not negligible -- a power formed at the top of a long chunk would have underflowed to zero for good.
template <typename T>
__host__ __device__ inline void zl_backward(Cx<T> *s, long long ps, int mv, const ZlMode &m, double cs, double ginx, double giny,
                                            double qx, double qy) {
  const double rho = m.rho, rinv = exp(-m.lr);
  const double lim = kZlTiny / -m.lr;  // planes until rho^(j+1) < exp(-kZlTiny)
  const int jstar = lim >= (double)mv ? mv - 1 : (int)lim;
  const double pstar = exp(m.lr * (double)(jstar + 1));
  double pw = 0;
  int j = mv;
  for (; j - kZlU >= 0; j -= kZlU) {
    Cx<T> v[kZlU];
#pragma unroll
    for (int u = 0; u < kZlU; u++) v[u] = zl_ld<T>(s + (long long)(j - 1 - u) * ps);
#pragma unroll
    for (int u = 0; u < kZlU; u++) {
      const int jj = j - 1 - u;
      pw = jj == jstar ? pstar : pw * rinv;  // (0 * rinv stays 0 above jstar)
      qx = ((double)v[u].x + pw * ginx) + rho * qx; qy = ((double)v[u].y + pw * giny) + rho * qy;
      zl_st<T>(s + (long long)jj * ps, cs * qx, cs * qy);
    }
  }
  for (; j > 0; j--) {
    const int jj = j - 1;
    const Cx<T> v = zl_ld<T>(s + (long long)jj * ps);
    pw = jj == jstar ? pstar : pw * rinv;
    qx = ((double)v.x + pw * ginx) + rho * qx; qy = ((double)v.y + pw * giny) + rho * qy;
    zl_st<T>(s + (long long)jj * ps, cs * qx, cs * qy);
  }
}

// what one thread (in-plane mode `idx` = ky * pitch + kx, sub-chunk c) does in each pass -------------------------------
template <typename T>
__host__ __device__ inline bool zl_thread_mode(const ZlGeo &g, long long idx, const T *ax, const T *ay, ZlMode &m, bool &zero) {
  const int ky = (int)(idx / g.pitch), kx = (int)(idx - (long long)ky * g.pitch);
  if (ky >= g.nky || kx >= g.nkx) return false;
  zero = kx == 0 && ky == 0;
  const T axy = ay ? (T)(ax[kx] + ay[ky]) : ax[kx];
  m = zl_mode<T>(axy, g.kappa);
  return true;
}

template <typename T>
__host__ __device__ inline void zl_passA(const ZlGeo &g, const PeerPtrs &X, Cx<T> *S, const T *ax, const T *ay, long long idx, int c) {
  ZlMode m;
  bool zero;
  if (!zl_thread_mode<T>(g, idx, ax, ay, m, zero)) return;
  Cx<T> *s = S + (long long)c * g.mv * g.pstride + idx;
  Cx<T> *xa = reinterpret_cast<Cx<T> *>(X.p[g.rank]) + g.xoff;
  if (zero) {  // the column of the constant in-plane mode goes to the exchange area as it is
    Cx<T> *zc = xa + (long long)g.C * 2 * g.pstride + (long long)c * g.mv;
    for (int j = 0; j < g.mv; j++) zc[j] = s[(long long)j * g.pstride];
    return;
  }
  double A[2], B[2];
  zl_forward<T>(s, g.pstride, g.mv, m.rho, A, B);
  zl_st<T>(xa + (long long)(c * 2 + 0) * g.pstride + idx, A[0], A[1]);
  zl_st<T>(xa + (long long)(c * 2 + 1) * g.pstride + idx, B[0], B[1]);
}

template <typename T>
__host__ __device__ inline void zl_passB(const ZlGeo &g, const PeerPtrs &X, Cx<T> *S, const T *ax, const T *ay, long long idx, int c) {
  ZlMode m;
  bool zero;
  if (!zl_thread_mode<T>(g, idx, ax, ay, m, zero) || zero) return;
  const int V = g.P * g.C, v = g.rank * g.C + c;
  const double rm = exp(m.lr * g.mv), inv1n = -1.0 / expm1(m.lr * g.n), inv1r2 = -1.0 / expm1(2.0 * m.lr);
  Cx<T> *s = S + (long long)c * g.mv * g.pstride + idx;
  double Gin[2], Bs[2];
  zl_chain<T>(X, g, idx, 0, (v + V - 1) % V, -1, rm, inv1n, Gin);  // G_(v-1)
  zl_chain<T>(X, g, idx, 1, (v + 1) % V, +1, rm, inv1n, Bs);
  const Cx<T> a = zl_ld<T>(s + (long long)(g.mv - 1) * g.pstride);  // A_v = g0[mv-1], as stored
  const double gex = (double)a.x + rm * Gin[0], gey = (double)a.y + rm * Gin[1];  // G_v
  const double qx = (m.rho * gex + Bs[0]) * inv1r2, qy = (m.rho * gey + Bs[1]) * inv1r2;
  zl_backward<T>(s, g.pstride, g.mv, m, -m.rho * g.scale, Gin[0], Gin[1], qx, qy);
}

// the constant in-plane mode: p[j] = scale * sum_i G((j - i) mod n) (h[i] - mean h), G(d) = d (n - d) / (2n); hcol = the
// whole column (n values), j = global index of the point; the sum over i = i0, i0 + istep, ..
template <typename T2>
__host__ __device__ inline void zl_zero_point(const T2 *hcol, int n, int j, int i0, int istep, double mean_x, double mean_y, double &sx, double &sy) {
  sx = 0; sy = 0;
  const double i2n = 0.5 / (double)n;
  for (int i = i0; i < n; i += istep) {
    int d = j - i;
    if (d < 0) d += n;
    const double G = (double)d * (double)(n - d) * i2n;
    sx += G * (hcol[i].x - mean_x);
    sy += G * (hcol[i].y - mean_y);
  }
}

#ifdef __CUDACC__
// ---- kernels: grid.x over the in-plane modes (ky * pitch + kx: consecutive threads, consecutive words), grid.y = sub-chunk
template <typename T>
__global__ void __launch_bounds__(256) k_zline_forward(ZlGeo g, PeerPtrs X, Cx<T> *S, const T *__restrict__ ax, const T *__restrict__ ay) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  zl_passA<T>(g, X, S, ax, ay, idx, blockIdx.y);
}
template <typename T>
__global__ void __launch_bounds__(256) k_zline_backward(ZlGeo g, PeerPtrs X, Cx<T> *S, const T *__restrict__ ax, const T *__restrict__ ay) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  zl_passB<T>(g, X, S, ax, ay, idx, blockIdx.y);
}
// the constant mode's column: every block gathers the n values from the ranks' exchange areas; one WARP per point of
// this rank (8 per block), the lanes share the sum over i
constexpr int kZlZeroWarps = 8;
template <typename T>
__global__ void __launch_bounds__(32 * kZlZeroWarps) k_zline_zero(ZlGeo g, PeerPtrs X, Cx<T> *S) {
  extern __shared__ double2 zl_col[];  // [n]
  __shared__ double red[2][kZlZeroWarps];
  __shared__ double mean[2];
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
  double sx = 0, sy = 0;
  for (int i = tid; i < g.n; i += nt) {
    const int q = i / g.mloc, jl = i - q * g.mloc;
    const Cx<T> h = zl_ldx<T>(reinterpret_cast<const Cx<T> *>(X.p[q]) + g.xoff + (long long)g.C * 2 * g.pstride + jl);
    zl_col[i] = make_double2((double)h.x, (double)h.y);
    sx += (double)h.x; sy += (double)h.y;
  }
  for (int o = 16; o; o >>= 1) { sx += __shfl_xor_sync(0xffffffffu, sx, o); sy += __shfl_xor_sync(0xffffffffu, sy, o); }
  if (lane == 0) { red[0][warp] = sx; red[1][warp] = sy; }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0;
    for (int w = 0; w < kZlZeroWarps; w++) { a += red[0][w]; b += red[1][w]; }
    mean[0] = a / g.n; mean[1] = b / g.n;
  }
  __syncthreads();
  const int jl = blockIdx.x * kZlZeroWarps + warp;
  if (jl >= g.mloc) return;
  zl_zero_point<double2>(zl_col, g.n, g.rank * g.mloc + jl, lane, 32, mean[0], mean[1], sx, sy);
  for (int o = 16; o; o >>= 1) { sx += __shfl_xor_sync(0xffffffffu, sx, o); sy += __shfl_xor_sync(0xffffffffu, sy, o); }
  if (lane == 0) zl_st<T>(S + (long long)jl * g.pstride, g.scale * sx, g.scale * sy);
}
#endif

}  // namespace insb200
