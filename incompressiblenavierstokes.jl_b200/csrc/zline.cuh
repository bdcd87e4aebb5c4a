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
// others only the state the recursion enters it with, and that follows from two weighted sums per chunk and mode,
//
//        A_v = sum_i rho^(mv-1-i) h[i],      B_v = sum_i rho^i h[i]         (pass 0, one read of the spectrum)
//
//        state entering chunk v from below:  G_(v-1) = 1/(1-rho^n) sum_k rho^(mv k) A_(v-1-k)
//        state entering chunk v from above:  Q_(v+1) = (rho G_v + 1/(1-rho^n) sum_k rho^(mv k) B_(v+1+k)) / (1 - rho^2)
//
// whose terms decay like rho^(mv k): all but a few hundred of the lowest modes need their nearest neighbour only, and the
// kernels read exactly those words from the peers' exchange areas over NVLink (peer loads inside the solve kernel: the
// collective is fused into it).  Per solve a GPU moves 2 words per mode and chunk instead of 2 x 7/8 of its spectrum
// (8 GPUs, 1024^3, Float32: 8 MB instead of 824 MB), and the slab passes run at HBM speed: 5 sweeps over the spectrum.
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

constexpr int kZlU = 8;  // planes per batch of loads

// a word of an exchange area (possibly a peer's, over NVLink): always fetched from its home L2, never from a stale L1 line
template <typename T>
__host__ __device__ inline Cx<T> zl_ldx(const Cx<T> *p) {
#ifdef __CUDA_ARCH__
  if constexpr (sizeof(T) == 4) { const float2 v = __ldcg(reinterpret_cast<const float2 *>(p)); return Cx<T>{v.x, v.y}; }
  else { const double2 v = __ldcg(reinterpret_cast<const double2 *>(p)); return Cx<T>{v.x, v.y}; }
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

// ---- pass 0: the two weighted sums of a chunk --------------------------------------------------------------------
template <typename T>
__host__ __device__ inline void zl_sums(const Cx<T> *__restrict__ s, long long ps, int mv, double rho, double (&A)[2], double (&B)[2]) {
  double ax = 0, ay = 0, bx = 0, by = 0, pw = 1;
  int j = 0;
  for (; j + kZlU <= mv; j += kZlU) {
    Cx<T> v[kZlU];
#pragma unroll
    for (int u = 0; u < kZlU; u++) v[u] = zl_ld<T>(s + (long long)(j + u) * ps);
#pragma unroll
    for (int u = 0; u < kZlU; u++) {
      const double hx = (double)v[u].x, hy = (double)v[u].y;
      ax = ax * rho + hx; ay = ay * rho + hy;
      bx += pw * hx; by += pw * hy;
      pw *= rho;
    }
  }
  for (; j < mv; j++) {
    const Cx<T> v = zl_ld<T>(s + (long long)j * ps);
    ax = ax * rho + (double)v.x; ay = ay * rho + (double)v.y;
    bx += pw * (double)v.x; by += pw * (double)v.y;
    pw *= rho;
  }
  A[0] = ax; A[1] = ay; B[0] = bx; B[1] = by;
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

// ---- pass 1: causal recursion over a chunk, in place ------------------------------------------------------------------
template <typename T>
__host__ __device__ inline void zl_causal(Cx<T> *s, long long ps, int mv, double rho, double gx, double gy) {
  int j = 0;
  for (; j + kZlU <= mv; j += kZlU) {
    Cx<T> v[kZlU];
#pragma unroll
    for (int u = 0; u < kZlU; u++) v[u] = zl_ld<T>(s + (long long)(j + u) * ps);
#pragma unroll
    for (int u = 0; u < kZlU; u++) {
      gx = (double)v[u].x + rho * gx; gy = (double)v[u].y + rho * gy;
      zl_st<T>(s + (long long)(j + u) * ps, gx, gy);
    }
  }
  for (; j < mv; j++) {
    const Cx<T> v = zl_ld<T>(s + (long long)j * ps);
    gx = (double)v.x + rho * gx; gy = (double)v.y + rho * gy;
    zl_st<T>(s + (long long)j * ps, gx, gy);
  }
}

// ---- pass 2: anticausal recursion over a chunk, in place; writes p = cs * q --------------------------------------------
template <typename T>
__host__ __device__ inline void zl_anti(Cx<T> *s, long long ps, int mv, double rho, double cs, double qx, double qy) {
  int j = mv;
  for (; j - kZlU >= 0; j -= kZlU) {
    Cx<T> v[kZlU];
#pragma unroll
    for (int u = 0; u < kZlU; u++) v[u] = zl_ld<T>(s + (long long)(j - 1 - u) * ps);
#pragma unroll
    for (int u = 0; u < kZlU; u++) {
      qx = (double)v[u].x + rho * qx; qy = (double)v[u].y + rho * qy;
      zl_st<T>(s + (long long)(j - 1 - u) * ps, cs * qx, cs * qy);
    }
  }
  for (; j > 0; j--) {
    const Cx<T> v = zl_ld<T>(s + (long long)(j - 1) * ps);
    qx = (double)v.x + rho * qx; qy = (double)v.y + rho * qy;
    zl_st<T>(s + (long long)(j - 1) * ps, cs * qx, cs * qy);
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
__host__ __device__ inline void zl_pass0(const ZlGeo &g, const PeerPtrs &X, const Cx<T> *S, const T *ax, const T *ay, long long idx, int c) {
  ZlMode m;
  bool zero;
  if (!zl_thread_mode<T>(g, idx, ax, ay, m, zero)) return;
  const Cx<T> *s = S + (long long)c * g.mv * g.pstride + idx;
  Cx<T> *xa = reinterpret_cast<Cx<T> *>(X.p[g.rank]) + g.xoff;
  if (zero) {  // the column of the constant in-plane mode goes to the exchange area as it is
    Cx<T> *zc = xa + (long long)g.C * 2 * g.pstride + (long long)c * g.mv;
    for (int j = 0; j < g.mv; j++) zc[j] = s[(long long)j * g.pstride];
    return;
  }
  double A[2], B[2];
  zl_sums<T>(s, g.pstride, g.mv, m.rho, A, B);
  zl_st<T>(xa + (long long)(c * 2 + 0) * g.pstride + idx, A[0], A[1]);
  zl_st<T>(xa + (long long)(c * 2 + 1) * g.pstride + idx, B[0], B[1]);
}

template <typename T>
__host__ __device__ inline void zl_pass1(const ZlGeo &g, const PeerPtrs &X, Cx<T> *S, const T *ax, const T *ay, long long idx, int c) {
  ZlMode m;
  bool zero;
  if (!zl_thread_mode<T>(g, idx, ax, ay, m, zero) || zero) return;
  const int V = g.P * g.C, v = g.rank * g.C + c;
  const double rm = exp(m.lr * g.mv), inv1n = -1.0 / expm1(m.lr * g.n);
  double G[2];
  zl_chain<T>(X, g, idx, 0, (v + V - 1) % V, -1, rm, inv1n, G);
  zl_causal<T>(S + (long long)c * g.mv * g.pstride + idx, g.pstride, g.mv, m.rho, G[0], G[1]);
}

template <typename T>
__host__ __device__ inline void zl_pass2(const ZlGeo &g, const PeerPtrs &X, Cx<T> *S, const T *ax, const T *ay, long long idx, int c) {
  ZlMode m;
  bool zero;
  if (!zl_thread_mode<T>(g, idx, ax, ay, m, zero) || zero) return;
  const int V = g.P * g.C, v = g.rank * g.C + c;
  const double rm = exp(m.lr * g.mv), inv1n = -1.0 / expm1(m.lr * g.n), inv1r2 = -1.0 / expm1(2.0 * m.lr);
  Cx<T> *s = S + (long long)c * g.mv * g.pstride + idx;
  const Cx<T> ge = zl_ld<T>(s + (long long)(g.mv - 1) * g.pstride);  // G_v: the causal state this chunk ended with
  double Bs[2];
  zl_chain<T>(X, g, idx, 1, (v + 1) % V, +1, rm, inv1n, Bs);
  const double qx = (m.rho * (double)ge.x + Bs[0]) * inv1r2, qy = (m.rho * (double)ge.y + Bs[1]) * inv1r2;
  zl_anti<T>(s, g.pstride, g.mv, m.rho, -m.rho * g.scale, qx, qy);
}

// the constant in-plane mode: p[j] = scale * sum_i G((j - i) mod n) (h[i] - mean h), G(d) = d (n - d) / (2n); hcol = the
// whole column (n values), j = global index of the point
template <typename T2>
__host__ __device__ inline void zl_zero_point(const T2 *hcol, int n, int j, double mean_x, double mean_y, double scale, double &px, double &py) {
  double sx = 0, sy = 0;
  const double i2n = 0.5 / (double)n;
  for (int i = 0; i < n; i++) {
    int d = j - i;
    if (d < 0) d += n;
    const double G = (double)d * (double)(n - d) * i2n;
    sx += G * (hcol[i].x - mean_x);
    sy += G * (hcol[i].y - mean_y);
  }
  px = scale * sx; py = scale * sy;
}

#ifdef __CUDACC__
// ---- kernels: grid.x over the in-plane modes (ky * pitch + kx: consecutive threads, consecutive words), grid.y = sub-chunk
template <typename T>
__global__ void __launch_bounds__(256) k_zline_sums(ZlGeo g, PeerPtrs X, const Cx<T> *__restrict__ S, const T *__restrict__ ax,
                                                    const T *__restrict__ ay) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  zl_pass0<T>(g, X, S, ax, ay, idx, blockIdx.y);
}
template <typename T>
__global__ void __launch_bounds__(256) k_zline_causal(ZlGeo g, PeerPtrs X, Cx<T> *S, const T *__restrict__ ax, const T *__restrict__ ay) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  zl_pass1<T>(g, X, S, ax, ay, idx, blockIdx.y);
}
template <typename T>
__global__ void __launch_bounds__(256) k_zline_anti(ZlGeo g, PeerPtrs X, Cx<T> *S, const T *__restrict__ ax, const T *__restrict__ ay) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  zl_pass2<T>(g, X, S, ax, ay, idx, blockIdx.y);
}
// one block: gathers the column of the constant mode from every rank's exchange area and solves this rank's points
template <typename T>
__global__ void __launch_bounds__(256) k_zline_zero(ZlGeo g, PeerPtrs X, Cx<T> *S) {
  extern __shared__ double2 zl_col[];  // [n]
  __shared__ double red[2][8];
  __shared__ double mean[2];
  const int tid = threadIdx.x, nt = blockDim.x;
  double sx = 0, sy = 0;
  for (int i = tid; i < g.n; i += nt) {
    const int q = i / g.mloc, jl = i - q * g.mloc;
    const Cx<T> h = zl_ldx<T>(reinterpret_cast<const Cx<T> *>(X.p[q]) + g.xoff + (long long)g.C * 2 * g.pstride + jl);
    zl_col[i] = make_double2((double)h.x, (double)h.y);
    sx += (double)h.x; sy += (double)h.y;
  }
  for (int o = 16; o; o >>= 1) { sx += __shfl_xor_sync(0xffffffffu, sx, o); sy += __shfl_xor_sync(0xffffffffu, sy, o); }
  if ((tid & 31) == 0) { red[0][tid >> 5] = sx; red[1][tid >> 5] = sy; }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0;
    for (int w = 0; w < (nt + 31) / 32; w++) { a += red[0][w]; b += red[1][w]; }
    mean[0] = a / g.n; mean[1] = b / g.n;
  }
  __syncthreads();
  for (int jl = tid; jl < g.mloc; jl += nt) {
    double px, py;
    zl_zero_point<double2>(zl_col, g.n, g.rank * g.mloc + jl, mean[0], mean[1], g.scale, px, py);
    S[(long long)jl * g.pstride] = Cx<T>{(T)px, (T)py};
  }
}
#endif

}  // namespace insb200
