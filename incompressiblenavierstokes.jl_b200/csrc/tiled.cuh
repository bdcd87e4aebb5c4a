// tiled.cuh -- shared-memory kernels of the RK stage (compact layout, 3-D): the hot kernels.
//
// k_stage3d: one launch per RK stage (step_explicit_runge_kutta.jl:17-50 up to the Poisson solve)
//   input   u      = stage velocity.  With PGRAD the projection of the PREVIOUS stage is applied while
//                    loading: u = u' - grad p (applypressure!, operators.jl:214-233), so the projected
//                    intermediate velocity is never written to HBM.
//   output  k_i    = momentum!(u)            (convection+diffusion [+ force] [+ div sigma])
//           u'     = ustart + sum_j (dt a_ij) k_j                                  (:35-38)
//           div    = divergence!(u') * Omega   (operators.jl:106-125, 81-95): the Poisson right-hand side
//
// Tiling: a block owns a TX x TY column of cells and marches along z over `zc` planes.  Four plane
// slots of (TX+3) x (TY+3) points x 3 components live in shared memory (halo -2..+1 in x and y:
// -1..+1 for the stencil plus one more on the minus side, where u'_x(i-1) and u'_y(j-1) are
// recomputed for the divergence).  Plane k+2 is fetched from HBM (coalesced rows) into registers
// while plane k is computed, then stored to the slot plane k-2 occupied: one barrier per plane.
// The z-difference of the divergence uses the thread's own u'_z of the previous plane; the first
// plane of every z-chunk is completed by k_divergence_planes.
//
// Arithmetic is the same device code as the one-thread-per-cell kernels (stencil.cuh), so results
// are bit-identical between the variants.
#pragma once
#include "stencil.cuh"

namespace insb200 {

template <int TX, int TY>
struct Tile3 {
  static constexpr int PW = TX + 3, PH = TY + 3, NP = PW * PH, NT = TX * TY, NL = (NP + NT - 1) / NT;
};

template <typename T>
struct SmemAcc {  // U(c,dx,dy,dz) on the plane slots; `c0` = index of the own cell inside a plane
  const T *s;     // [4][3][NP]
  int np, pw, c0, k;
  __device__ __forceinline__ T operator()(int c, int dx, int dy, int dz) const {
    return s[(((k + dz) & 3) * 3 + c) * np + c0 + dy * pw + dx];
  }
};

__device__ __forceinline__ int wrapi(int v, int n, int wrap) {
  if (wrap) { if (v < 0) v += n; else if (v >= n) v -= n; }
  return v;
}

// one component `a` of the stage at the cell (ci, cj, k): f = momentum!, return u' (the combination).
// `U` reads the velocity tile around that cell; oc = in-plane offset of the cell in global memory.
template <typename T, bool SMAG, bool PGRAD, class ACC>
__device__ __forceinline__ T stage_comp(const StageArgs<T> &A, const Geo &g, const ACC &U, const int a, const i64 oc,
                                        const int ci, const int cj, const int k, const T visc, const T (&rh)[3],
                                        T &f_out) {
  T f = convdiff_comp<T, 3, true, true>(U, a, visc, rh);
  const i64 og = (i64)k * A.sz + oc + a * A.sc;
  if (A.force) f = f + __ldg(A.force + og);
  if (SMAG) {
    Sten<3, 1> st(g, A.sy, A.sz, ci, cj, k);
    Acc<T, 3, 1> Sg{A.sigma, A.ssc, st};
    f = f + divsigma_comp<T, 3>(Sg, a, rh);
  }
  f_out = f;
  T acc = (!PGRAD && A.u0 == A.u) ? U(a, 0, 0, 0) : __ldg(A.u0 + og);
  for (int q = 0; q < A.nk; q++) acc = acc + A.coef[q] * __ldg(A.kin[q] + og);
  return acc + A.coef_self * f;
}

template <typename T, bool SMAG, bool PGRAD, int TX, int TY>
__global__ void __launch_bounds__(TX *TY) k_stage3d(StageArgs<T> A, Geo g, T visc, Rh<T> rh, T vol, int zc) {
  typedef Tile3<TX, TY> TL;
  constexpr int PW = TL::PW, NP = TL::NP, NT = TL::NT, NL = TL::NL;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T *su = reinterpret_cast<T *>(smem_raw);           // [4][3][NP] plane slots
  T *sdx = su + 4 * 3 * NP;                           // [2][TY][TX+1]  u'_x incl. column -1
  T *sdy = sdx + 2 * TY * (TX + 1);                   // [2][TY+1][TX]  u'_y incl. row -1

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TX + tx;
  const int i0 = blockIdx.x * TX, j0 = blockIdx.y * TY;
  const int kbeg = blockIdx.z * zc, kend = min(kbeg + zc, g.n[2]);
  const int n0 = g.n[0], n1 = g.n[1], n2 = g.n[2];
  const i64 sy = A.sy, sz = A.sz, sc = A.sc;

  // in-plane offsets of the points this thread stages (constant along z)
  i64 o[NL], ox1[NL], oy1[NL];
#pragma unroll
  for (int q = 0; q < NL; q++) {
    const int l = tid + q * NT;
    const int lx = l % PW, ly = l / PW;
    const int gi = wrapi(i0 - 2 + lx, n0, 1), gj = wrapi(j0 - 2 + ly, n1, 1);
    o[q] = (i64)gj * sy + gi;
    if (PGRAD) {
      ox1[q] = (i64)gj * sy + wrapi(gi + 1, n0, 1);
      oy1[q] = (i64)wrapi(gj + 1, n1, 1) * sy + gi;
    }
  }
  auto fetch = [&](int kk, T(&r)[NL][3]) {
    const i64 zo = (i64)wrapi(kk, n2, g.wrap[2]) * sz;
    const i64 zo1 = PGRAD ? (i64)wrapi(kk + 1, n2, g.wrap[2]) * sz : 0;
#pragma unroll
    for (int q = 0; q < NL; q++) {
      if (tid + q * NT < NP) {
        const i64 b = zo + o[q];
        if (PGRAD) {
          const T p0 = __ldg(A.pin + b);
          const T gx = (__ldg(A.pin + zo + ox1[q]) - p0) * rh.v[0];
          const T gy = (__ldg(A.pin + zo + oy1[q]) - p0) * rh.v[1];
          const T gz = (__ldg(A.pin + zo1 + o[q]) - p0) * rh.v[2];
          r[q][0] = __ldg(A.u + b) - gx;
          r[q][1] = __ldg(A.u + sc + b) - gy;
          r[q][2] = __ldg(A.u + 2 * sc + b) - gz;
        } else {
          r[q][0] = __ldg(A.u + b);
          r[q][1] = __ldg(A.u + sc + b);
          r[q][2] = __ldg(A.u + 2 * sc + b);
        }
      }
    }
  };
  auto stash = [&](int kk, const T(&r)[NL][3]) {
    T *dst = su + ((kk & 3) * 3) * NP;
#pragma unroll
    for (int q = 0; q < NL; q++) {
      const int l = tid + q * NT;
      if (l < NP) { dst[l] = r[q][0]; dst[NP + l] = r[q][1]; dst[2 * NP + l] = r[q][2]; }
    }
  };

  {  // prologue: planes kbeg-1, kbeg, kbeg+1
    T r0[NL][3], r1[NL][3], r2[NL][3];
    fetch(kbeg - 1, r0);
    fetch(kbeg, r1);
    fetch(kbeg + 1, r2);
    stash(kbeg - 1, r0);
    stash(kbeg, r1);
    stash(kbeg + 1, r2);
  }
  __syncthreads();

  const int gi = i0 + tx, gj = j0 + ty;
  const i64 ocell = (i64)gj * sy + gi;
  // the extra point of this thread on the minus halo: threads [0,TY): (-1, tid) component x;
  // threads [TY, TY+TX): (tid-TY, -1) component y
  const bool hx = tid < TY, hy = !hx && tid < TY + TX;
  const int hlx = hx ? -1 : tid - TY, hly = hx ? tid : -1;
  const int hgi = wrapi(i0 + hlx, n0, 1), hgj = wrapi(j0 + hly, n1, 1);
  const i64 ohalo = (i64)hgj * sy + hgi;

  T uz_prev = T(0);
  for (int k = kbeg; k < kend; k++) {
    T rn[NL][3];
    const bool more = k + 2 <= kend;  // plane k+2 is needed by iteration k+1 <= kend-1
    if (more) fetch(k + 2, rn);

    SmemAcc<T> U{su, NP, PW, (ty + 2) * PW + tx + 2, k};
    T un[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
      T f;
      un[a] = stage_comp<T, SMAG, PGRAD>(A, g, U, a, ocell, gi, gj, k, visc, rh.v, f);
      const i64 og = (i64)k * sz + ocell + a * sc;
      if (A.kout) A.kout[og] = f;
      A.unew[og] = un[a];
    }
    if (A.div) {
      T *bx = sdx + (k & 1) * TY * (TX + 1), *by = sdy + (k & 1) * (TY + 1) * TX;
      bx[ty * (TX + 1) + tx + 1] = un[0];
      by[(ty + 1) * TX + tx] = un[1];
      if (hx || hy) {
        SmemAcc<T> Uh{su, NP, PW, (hly + 2) * PW + hlx + 2, k};
        T f;
        if (hx) bx[hly * (TX + 1)] = stage_comp<T, SMAG, PGRAD>(A, g, Uh, 0, ohalo, hgi, hgj, k, visc, rh.v, f);
        else by[hlx] = stage_comp<T, SMAG, PGRAD>(A, g, Uh, 1, ohalo, hgi, hgj, k, visc, rh.v, f);
      }
    }
    if (more) stash(k + 2, rn);
    __syncthreads();
    if (A.div && k > kbeg) {
      const T *bx = sdx + (k & 1) * TY * (TX + 1), *by = sdy + (k & 1) * (TY + 1) * TX;
      T d = T(0);
      d += (un[0] - bx[ty * (TX + 1) + tx]) * rh.v[0];
      d += (un[1] - by[ty * TX + tx]) * rh.v[1];
      d += (un[2] - uz_prev) * rh.v[2];
      A.div[(i64)k * sz + ocell] = d * vol;
    }
    uz_prev = un[2];
  }
}

// divergence!(u') * Omega on the planes k = q*zc (first plane of every z-chunk of k_stage3d)
template <typename T>
__global__ void k_divergence_planes(T *__restrict__ out, const T *__restrict__ u, i64 sy, i64 sz, i64 sc, Geo g,
                                    Rh<T> rh, T vol, int zc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * zc;
  if (i >= g.n[0] || j >= g.n[1]) return;
  Sten<3, 1> s(g, sy, sz, i, j, k);
  Acc<T, 3, 1> U{u, sc, s};
  T d = T(0);
#pragma unroll
  for (int a = 0; a < 3; a++) d += (U(a, 0, 0, 0) - U(a, -(a == 0), -(a == 1), -(a == 2))) * rh.v[a];
  out[(i64)k * sz + (i64)j * sy + i] = d * vol;
}

}  // namespace insb200
