// stencil.cuh -- device arithmetic of the staggered-grid operators and the "one thread per cell"
// kernels that work on ANY layout (reference/ghosted or compact) through a View + Geo pair.
//
// Arithmetic contract (checked bit-for-bit against the oracle when h is a power of two):
//   * same operation order as the reference kernels (cited per function),
//   * the library is compiled with -fmad=false, so a*b+c is never contracted (Julia does not
//     contract either),
//   * x / h is evaluated as x * (1/h): exact when h is a power of two (every BASELINE config:
//     unit box, n = 2^m), otherwise within 1 ulp per division;  x/2, x/4 -> x*0.5, x*0.25 (exact).
#pragma once
#include "ctx.cuh"

namespace insb200 {

// ---------------------------------------------------------------------------------------------
// neighbour addressing
// ---------------------------------------------------------------------------------------------
template <int D, int R>
struct Sten {
  i64 ox[2 * R + 1], oy[2 * R + 1], oz[2 * R + 1];
  __device__ __forceinline__ Sten(const Geo &g, i64 sy, i64 sz, int i, int j, int k) {
#pragma unroll
    for (int d = -R; d <= R; d++) {
      int ii = i + d, jj = j + d, kk = k + d;
      if (g.wrap[0]) { if (ii < 0) ii += g.n[0]; else if (ii >= g.n[0]) ii -= g.n[0]; }
      if (g.wrap[1]) { if (jj < 0) jj += g.n[1]; else if (jj >= g.n[1]) jj -= g.n[1]; }
      ox[d + R] = ii;
      oy[d + R] = (i64)jj * sy;
      if (D == 3) {
        if (g.wrap[2]) { if (kk < 0) kk += g.n[2]; else if (kk >= g.n[2]) kk -= g.n[2]; }
        oz[d + R] = (i64)kk * sz;
      } else {
        oz[d + R] = 0;
      }
    }
  }
  __device__ __forceinline__ i64 at(int dx, int dy, int dz) const {
    return ox[dx + R] + oy[dy + R] + (D == 3 ? oz[dz + R] : 0);
  }
};

template <typename T, int D, int R>
struct Acc {  // read accessor of a multi-component field at a stencil
  const T *__restrict__ p;
  i64 sc;
  const Sten<D, R> &s;
  __device__ __forceinline__ T operator()(int c, int dx, int dy, int dz) const {
    return __ldg(p + c * sc + s.at(dx, dy, dz));
  }
};

// read accessor of u = u' - grad p (applypressure! fused into the read, operators.jl:225-233);
// the stencil `s` has radius 2 because p is read one cell beyond the velocity stencil.
template <typename T, int D>
struct GradAcc {
  const T *__restrict__ u;
  const T *__restrict__ p;
  i64 sc;
  T rh[3];
  const Sten<D, 2> &s;
  __device__ __forceinline__ T operator()(int c, int dx, int dy, int dz) const {
    const T p0 = __ldg(p + s.at(dx, dy, dz));
    const T p1 = __ldg(p + s.at(dx + (c == 0), dy + (c == 1), dz + (c == 2)));
    return __ldg(u + c * sc + s.at(dx, dy, dz)) - (p1 - p0) * rh[c];
  }
};

// ---------------------------------------------------------------------------------------------
// convection_diffusion_kernel!  src/operators.jl:647-690 (also convection_kernel! :392-415 and
// diffusion_kernel! :551-573 through the CONV / DIFF switches), one component `a` at one cell.
// `U(c,dx,dy,dz)` returns u[I + (dx,dy,dz), c].
// ---------------------------------------------------------------------------------------------
template <typename T, int D, bool CONV, bool DIFF, class A>
__device__ __forceinline__ T convdiff_comp(const A &U, const int a, const T visc, const T (&rh)[3]) {
  // Algebra: (x/2)*(y/2+z/2) == ((x)*(y+z))/4 and (a/h - b/h) == (a-b)/h hold bit-for-bit when h is a
  // power of two (scaling by 2^k commutes with rounding), so the forms below reproduce the reference's
  // operation order exactly there with 16 instead of 22 flops per (a, b) pair.
  T f = T(0);
  const T uI = U(a, 0, 0, 0);
  const int ax = (a == 0), ay = (a == 1), az = (a == 2);
#pragma unroll
  for (int b = 0; b < D; b++) {
    const int bx = (b == 0), by = (b == 1), bz = (b == 2);
    const T um = U(a, -bx, -by, -bz);
    const T up = U(a, bx, by, bz);
    T conv = T(0), diff = T(0);
    if (CONV) {
      const T x1 = (um + uI) * (U(b, -bx, -by, -bz) + U(b, ax - bx, ay - by, az - bz));
      const T x2 = (uI + up) * (U(b, 0, 0, 0) + U(b, ax, ay, az));
      conv = (x2 - x1) * T(0.25);
    }
    if (DIFF) diff = ((up - uI) - (uI - um)) * rh[b];
    if (CONV && DIFF) f += (visc * diff - conv) * rh[b];
    else if (CONV) f -= conv * rh[b];
    else f += visc * diff * rh[b];
  }
  return f;
}

// ---------------------------------------------------------------------------------------------
// Smagorinsky: dx / grad / smagtensor!  src/operators.jl:1023-1033, 1072-1088, 1135-1151.
// sigma is symmetric; stored components (index): 2-D: xx(0) xy(1) yy(2); 3-D: xx(0) xy(1) xz(2)
// yy(3) yz(4) zz(5).
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ int symidx(int D, int a, int b) {
  if (a > b) { int t = a; a = b; b = t; }
  return D == 2 ? a + b : (a == 0 ? b : (a == 1 ? 2 + b : 5));
}

template <typename T, int D, class A>
__device__ __forceinline__ T ddx(const A &U, const int a, const int b, const T (&rh)[3]) {
  const int ax = (a == 0), ay = (a == 1), az = (a == 2);
  const int bx = (b == 0), by = (b == 1), bz = (b == 2);
  if (a == b) return (U(a, 0, 0, 0) - U(a, -bx, -by, -bz)) * rh[b];
  // ((A/h + B/h) + C/h + D/h)/4 == (((A+B)+C)+D) * (1/h * 0.25) bit-for-bit for power-of-two h
  return ((((U(a, bx, by, bz) - U(a, 0, 0, 0)) + (U(a, bx - ax, by - ay, bz - az) - U(a, -ax, -ay, -az))) +
           (U(a, 0, 0, 0) - U(a, -bx, -by, -bz))) +
          (U(a, -ax, -ay, -az) - U(a, -ax - bx, -ay - by, -az - bz))) * (rh[b] * T(0.25));
}

// sig[symidx] at one p-cell; c_eddy = theta^2 * d^2 (formed on the host in T, left to right)
template <typename T, int D, class A>
__device__ __forceinline__ void smag_sigma(const A &U, const T c_eddy, const T (&rh)[3], T (&sig)[6]) {
  T G[3][3];
#pragma unroll
  for (int a = 0; a < D; a++)
#pragma unroll
    for (int b = 0; b < D; b++) G[a][b] = ddx<T, D>(U, a, b, rh);
  T S[3][3];
#pragma unroll
  for (int a = 0; a < D; a++)
#pragma unroll
    for (int b = 0; b < D; b++) S[a][b] = (G[a][b] + G[b][a]) * T(0.5);
  // sum(S .* S): column-major order of the SMatrix (row index fastest), left fold
  T ss = T(0);
  bool first = true;
#pragma unroll
  for (int b = 0; b < D; b++)
#pragma unroll
    for (int a = 0; a < D; a++) {
      const T t = S[a][b] * S[a][b];
      ss = first ? t : ss + t;
      first = false;
    }
  const T eddy = c_eddy * sqrt(T(2) * ss);
  const T two_eddy = T(2) * eddy;
#pragma unroll
  for (int a = 0; a < D; a++)
#pragma unroll
    for (int b = a; b < D; b++) sig[symidx(D, a, b)] = two_eddy * S[a][b];
}

// divoftensor_kernel!  src/operators.jl:1205-1238, one component at one u-point.
// `Sg(idx,dx,dy,dz)` returns sigma[I + (dx,dy,dz)][component idx].
template <typename T, int D, class A>
__device__ __forceinline__ T divsigma_comp(const A &Sg, const int a, const T (&rh)[3]) {
  const int ax = (a == 0), ay = (a == 1), az = (a == 2);
  T s = T(0);
#pragma unroll
  for (int b = 0; b < D; b++) {
    const int bx = (b == 0), by = (b == 1), bz = (b == 2);
    const int c = symidx(D, a, b);
    T s1, s2;
    if (a == b) {
      s2 = Sg(c, bx, by, bz);
      s1 = Sg(c, 0, 0, 0);
    } else {
      s2 = (Sg(c, 0, 0, 0) + Sg(c, bx, by, bz) + Sg(c, ax + bx, ay + by, az + bz) + Sg(c, ax, ay, az)) * T(0.25);
      s1 = (Sg(c, -bx, -by, -bz) + Sg(c, 0, 0, 0) + Sg(c, ax - bx, ay - by, az - bz) + Sg(c, ax, ay, az)) * T(0.25);
    }
    s += (s2 - s1) * rh[b];
  }
  return s;
}

// ---------------------------------------------------------------------------------------------
// kernels: one thread per cell, x fastest -> coalesced; neighbours through L1/L2
// ---------------------------------------------------------------------------------------------
#define INS_CELL_IDX(g)                                           \
  const int i = blockIdx.x * blockDim.x + threadIdx.x;            \
  const int j = blockIdx.y * blockDim.y + threadIdx.y;            \
  const int k = blockIdx.z;                                       \
  if (i >= (g).n[0] || j >= (g).n[1]) return;

template <typename T>
struct Rh {
  T v[3];
};

// copy the interior of one view into another (import / export between layouts); nc components
template <typename T>
__global__ void k_copy_view(View<T> dst, View<const T> src, Geo g, int nc) {
  INS_CELL_IDX(g)
  for (int c = 0; c < nc; c++)
    dst.p[c * dst.sc + k * dst.sz + j * dst.sy + i] = src.p[c * src.sc + k * src.sz + j * src.sy + i];
}

// the same copy with four cells per thread: 16-byte accesses on the side whose rows are 16-byte aligned (the internal
// padded / compact layout; the reference layout's rows start 4 or 8 bytes off), scalar accesses on the other.
// VDST: the destination is the aligned side (import), else the source is (export).  Needs n[0] % 4 == 0.
template <typename T, bool VDST>
__global__ void k_copy_view4(View<T> dst, View<const T> src, Geo g, int nc) {
  const int i = 4 * (blockIdx.x * blockDim.x + threadIdx.x);
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z;
  if (i >= g.n[0] || j >= g.n[1]) return;
  struct alignas(4 * sizeof(T)) V4 { T v[4]; };
  for (int c = 0; c < nc; c++) {
    T *d = dst.p + c * dst.sc + k * dst.sz + j * dst.sy + i;
    const T *s = src.p + c * src.sc + k * src.sz + j * src.sy + i;
    V4 v;
    if (VDST) {
#pragma unroll
      for (int q = 0; q < 4; q++) v.v[q] = __ldg(s + q);
      *reinterpret_cast<V4 *>(d) = v;
    } else {
      v = *reinterpret_cast<const V4 *>(s);
#pragma unroll
      for (int q = 0; q < 4; q++) d[q] = v.v[q];
    }
  }
}

// periodic ghost fill of a reference-layout array: the two ghost layers normal to `axis`, over
// the FULL extent of the other axes; the source index is wrapped in every axis, which equals the
// reference's axis-by-axis slab copies (src/boundary_conditions.jl:276-288, 306-318).
// base = pointer to element (1,1,1) Julia of component 0; N = n + 2.  If `only_local` the axis
// given by skip_axis is not wrapped (its ghosts come from the neighbour ranks).
template <typename T>
__global__ void k_bc_axis(T *base, i64 sy, i64 sz, i64 sc, int n0, int n1, int n2, int D, int nc, int axis,
                          int skip_axis) {
  // thread space: the other two axes cover -1..n (N points), the `axis` dimension has 2 layers
  int t0 = blockIdx.x * blockDim.x + threadIdx.x;
  int t1 = blockIdx.y * blockDim.y + threadIdx.y;
  int side = blockIdx.z & 1;
  int c = blockIdx.z >> 1;
  int n[3] = {n0, n1, n2};
  int idx[3];
  int o0 = axis == 0 ? 1 : 0, o1 = axis == 2 ? 1 : 2;  // the two non-axis dims
  if (D == 2) { o0 = axis == 0 ? 1 : 0; o1 = 2; }
  if (t0 >= n[o0] + 2) return;
  if (D == 3 ? (t1 >= n[o1] + 2) : (t1 >= 1)) return;
  idx[o0] = t0 - 1;
  idx[o1] = D == 3 ? t1 - 1 : 0;
  idx[axis] = side ? n[axis] : -1;
  if (D == 2) idx[2] = 0;
  int s[3];
  for (int d = 0; d < 3; d++) {
    s[d] = idx[d];
    if (d < D && d != skip_axis) { if (s[d] < 0) s[d] += n[d]; else if (s[d] >= n[d]) s[d] -= n[d]; }
  }
  T *pc = base + c * sc;
  pc[idx[2] * sz + idx[1] * sy + idx[0]] = pc[s[2] * sz + s[1] * sy + s[0]];
}

// momentum! (src/operators.jl:967-976) = convectiondiffusion! + applybodyforce! written (not
// accumulated) into F's interior.  force may be null.
template <typename T, int D, bool CONV, bool DIFF>
__global__ void k_momentum(View<T> F, View<const T> u, View<const T> force, Geo g, T visc, Rh<T> rh) {
  INS_CELL_IDX(g)
  Sten<D, 1> s(g, u.sy, u.sz, i, j, k);
  Acc<T, D, 1> U{u.p, u.sc, s};
  const i64 o = k * F.sz + j * F.sy + i;
#pragma unroll
  for (int a = 0; a < D; a++) {
    T f = convdiff_comp<T, D, CONV, DIFF>(U, a, visc, rh.v);
    if (force.p) f = f + force.p[a * force.sc + k * force.sz + j * force.sy + i];
    F.p[a * F.sc + o] = f;
  }
}

// Fused momentum + RK stage combination (step_explicit_runge_kutta.jl:21-38), compact layout.
//   k_i  = convdiff(u) [+ force] [+ div sigma]                 -> kout (if non-null)
//   unew = u0 + sum_j coef[j]*kin[j] (ascending j) + coef_self*k_i
template <typename T>
struct StageArgs {
  const T *u;       // stage input (ghost planes valid when not wrapped)
  const T *u0;      // ustart
  const T *kin[INSB200_MAX_STAGES];
  T coef[INSB200_MAX_STAGES];
  int nk;
  T coef_self;
  T *kout;          // nullable
  T *unew;
  const T *force;   // nullable (compact)
  const T *sigma;   // nullable (compact, ghost planes valid when not wrapped)
  const T *pin;     // tiled kernels with PGRAD: multiplier p of the previous stage, u = u' - grad p on load
  T *div;           // tiled kernels: nullable, receives divergence(unew) * Omega (the Poisson right-hand side)
  i64 sy, sz, sc;   // strides shared by all compact vector buffers
  i64 ssc;          // component stride of sigma
  const T *sadd;    // TMA kernels: closure term s = div sigma added to k_i (nullable)
  int u0_is_u;      // TMA kernels: ustart is the stage input itself (first stage): take it from the tile
};

template <typename T, int D, bool SMAG>
__global__ void k_momentum_stage(StageArgs<T> A, Geo g, T visc, Rh<T> rh) {
  INS_CELL_IDX(g)
  Sten<D, 1> s(g, A.sy, A.sz, i, j, k);
  Acc<T, D, 1> U{A.u, A.sc, s};
  Acc<T, D, 1> Sg{A.sigma, A.ssc, s};
  const i64 o = k * A.sz + j * A.sy + i;
#pragma unroll
  for (int a = 0; a < D; a++) {
    T f = convdiff_comp<T, D, true, true>(U, a, visc, rh.v);
    if (A.force) f = f + A.force[a * A.sc + o];
    if (SMAG) f = f + divsigma_comp<T, D>(Sg, a, rh.v);
    if (A.kout) A.kout[a * A.sc + o] = f;
    T acc = A.u0[a * A.sc + o];
    for (int q = 0; q < A.nk; q++) acc = acc + A.coef[q] * A.kin[q][a * A.sc + o];
    acc = acc + A.coef_self * f;
    A.unew[a * A.sc + o] = acc;
  }
}

// smagtensor! on all local p-cells (src/operators.jl:1135-1151) into the compact sigma buffer
// `pin` non-null: the velocity is u - grad pin (projection of the previous stage fused into the read)
template <typename T, int D>
__global__ void k_smagtensor(View<T> sg, View<const T> u, const T *__restrict__ pin, Geo g, T c_eddy, Rh<T> rh) {
  INS_CELL_IDX(g)
  T sig[6];
  if (pin) {
    Sten<D, 2> s(g, u.sy, u.sz, i, j, k);
    GradAcc<T, D> U{u.p, pin, u.sc, {rh.v[0], rh.v[1], rh.v[2]}, s};
    smag_sigma<T, D>(U, c_eddy, rh.v, sig);
  } else {
    Sten<D, 1> s(g, u.sy, u.sz, i, j, k);
    Acc<T, D, 1> U{u.p, u.sc, s};
    smag_sigma<T, D>(U, c_eddy, rh.v, sig);
  }
  const i64 o = k * sg.sz + j * sg.sy + i;
#pragma unroll
  for (int c = 0; c < D * (D + 1) / 2; c++) sg.p[c * sg.sc + o] = sig[c];
}

// divoftensor! (src/operators.jl:1166-1238): sigma view geometry gs (compact), output any layout
template <typename T, int D>
__global__ void k_divtensor(View<T> out, View<const T> sg, Geo gs, Rh<T> rh) {
  INS_CELL_IDX(gs)
  Sten<D, 1> s(gs, sg.sy, sg.sz, i, j, k);
  Acc<T, D, 1> Sg{sg.p, sg.sc, s};
  const i64 o = k * out.sz + j * out.sy + i;
#pragma unroll
  for (int a = 0; a < D; a++) out.p[a * out.sc + o] = divsigma_comp<T, D>(Sg, a, rh.v);
}

// divergence! (+ scalewithvolume!)  src/operators.jl:106-125, 81-95:  out = (sum_a (u-um)/h_a) * vol
template <typename T, int D>
__global__ void k_divergence(View<T> out, View<const T> u, Geo g, Rh<T> rh, T vol, int scale) {
  INS_CELL_IDX(g)
  Sten<D, 1> s(g, u.sy, u.sz, i, j, k);
  Acc<T, D, 1> U{u.p, u.sc, s};
  T d = T(0);
#pragma unroll
  for (int a = 0; a < D; a++) d += (U(a, 0, 0, 0) - U(a, -(a == 0), -(a == 1), -(a == 2))) * rh.v[a];
  if (scale) d = d * vol;
  out.p[k * out.sz + j * out.sy + i] = d;
}

// pressuregradient! / applypressure!  src/operators.jl:159-178, 214-233.
// MODE 0: G = grad p ; MODE 1: u -= grad p.  gp = geometry (wrap flags) of the p view.
template <typename T, int D, int MODE>
__global__ void k_gradient(View<T> out, View<const T> p, Geo gp, Rh<T> rh) {
  INS_CELL_IDX(gp)
  Sten<D, 1> s(gp, p.sy, p.sz, i, j, k);
  const T p0 = __ldg(p.p + s.at(0, 0, 0));
  const i64 o = k * out.sz + j * out.sy + i;
#pragma unroll
  for (int a = 0; a < D; a++) {
    const T gr = (__ldg(p.p + s.at(a == 0, a == 1, a == 2)) - p0) * rh.v[a];
    if (MODE == 0) out.p[a * out.sc + o] = gr;
    else out.p[a * out.sc + o] = out.p[a * out.sc + o] - gr;
  }
}

// 3-D form marching ZPT planes per thread: p(k+1) of one plane is p(k) of the next (one load less per
// cell), the x / y wrap is resolved once, and the grid has ZPT times fewer blocks.
template <typename T, int MODE, int ZPT>
__global__ void k_gradient_z(View<T> out, View<const T> p, Geo gp, Rh<T> rh) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k0 = blockIdx.z * ZPT;
  if (i >= gp.n[0] || j >= gp.n[1]) return;
  int ip = i + 1, jp = j + 1;
  if (gp.wrap[0] && ip >= gp.n[0]) ip -= gp.n[0];
  if (gp.wrap[1] && jp >= gp.n[1]) jp -= gp.n[1];
  const i64 pc = (i64)j * p.sy + i, px = (i64)j * p.sy + ip, py = (i64)jp * p.sy + i;
  const int nk = min(ZPT, gp.n[2] - k0);
  T pz[ZPT + 1], pxv[ZPT], pyv[ZPT];
#pragma unroll
  for (int m = 0; m <= ZPT; m++) {
    int kk = k0 + m;
    if (gp.wrap[2] && kk >= gp.n[2]) kk -= gp.n[2];
    if (m <= nk) pz[m] = __ldg(p.p + (i64)kk * p.sz + pc);
    if (m < nk) {
      pxv[m] = __ldg(p.p + (i64)kk * p.sz + px);
      pyv[m] = __ldg(p.p + (i64)kk * p.sz + py);
    }
  }
  T *o = out.p + (i64)k0 * out.sz + (i64)j * out.sy + i;
  if (MODE == 0) {
#pragma unroll
    for (int m = 0; m < ZPT; m++)
      if (m < nk) {
        o[m * out.sz] = (pxv[m] - pz[m]) * rh.v[0];
        o[m * out.sz + out.sc] = (pyv[m] - pz[m]) * rh.v[1];
        o[m * out.sz + 2 * out.sc] = (pz[m + 1] - pz[m]) * rh.v[2];
      }
  } else {
    T u[ZPT][3];
#pragma unroll
    for (int m = 0; m < ZPT; m++)
      if (m < nk) {
#pragma unroll
        for (int a = 0; a < 3; a++) u[m][a] = o[m * out.sz + a * out.sc];
      }
#pragma unroll
    for (int m = 0; m < ZPT; m++)
      if (m < nk) {
        o[m * out.sz] = u[m][0] - (pxv[m] - pz[m]) * rh.v[0];
        o[m * out.sz + out.sc] = u[m][1] - (pyv[m] - pz[m]) * rh.v[1];
        o[m * out.sz + 2 * out.sc] = u[m][2] - (pz[m + 1] - pz[m]) * rh.v[2];
      }
  }
}

// ---------------------------------------------------------------------------------------------
// 2-D forms marching RPT rows per thread (the one-cell-per-thread kernels reach about half of the HBM bandwidth
// in 2-D: too few bytes in flight per thread, 4096^2 measured).  All loads of a thread's rows are issued before its
// first store; values shared by consecutive rows (p(j+1), u_y(j-1)) are loaded once.  Same arithmetic, same order.
// Measured at 4096^2 Float32: divergence 0.072 -> 0.041 ms, gradient 0.103 -> 0.059 ms.  (The same treatment of the
// stage kernel, two rows per thread, was slower -- 0.199 -> 0.235 ms -- and is not kept.)
// ---------------------------------------------------------------------------------------------
template <typename T, int MODE, int RPT>
__global__ void k_gradient_y(View<T> out, View<const T> p, Geo gp, Rh<T> rh) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j0 = (blockIdx.y * blockDim.y + threadIdx.y) * RPT;
  if (i >= gp.n[0] || j0 >= gp.n[1]) return;
  int ip = i + 1;
  if (gp.wrap[0] && ip >= gp.n[0]) ip -= gp.n[0];
  const int nr = min(RPT, gp.n[1] - j0);
  T pc[RPT + 1], px[RPT];
#pragma unroll
  for (int m = 0; m <= RPT; m++) {
    int jj = j0 + m;
    if (gp.wrap[1] && jj >= gp.n[1]) jj -= gp.n[1];
    if (m <= nr) pc[m] = __ldg(p.p + (i64)jj * p.sy + i);
    if (m < nr) px[m] = __ldg(p.p + (i64)jj * p.sy + ip);
  }
  T *o = out.p + (i64)j0 * out.sy + i;
  if (MODE == 0) {
#pragma unroll
    for (int m = 0; m < RPT; m++)
      if (m < nr) {
        o[m * out.sy] = (px[m] - pc[m]) * rh.v[0];
        o[m * out.sy + out.sc] = (pc[m + 1] - pc[m]) * rh.v[1];
      }
  } else {
    T u[RPT][2];
#pragma unroll
    for (int m = 0; m < RPT; m++)
      if (m < nr) { u[m][0] = o[m * out.sy]; u[m][1] = o[m * out.sy + out.sc]; }
#pragma unroll
    for (int m = 0; m < RPT; m++)
      if (m < nr) {
        o[m * out.sy] = u[m][0] - (px[m] - pc[m]) * rh.v[0];
        o[m * out.sy + out.sc] = u[m][1] - (pc[m + 1] - pc[m]) * rh.v[1];
      }
  }
}

template <typename T, int RPT>
__global__ void k_divergence_y(View<T> out, View<const T> u, Geo g, Rh<T> rh, T vol, int scale) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j0 = (blockIdx.y * blockDim.y + threadIdx.y) * RPT;
  if (i >= g.n[0] || j0 >= g.n[1]) return;
  int im = i - 1;
  if (g.wrap[0] && im < 0) im += g.n[0];
  const int nr = min(RPT, g.n[1] - j0);
  T ux[RPT], uxm[RPT], uy[RPT + 1];  // uy[m] = u_y(i, j0 + m - 1)
#pragma unroll
  for (int m = 0; m <= RPT; m++) {
    int jj = j0 + m - 1;
    if (g.wrap[1] && jj < 0) jj += g.n[1];
    if (m <= nr) uy[m] = __ldg(u.p + u.sc + (i64)jj * u.sy + i);
    if (m >= 1 && m <= nr) {
      ux[m - 1] = __ldg(u.p + (i64)jj * u.sy + i);
      uxm[m - 1] = __ldg(u.p + (i64)jj * u.sy + im);
    }
  }
#pragma unroll
  for (int m = 0; m < RPT; m++)
    if (m < nr) {
      T d = T(0);
      d += (ux[m] - uxm[m]) * rh.v[0];
      d += (uy[m + 1] - uy[m]) * rh.v[1];
      if (scale) d = d * vol;
      out.p[(i64)(j0 + m) * out.sy + i] = d;
    }
}

// ---------------------------------------------------------------------------------------------
// observers (processors.jl:77-198): fields derived from u, on the reference layout (raw base pointer,
// N = n+2 points per axis incl. ghosts; the ghosts of u must be valid).
// ---------------------------------------------------------------------------------------------
// vorticity!  src/operators.jl:989-1021: over ndrange N-1 (0-based indices 0 .. N-2 on every axis)
//   2-D: w[I]   = (u[I+e1,2] - u[I,2])/h1 - (u[I+e2,1] - u[I,1])/h2
//   3-D: w[I,a] = (u[I+e(a+),a-] - u[I,a-])/h(a+) - (u[I+e(a-),a+] - u[I,a+])/h(a-),  (a,a+,a-) cyclic
template <typename T, int D>
__global__ void k_vorticity(T *__restrict__ w, const T *__restrict__ u, int N0, int N1, int N2, Rh<T> rh) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z;
  if (i >= N0 - 1 || j >= N1 - 1 || (D == 3 && k >= N2 - 1)) return;
  const i64 sy = N0, sz = (i64)N0 * N1, sc = (i64)N0 * N1 * (D == 3 ? N2 : 1);
  const i64 o = i + j * sy + (D == 3 ? k * sz : 0);
  const i64 e[3] = {1, sy, sz};
  if (D == 2) {
    w[o] = (u[sc + o + e[0]] - u[sc + o]) * rh.v[0] - (u[o + e[1]] - u[o]) * rh.v[1];
  } else {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      const int ap = (a + 1) % 3, am = (a + 2) % 3;
      w[a * sc + o] = (u[am * sc + o + e[ap]] - u[am * sc + o]) * rh.v[ap] - (u[ap * sc + o + e[am]] - u[ap * sc + o]) * rh.v[am];
    }
  }
}

// MODE 0: interpolate_u_p!  src/operators.jl:1311-1326:  up[I,a] = (u[I-e(a),a] + u[I,a]) / 2 on Ip
// MODE 1: interpolate_w_p!  src/operators.jl:1338-1372:  2-D: wp[I] = (w[I-e1-e2] + w[I]) / 2;
//                                                        3-D: wp[I,a] = (w[I-e(a+)-e(a-),a] + w[I,a]) / 2 on Ip
template <typename T, int D, int MODE>
__global__ void k_interp_p(View<T> out, View<const T> in, Geo g) {
  INS_CELL_IDX(g)
  const i64 o = k * in.sz + j * in.sy + i;
  const i64 e[3] = {1, in.sy, in.sz};
  const int nc = (MODE == 1 && D == 2) ? 1 : D;
  for (int a = 0; a < nc; a++) {
    i64 back;
    if (MODE == 0) back = e[a];
    else if (D == 2) back = e[0] + e[1];
    else back = e[(a + 1) % 3] + e[(a + 2) % 3];
    out.p[a * out.sc + o] = (in.p[a * in.sc + o - back] + in.p[a * in.sc + o]) * T(0.5);
  }
}

// velocity norm of observefield(:velocitynorm), processors.jl:141-151: sqrt(up1^2 + up2^2 [+ up3^2]) over the whole array
template <typename T, int D>
__global__ void k_velocitynorm(T *__restrict__ out, const T *__restrict__ up, i64 n) {
  const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  T s = up[t] * up[t] + up[n + t] * up[n + t];
  if (D == 3) s = s + up[2 * n + t] * up[2 * n + t];
  out[t] = sqrt(s);
}

// whole-array scale (scalewithvolume! on the reference layout, ghosts included)
template <typename T>
__global__ void k_scale(T *p, i64 n, T f) {
  i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) p[t] = p[t] * f;
}

// spectral Poisson: phat = -phat / ((ax+ay)+az) * invN ; zero mode -> 0  (src/pressure.jl:326-341;
// the 1/N of ldiv! :344 is folded in here).  Layout (kx fastest, ky, kz); kx0/ky0 = global offsets.
template <typename T, typename C, int D>
__global__ void k_poisson_scale(C *phat, const T *__restrict__ ax, const T *__restrict__ ay,
                                const T *__restrict__ az, int nkx, int nky, int nkz, int kx0, T invN) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z;
  if (i >= nkx || j >= nky) return;
  T den = ax[kx0 + i] + ay[j];
  if (D == 3) den = den + az[k];
  const i64 o = ((i64)k * nky + j) * nkx + i;
  C v = phat[o];
  if (kx0 + i == 0 && j == 0 && k == 0) {
    v.x = T(0);
    v.y = T(0);
  } else {
    v.x = (-v.x / den) * invN;
    v.y = (-v.y / den) * invN;
  }
  phat[o] = v;
}

// ---------------------------------------------------------------------------------------------
// reductions: per-block partials with warp shuffles, finished by a second tiny kernel
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}

// partial[block] = sum over the block's cells of  (sum_a u[I,a]^2 + u[I-e_a,a]^2)/4 * vol
// (kinetic_energy! elast! src/operators.jl:1531-1539 + scalewithvolume! + sum); per-cell value in
// T as the reference, accumulation in double.  Also per-component max |u| (get_cfl_timestep!
// src/solver.jl:110-122: min_I h/|u| == h / max_I |u| because fl(h/x) is monotone).
template <typename T, int D>
__global__ void k_reduce_partials(View<const T> u, Geo g, T vol, double *part_ke, double *part_max) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z;
  double ke = 0.0;
  double mx[3] = {0.0, 0.0, 0.0};
  if (i < g.n[0] && j < g.n[1]) {
    Sten<D, 1> s(g, u.sy, u.sz, i, j, k);
    Acc<T, D, 1> U{u.p, u.sc, s};
    T kk = T(0);
#pragma unroll
    for (int a = 0; a < D; a++) {
      const T v = U(a, 0, 0, 0), vm = U(a, -(a == 0), -(a == 1), -(a == 2));
      kk += v * v + vm * vm;
      mx[a] = fabs((double)v);
    }
    kk = kk * T(0.25);
    ke = (double)(kk * vol);
  }
  __shared__ double sh[4][32];
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  const int lane = tid & 31, w = tid >> 5, nw = (blockDim.x * blockDim.y + 31) >> 5;
  ke = warp_sum(ke);
#pragma unroll
  for (int a = 0; a < 3; a++) mx[a] = warp_max(mx[a]);
  if (lane == 0) { sh[0][w] = ke; sh[1][w] = mx[0]; sh[2][w] = mx[1]; sh[3][w] = mx[2]; }
  __syncthreads();
  if (w == 0) {
    ke = lane < nw ? sh[0][lane] : 0.0;
    double m0 = lane < nw ? sh[1][lane] : 0.0, m1 = lane < nw ? sh[2][lane] : 0.0, m2 = lane < nw ? sh[3][lane] : 0.0;
    ke = warp_sum(ke); m0 = warp_max(m0); m1 = warp_max(m1); m2 = warp_max(m2);
    if (lane == 0) {
      const i64 b = ((i64)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
      part_ke[b] = ke;
      part_max[3 * b + 0] = m0; part_max[3 * b + 1] = m1; part_max[3 * b + 2] = m2;
    }
  }
}

// out[0] = sum part_ke ; out[1..3] = max part_max   (single block, fixed order -> deterministic)
static __global__ void k_reduce_final(const double *part_ke, const double *part_max, i64 nb, double *out) {
  double ke = 0.0, m0 = 0.0, m1 = 0.0, m2 = 0.0;
  for (i64 b = threadIdx.x; b < nb; b += blockDim.x) {
    ke += part_ke[b];
    m0 = fmax(m0, part_max[3 * b]); m1 = fmax(m1, part_max[3 * b + 1]); m2 = fmax(m2, part_max[3 * b + 2]);
  }
  __shared__ double sh[4][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  ke = warp_sum(ke); m0 = warp_max(m0); m1 = warp_max(m1); m2 = warp_max(m2);
  if (lane == 0) { sh[0][w] = ke; sh[1][w] = m0; sh[2][w] = m1; sh[3][w] = m2; }
  __syncthreads();
  if (w == 0) {
    ke = lane < nw ? sh[0][lane] : 0.0; m0 = lane < nw ? sh[1][lane] : 0.0;
    m1 = lane < nw ? sh[2][lane] : 0.0; m2 = lane < nw ? sh[3][lane] : 0.0;
    ke = warp_sum(ke); m0 = warp_max(m0); m1 = warp_max(m1); m2 = warp_max(m2);
    if (lane == 0) { out[0] = ke; out[1] = m0; out[2] = m1; out[3] = m2; }
  }
}

// ---------------------------------------------------------------------------------------------
// distributed-FFT pack / unpack (nranks > 1).  Spectrum after the in-slab transform:
//   S[zl][ky][kx], kx in [0,nxh) fastest.  Peer q owns kx in [q*kxc, min((q+1)*kxc, nxh)).
// pack:   B[q][zl][ky][kxq]   = S[zl][ky][q*kxc + kxq]      (block stride = kxc*ny*nzl, dense inside)
// after the all-to-all rank r holds R[q][zl_q][ky][kxl] = all z for its kx chunk: (kxl fastest, ky, z)
// unpack is the inverse of pack on the way back.
// ---------------------------------------------------------------------------------------------
template <typename C>
__global__ void k_fft_pack(C *B, const C *__restrict__ S, int nxh, int ps /* row pitch of S */, i64 rows /* ny*nzl (or nyl in 2-D) */,
                           int kxc, int nranks, int dir /*0: S->B, 1: B->S*/) {
  const int i = blockIdx.y * blockDim.x + threadIdx.x;                  // kx
  const i64 r = (i64)blockIdx.x * blockDim.y + threadIdx.y;              // row
  if (i >= nxh || r >= rows) return;
  const int q = min(i / max(kxc, 1), nranks - 1);
  const int kq = i - q * kxc;
  const int w = q == nranks - 1 ? nxh - q * kxc : kxc;  // width of q's chunk (the last rank also takes the remainder)
  const i64 bo = (i64)q * kxc * rows + r * w + kq;
  if (dir == 0) B[bo] = S[r * ps + i];
  else ((C *)S)[r * ps + i] = B[bo];
}

}  // namespace insb200
