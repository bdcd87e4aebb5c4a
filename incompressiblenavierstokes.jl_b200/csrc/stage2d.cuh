// stage2d.cuh -- fused RK stage kernel of the 2-D path (compact layout): what k_stage3d_tma is to 3-D.  One rank: rows
// wrap by index; slabs (nranks > 1, g.wrap[1] == 0): the two ghost rows on each side of the slab are read instead
// (u: 2 below / 1 above, p: 2 / 2) and the divergence of the slab's FIRST row -- which needs u'_y of the neighbour rank's
// last row -- is completed by k_divergence_row0 after the halo exchange of u'.
//
// One launch per stage does   [u = u' - grad p of the previous stage, on load]  ->  k_i = momentum!(u) [+ force]  ->
// u' = ustart + sum_j (dt a_ij) k_j  ->  div = divergence!(u') * Omega   (step_explicit_runge_kutta.jl:17-50 up to the
// Poisson solve; pressure.jl:73-74, 81 of the previous projection), i.e. the stage kernel, the divergence kernel and the
// gradient kernel of the unfused path in one pass: 12 instead of 18 words per cell and stage.
//
// A 2-D stencil needs no shared memory: a WARP owns 32 consecutive columns and marches down the rows of its chunk,
// every lane keeps the rows j-1, j, j+1 of its own column in registers, and the x-neighbours come from the adjacent
// lanes by warp shuffle.  Lanes 0, 1 and 31 are halo lanes (lane 0 and 31 only supply neighbours, lane 1 also forms
// u'_x for the divergence of lane 2), lanes 2..30 store: 29 output columns per warp, no barrier, no bank conflicts.
// The row j_beg - 1 is run as a pre-iteration that only forms u'_y for the divergence of the chunk's first row.
// Arithmetic: convdiff_comp and the projection / divergence expressions of the unfused kernels, same order:
// bit-identical results (tests/test_gpu_stepping.py).
#pragma once
#include "stencil.cuh"
#include "tiled.cuh"

namespace insb200 {

constexpr int kS2Cols = 29;   // output columns per warp
constexpr int kS2Warps = 4;   // warps per block, side by side in x

// velocity accessor on registers: own column on rows -1, 0, +1 plus the six neighbour values the 2-D stencil reads
template <typename T>
struct RowAcc {
  T xm, xc, xp, ym, yc, yp;  // u_x, u_y of the own column on rows j-1, j, j+1
  T xc_p, xc_m, xp_m;        // u_x(+1, 0), u_x(-1, 0), u_x(-1, +1)
  T yc_p, yc_m, ym_p;        // u_y(+1, 0), u_y(-1, 0), u_y(+1, -1)
  __device__ __forceinline__ T operator()(int c, int dx, int dy, int) const {
    if (c == 0) {
      if (dx == 0) return dy < 0 ? xm : (dy > 0 ? xp : xc);
      if (dx > 0) return xc_p;
      return dy > 0 ? xp_m : xc_m;
    }
    if (dx == 0) return dy < 0 ? ym : (dy > 0 ? yp : yc);
    if (dx < 0) return yc_m;
    return dy < 0 ? ym_p : yc_p;
  }
};

template <typename T, bool PGRAD>
__global__ void __launch_bounds__(32 * kS2Warps)
    k_stage2d_warp(StageArgs<T> A, Geo g, T visc, Rh<T> rh, T vol, int rc) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, wx = blockIdx.x * kS2Warps + (threadIdx.x >> 5);
  const int n0 = g.n[0], n1 = g.n[1];
  const int c0 = wx * kS2Cols;
  if (c0 >= n0) return;  // whole warp
  const int col = c0 - 2 + lane;
  const int cw = ((col % n0) + n0) % n0;       // halo lanes wrap around the periodic box (any n0 >= 2)
  const int cwp = cw + 1 == n0 ? 0 : cw + 1;   // lane 31: column of its extra pressure value
  const bool store = lane >= 2 && lane <= 30 && col < n0;
  const bool need = lane >= 1 && lane <= 30;   // lanes whose u' enters a stored value
  const int jbeg = blockIdx.y * rc, jend = min(jbeg + rc, n1);
  const int wy = g.wrap[1];                     // 0: ghost rows are addressable (slab of a decomposed grid)
  const bool have_prev = jbeg > 0 || wy != 0;   // otherwise row 0's divergence is completed by k_divergence_row0
  const i64 sy = A.sy, sc = A.sc;
  const bool u0_is_u = A.u0 == A.u;

  // raw row r of the stage input: u' (and p for the projection)
  auto row = [&](int r) -> i64 { return (i64)wrapi(r, n1, wy) * sy + cw; };
  T pn = T(0);  // p(col, r) of the next row to bring in (carried: the upper p of one row is the lower p of the next)
  auto load_p = [&](int r) -> T { return PGRAD ? __ldg(A.pin + row(r)) : T(0); };
  auto load_pe = [&](int r) -> T { return (PGRAD && lane == 31) ? __ldg(A.pin + (i64)wrapi(r, n1, wy) * sy + cwp) : T(0); };
  // projected row from raw values: u_x - (p(+1,0) - p) / h_x,  u_y - (p(0,+1) - p) / h_y   (operators.jl:225-233)
  auto project = [&](T &ux, T &uy, T p0, T p1, T pe) {
    if (!PGRAD) return;
    T px = __shfl_down_sync(FULL, p0, 1);
    if (lane == 31) px = pe;
    ux = ux - (px - p0) * rh.v[0];
    uy = uy - (p1 - p0) * rh.v[1];
  };

  constexpr int NKP = 3;  // earlier k_j prefetched with the other operands (RK44's last row has 3)
  const int nk = A.nk;
  T n_u0[2], n_f[2], n_k[2][NKP];
  auto fetch_operands = [&](int r) {  // operands of row r, consumed one iteration later
    const i64 o1 = row(r);
#pragma unroll
    for (int a = 0; a < 2; a++) {
      n_u0[a] = (need && !u0_is_u) ? __ldg(A.u0 + a * sc + o1) : T(0);
      n_f[a] = (need && A.force) ? __ldg(A.force + a * sc + o1) : T(0);
#pragma unroll
      for (int q = 0; q < NKP; q++) n_k[a][q] = (need && q < nk) ? __ldg(A.kin[q] + a * sc + o1) : T(0);
    }
  };

  RowAcc<T> U;
  T rx, ry, rp1, re;  // RAW row j+1 of the coming iteration (loaded one iteration ahead), its upper p and lane 31's extra p
  // prologue: rows jbeg-2, jbeg-1 projected into (c, p) -- the first iteration rotates them into (m, c) --,
  // raw row jbeg; pn = p(jbeg)
  {
    const int r0 = jbeg - 2;
    T x0 = __ldg(A.u + row(r0)), y0 = __ldg(A.u + sc + row(r0));
    T x1 = __ldg(A.u + row(r0 + 1)), y1 = __ldg(A.u + sc + row(r0 + 1));
    rx = __ldg(A.u + row(r0 + 2)); ry = __ldg(A.u + sc + row(r0 + 2));
    const T p0 = load_p(r0), p1 = load_p(r0 + 1), p2 = load_p(r0 + 2);
    rp1 = load_p(r0 + 3);
    const T e0 = load_pe(r0), e1 = load_pe(r0 + 1);
    re = load_pe(r0 + 2);
    project(x0, y0, p0, p1, e0);
    project(x1, y1, p1, p2, e1);
    U.xc = x0; U.yc = y0; U.xp = x1; U.yp = y1;  // rotated into (m, c) by the first iteration below
    U.xm = x0; U.ym = y0;
    pn = p2;
    U.yc_p = __shfl_down_sync(FULL, y0, 1);
    U.xp_m = __shfl_up_sync(FULL, x1, 1);
    fetch_operands(jbeg - 1);
  }
  T uy_prev = T(0);  // u'_y of the row above (divergence)
  for (int j = jbeg - 1; j < jend; j++) {
    const bool pre = j < jbeg;
    // rotate: (c, p) -> (m, c); the raw row loaded during the previous iteration is projected into p
    U.ym_p = U.yc_p;  // u_y(+1, -1) of this row = u_y(+1, 0) of the previous one
    U.xc_m = U.xp_m;  // u_x(-1, 0) of this row = u_x(-1, +1) of the previous one
    U.xm = U.xc; U.ym = U.yc;
    U.xc = U.xp; U.yc = U.yp;
    project(rx, ry, pn, rp1, re);
    U.xp = rx; U.yp = ry;
    pn = rp1;
    // operands of this row (requested one iteration ago)
    T l_u0[2], l_f[2], l_k[2][NKP];
#pragma unroll
    for (int a = 0; a < 2; a++) {
      l_u0[a] = n_u0[a]; l_f[a] = n_f[a];
#pragma unroll
      for (int q = 0; q < NKP; q++) l_k[a][q] = n_k[a][q];
    }
    const i64 o = row(j);
    // requests for the next iteration: raw row j+2 (+ its upper p), operands of row j+1
    if (j + 1 < jend) {
      rx = __ldg(A.u + row(j + 2)); ry = __ldg(A.u + sc + row(j + 2));
      rp1 = load_p(j + 3); re = load_pe(j + 2);
      fetch_operands(j + 1);
    }
    // neighbour values of this row set
    U.xc_p = __shfl_down_sync(FULL, U.xc, 1);
    U.yc_p = __shfl_down_sync(FULL, U.yc, 1);
    U.yc_m = __shfl_up_sync(FULL, U.yc, 1);
    U.xp_m = __shfl_up_sync(FULL, U.xp, 1);
    T un[2];
#pragma unroll
    for (int a = 0; a < 2; a++) {
      if (pre && a == 0) { un[0] = T(0); continue; }
      T f = convdiff_comp<T, 2, true, true>(U, a, visc, rh.v);
      if (A.force) f = f + l_f[a];
      T acc = u0_is_u ? (a == 0 ? U.xc : U.yc) : l_u0[a];
#pragma unroll
      for (int q = 0; q < NKP; q++)
        if (q < nk) acc = acc + A.coef[q] * l_k[a][q];
      for (int q = NKP; q < nk; q++) acc = acc + A.coef[q] * (need ? __ldg(A.kin[q] + a * sc + o) : T(0));
      un[a] = acc + A.coef_self * f;
      if (store && !pre) {
        if (A.kout) A.kout[a * sc + o] = f;
        A.unew[a * sc + o] = un[a];
      }
    }
    const T unx_m = __shfl_up_sync(FULL, un[0], 1);
    if (store && !pre && (j > jbeg || have_prev)) {
      T d = T(0);
      d += (un[0] - unx_m) * rh.v[0];
      d += (un[1] - uy_prev) * rh.v[1];
      A.div[o] = d * vol;
    }
    uy_prev = un[1];
  }
}

// divergence(u') * Omega on row 0 of a slab (2-D, nranks > 1): the fused kernel cannot form u'_y of the neighbour rank's
// last row; after the halo exchange it sits in ghost row -1 (operators.jl:106-125, 81-95, same expression order)
template <typename T>
__global__ void k_divergence_row0(T *__restrict__ out, const T *__restrict__ u, i64 sy, i64 sc, int n0, Rh<T> rh, T vol) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n0) return;
  const int im = i == 0 ? n0 - 1 : i - 1;
  T d = T(0);
  d += (u[i] - u[im]) * rh.v[0];
  d += (u[sc + i] - u[sc + i - sy]) * rh.v[1];
  out[i] = d * vol;
}

}  // namespace insb200
