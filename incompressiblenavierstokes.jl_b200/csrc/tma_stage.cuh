// tma_stage.cuh -- the hot kernels of the 3-D RK stage on the PADDED layout (sm_100a):
// TMA (cp.async.bulk.tensor) plane pipeline into shared memory, z-marching 2.5-D tiles.
//
// Padded layout (ctx.cuh): every internal 3-D field is (PX, PY, PZ[, comp]) with real ghost cells
// (2 on the minus side, 3 on the plus side in y and z, >= 4 in x), row pitch PX*sizeof(T) a multiple
// of 128 B and the interior origin 64-B aligned, so a (x, y, 1, comp) box of any tile including its
// halo is ONE bulk tensor copy: no index wrap, no per-thread address arithmetic, no register staging.
//
// k_stage3d_tma: one launch per RK stage (step_explicit_runge_kutta.jl:17-50 up to the Poisson solve)
//   u      = stage velocity.  With PGRAD the projection of the PREVIOUS stage is applied to each plane
//            once it has landed in shared memory: u = u' - grad p (applypressure!, operators.jl:214-233),
//            so the projected intermediate velocity is never written to HBM.
//   k_i    = momentum!(u) [+ s]   (convection+diffusion [+ force] [+ closure term s = div sigma])
//   u'     = ustart + sum_j (dt a_ij) k_j                                  (:35-38)
//   div    = divergence!(u') * Omega   (operators.jl:106-125, 81-95): the Poisson right-hand side
// k_smag3d_tma: s = divoftensor(smagtensor(u)) (operators.jl:1135-1151, 1205-1238, 1294-1305) with the
//   tensor sigma living only in a 3-plane shared-memory ring (the reference stores 9 words/cell).
//
// Pipeline per block (tile TX x TY, z-chunk [kbeg, kend)): NS plane slots; plane m+NS-1 is requested
// by one thread right after the barrier that retires plane m-1; consumers wait on the slot's mbarrier.
#pragma once
#include <cuda.h>

#include "stencil.cuh"

namespace insb200 {

// ---- PTX wrappers -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void *dst, const CUtensorMap *tm, uint64_t *bar, int x, int y, int z, int c) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(
          smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(x), "r"(y), "r"(z), "r"(c), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// ---- plane pipeline -----------------------------------------------------------------------------
// Slots hold planes q = 0, 1, 2, ... (q = z - zfirst).  With PGRAD the slot's barrier also covers the
// pressure plane q+1 (and q for the very first request), which the in-place projection of plane q needs.
template <typename T, int TX, int BYU, bool PGRAD, int NS>
struct PlanePipe {
  static constexpr int V = 16 / (int)sizeof(T);   // elements per 16 B
  static constexpr int XL = V;                    // left x margin of the box (>= 2 halo cells)
  static constexpr int BX = TX + 2 * V;           // box width: i0-XL .. i0+TX+V-1
  static constexpr int BYP = BYU + 1;
  static constexpr int NPS = NS + 1;              // pressure slots (the prologue holds planes 0 .. NS at once)
  static constexpr int UPL = BYU * BX;            // elements of one component plane
  static constexpr int USLOT = ((3 * UPL * (int)sizeof(T) + 127) / 128) * 128 / (int)sizeof(T);
  static constexpr int PSLOT = ((BYP * BX * (int)sizeof(T) + 127) / 128) * 128 / (int)sizeof(T);
  static constexpr int ELEMS = NS * USLOT + (PGRAD ? NPS * PSLOT : 0);
  static constexpr uint32_t UBYTES = 3u * UPL * sizeof(T), PBYTES = (uint32_t)(BYP * BX * sizeof(T));

  T *su, *sp;
  uint64_t *full;
  const CUtensorMap *tmu, *tmp;
  int x0, y0, z0;  // tensor coordinates of the box origin of plane q = 0

  __device__ __forceinline__ void init(T *base, uint64_t *bars, const CUtensorMap *tu, const CUtensorMap *tp, int x0_,
                                       int y0_, int z0_, int tid) {
    su = base;
    sp = base + NS * USLOT;
    full = bars;
    tmu = tu; tmp = tp;
    x0 = x0_; y0 = y0_; z0 = z0_;
    if (tid == 0) {
#pragma unroll
      for (int s = 0; s < NS; s++) mbar_init(&full[s], 1);
      fence_barrier_init();
    }
  }
  __device__ __forceinline__ T *uslot(int q) const { return su + (q % NS) * USLOT; }
  __device__ __forceinline__ T *pslot(int q) const { return sp + (q % NPS) * PSLOT; }

  // one thread: request plane q (all three velocity components; the pressure planes it needs)
  __device__ __forceinline__ void issue(int q) const {
    uint64_t *bar = &full[q % NS];
    fence_proxy_async();  // earlier generic-proxy accesses of the slot are ordered before the bulk write
    if (PGRAD) {
      mbar_expect_tx(bar, UBYTES + (q == 0 ? 2u : 1u) * PBYTES);
      if (q == 0) tma_load_4d(pslot(0), tmp, bar, x0, y0, z0, 0);
      tma_load_4d(pslot(q + 1), tmp, bar, x0, y0, z0 + q + 1, 0);
    } else {
      mbar_expect_tx(bar, UBYTES);
    }
    tma_load_4d(uslot(q), tmu, bar, x0, y0, z0 + q, 0);
  }
  __device__ __forceinline__ void wait(int q) const { mbar_wait(&full[q % NS], (uint32_t)((q / NS) & 1)); }

  // all threads: u(q) = u'(q) - grad p over the whole box (the last column / row of the box is never
  // consumed, so the neighbour reads p[l+1], p[l+BX] may run into the next row / the extra p row)
  __device__ __forceinline__ void project(int q, int tid, int nt, const T (&rh)[3]) const {
    if (!PGRAD) return;
    T *u = uslot(q);
    const T *p = pslot(q), *pn = pslot(q + 1);
    for (int l = tid; l < UPL; l += nt) {
      const T p0 = p[l];
      u[l] = u[l] - (p[l + 1] - p0) * rh[0];
      u[UPL + l] = u[UPL + l] - (p[l + BX] - p0) * rh[1];
      u[2 * UPL + l] = u[2 * UPL + l] - (pn[l] - p0) * rh[2];
    }
  }
};

// velocity accessor on three resident plane slots
template <typename T, int BX, int UPL>
struct PlaneAcc {
  const T *pm, *p0, *pp;  // planes k-1, k, k+1
  int c0;                 // offset of the cell inside a component plane
  __device__ __forceinline__ T operator()(int c, int dx, int dy, int dz) const {
    const T *b = dz < 0 ? pm : (dz > 0 ? pp : p0);
    return b[c * UPL + c0 + dy * BX + dx];
  }
};

// one component `a` of the stage at a cell: f = momentum!(u)[a] (+ force + s), return u' (the combination).
//   og = global element offset of the cell in the padded buffers (without the component stride)
template <typename T, bool SADD, class ACC>
__device__ __forceinline__ T stage_comp_p(const StageArgs<T> &A, const ACC &U, const int a, const i64 og, const T visc,
                                          const T (&rh)[3], T &f_out) {
  T f = convdiff_comp<T, 3, true, true>(U, a, visc, rh);
  const i64 o = og + a * A.sc;
  if (A.force) f = f + __ldg(A.force + o);
  if (SADD) f = f + __ldg(A.sadd + o);
  f_out = f;
  T acc = A.u0_is_u ? U(a, 0, 0, 0) : __ldg(A.u0 + o);
  for (int q = 0; q < A.nk; q++) acc = acc + A.coef[q] * __ldg(A.kin[q] + o);
  return acc + A.coef_self * f;
}

template <typename T, bool PGRAD, bool SADD, int TX, int TY, int NS>
struct StageCfg {
  typedef PlanePipe<T, TX, TY + 3, PGRAD, NS> Pipe;
  static constexpr int NT = TX * TY;
  static constexpr size_t SMEM = 128 + (size_t)Pipe::ELEMS * sizeof(T) + (size_t)(2 * TY * (TX + 1) + 2 * (TY + 1) * TX) * sizeof(T) + 8 * NS + 16;
};

template <typename T, bool PGRAD, bool SADD, int TX, int TY, int NS>
__global__ void __launch_bounds__(TX *TY, (sizeof(T) == 4 ? 2 : 1))
    k_stage3d_tma(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_p, StageArgs<T> A, Geo g,
                  T visc, Rh<T> rh, T vol, int zc, int ox, int oy, int oz) {
  typedef StageCfg<T, PGRAD, SADD, TX, TY, NS> Cfg;
  typedef typename Cfg::Pipe Pipe;
  constexpr int BX = Pipe::BX, UPL = Pipe::UPL, NT = Cfg::NT, XL = Pipe::XL;
  extern __shared__ unsigned char smem_raw[];
  unsigned char *sb = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  T *sbase = reinterpret_cast<T *>(sb);
  T *sdx = sbase + Pipe::ELEMS;                       // [2][TY][TX+1]  u'_x incl. column -1
  T *sdy = sdx + 2 * TY * (TX + 1);                   // [2][TY+1][TX]  u'_y incl. row -1
  uint64_t *bars = reinterpret_cast<uint64_t *>((reinterpret_cast<uintptr_t>(sdy + 2 * (TY + 1) * TX) + 7) & ~(uintptr_t)7);

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TX + tx;
  const int i0 = blockIdx.x * TX, j0 = blockIdx.y * TY;
  const int kbeg = blockIdx.z * zc, kend = min(kbeg + zc, g.n[2]);
  const int n0 = g.n[0], n1 = g.n[1], n2 = g.n[2];
  const i64 sy = A.sy, sz = A.sz;

  // plane q <-> z = kbeg - 2 + q
  Pipe pipe;
  pipe.init(sbase, bars, &tm_u, &tm_p, ox + i0 - XL, oy + j0 - 2, oz + kbeg - 2, tid);
  __syncthreads();
  const int qlast = kend - kbeg + 2;  // plane kend
  if (tid == 0) {
#pragma unroll
    for (int q = 0; q < NS; q++)
      if (q <= qlast) pipe.issue(q);
  }
  // planes q = 0 .. 3 (z = kbeg-2 .. kbeg+1) must be resident and projected before the first iteration
#pragma unroll
  for (int q = 0; q < 4; q++) {
    pipe.wait(q);
    pipe.project(q, tid, NT, rh.v);
  }
  __syncthreads();

  const int gi = i0 + tx, gj = j0 + ty;
  const i64 ocell = (i64)gj * sy + gi;
  // the extra point of some threads on the minus halo: warp 1 (first TY lanes): (-1, l) component x;
  // warp 2: (l, -1) component y  (different warps / schedulers than each other)
  const bool hx = tid >= 32 && tid < 32 + TY, hy = tid >= 64 && tid < 64 + TX;
  const int hlx = hx ? -1 : tid - 64, hly = hx ? tid - 32 : -1;
  const int hgi = wrapi(i0 + hlx, n0, 1), hgj = wrapi(j0 + hly, n1, 1);
  const i64 ohalo = (i64)hgj * sy + hgi;
  const int c_own = (ty + 2) * BX + tx + XL;
  const int c_halo = (hly + 2) * BX + hlx + XL;

  // u'_z of plane kbeg-1 (for the z-difference of the divergence on the first plane of the chunk)
  T uz_prev = T(0);
  const bool have_prev = kbeg > 0 || g.wrap[2];  // otherwise plane 0 is completed by k_divergence_planes
  if (A.div && have_prev) {
    const int kz = wrapi(kbeg - 1, n2, g.wrap[2]);
    PlaneAcc<T, BX, UPL> U{pipe.uslot(0), pipe.uslot(1), pipe.uslot(2), c_own};
    T f;
    uz_prev = stage_comp_p<T, SADD>(A, U, 2, (i64)kz * sz + ocell, visc, rh.v, f);
  }
  __syncthreads();  // plane 0 is retired: its slot takes plane NS
  if (tid == 0 && NS <= qlast) pipe.issue(NS);

  for (int k = kbeg; k < kend; k++) {
    const int q = k - kbeg + 2;  // plane number of k
    PlaneAcc<T, BX, UPL> U{pipe.uslot(q - 1), pipe.uslot(q), pipe.uslot(q + 1), c_own};
    const i64 ozk = (i64)k * sz;
    T un[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
      T f;
      un[a] = stage_comp_p<T, SADD>(A, U, a, ozk + ocell, visc, rh.v, f);
      const i64 og = ozk + ocell + a * A.sc;
      if (A.kout) A.kout[og] = f;
      A.unew[og] = un[a];
    }
    if (A.div) {
      T *bx = sdx + (k & 1) * TY * (TX + 1), *by = sdy + (k & 1) * (TY + 1) * TX;
      bx[ty * (TX + 1) + tx + 1] = un[0];
      by[(ty + 1) * TX + tx] = un[1];
      if (hx || hy) {
        PlaneAcc<T, BX, UPL> Uh{pipe.uslot(q - 1), pipe.uslot(q), pipe.uslot(q + 1), c_halo};
        T f;
        if (hx) bx[hly * (TX + 1)] = stage_comp_p<T, SADD>(A, Uh, 0, ozk + ohalo, visc, rh.v, f);
        else by[hlx] = stage_comp_p<T, SADD>(A, Uh, 1, ozk + ohalo, visc, rh.v, f);
      }
    }
    // plane k+2 for the next iteration
    if (k + 1 < kend) {
      pipe.wait(q + 2);
      pipe.project(q + 2, tid, NT, rh.v);
    }
    __syncthreads();
    // slot of plane k-1 is free now: request plane k+NS-1
    if (tid == 0 && q - 1 + NS <= qlast) pipe.issue(q - 1 + NS);
    if (A.div && (k > kbeg || have_prev)) {
      const T *bx = sdx + (k & 1) * TY * (TX + 1), *by = sdy + (k & 1) * (TY + 1) * TX;
      T d = T(0);
      d += (un[0] - bx[ty * (TX + 1) + tx]) * rh.v[0];
      d += (un[1] - by[ty * TX + tx]) * rh.v[1];
      d += (un[2] - uz_prev) * rh.v[2];
      A.div[ozk + ocell] = d * vol;
    }
    uz_prev = un[2];
  }
}

// -------------------------------------------------------------------------------------------------
// k_smag3d_tma: s = divoftensor!(apply_bc_p!(smagtensor!(u)))  (operators.jl:1294-1305) in one pass.
// Per z-plane the tile's sigma (6 symmetric components, p-cells x in [-1,TX], y in [-1,TY]) is computed
// from three resident velocity planes into a 3-slot shared-memory ring; s(k) is then the divergence of
// the ring's planes k-1, k, k+1.  Plane numbering: velocity q = 0 <-> z = kbeg-2, sigma r = 0 <-> z = kbeg-1.
// -------------------------------------------------------------------------------------------------
template <typename T, int SX, int SPL>
struct SigmaAcc {
  const T *sm, *s0, *sp;  // sigma planes k-1, k, k+1
  int c0;
  __device__ __forceinline__ T operator()(int c, int dx, int dy, int dz) const {
    const T *b = dz < 0 ? sm : (dz > 0 ? sp : s0);
    return b[c * SPL + c0 + dy * SX + dx];
  }
};

template <typename T, bool PGRAD, int TX, int TY, int NS>
struct SmagCfg {
  typedef PlanePipe<T, TX, TY + 4, PGRAD, NS> Pipe;
  static constexpr int NT = TX * TY;
  static constexpr int SX = TX + 2, SY = TY + 2, SPL = SX * SY, SSLOT = 6 * SPL;
  static constexpr size_t SMEM = 128 + (size_t)(Pipe::ELEMS + 3 * SSLOT) * sizeof(T) + 8 * NS + 16;
};

template <typename T, bool PGRAD, int TX, int TY, int NS>
__global__ void __launch_bounds__(TX *TY, (sizeof(T) == 4 ? 2 : 1))
    k_smag3d_tma(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_p, T *__restrict__ sout,
                 i64 sy, i64 sz, i64 sc, Geo g, T c_eddy, Rh<T> rh, int zc, int ox, int oy, int oz) {
  typedef SmagCfg<T, PGRAD, TX, TY, NS> Cfg;
  typedef typename Cfg::Pipe Pipe;
  constexpr int BX = Pipe::BX, UPL = Pipe::UPL, NT = Cfg::NT, XL = Pipe::XL;
  constexpr int SX = Cfg::SX, SPL = Cfg::SPL, SSLOT = Cfg::SSLOT;
  extern __shared__ unsigned char smem_raw[];
  unsigned char *sb = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  T *sbase = reinterpret_cast<T *>(sb);
  T *ssig = sbase + Pipe::ELEMS;  // [3][6][SY][SX]
  uint64_t *bars = reinterpret_cast<uint64_t *>((reinterpret_cast<uintptr_t>(ssig + 3 * SSLOT) + 7) & ~(uintptr_t)7);

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TX + tx;
  const int i0 = blockIdx.x * TX, j0 = blockIdx.y * TY;
  const int kbeg = blockIdx.z * zc, kend = min(kbeg + zc, g.n[2]);

  Pipe pipe;
  pipe.init(sbase, bars, &tm_u, &tm_p, ox + i0 - XL, oy + j0 - 2, oz + kbeg - 2, tid);
  __syncthreads();
  const int qlast = kend - kbeg + 3;  // velocity plane z = kend+1
  if (tid == 0) {
#pragma unroll
    for (int q = 0; q < NS; q++)
      if (q <= qlast) pipe.issue(q);
  }
#pragma unroll
  for (int q = 0; q < 4; q++) {
    pipe.wait(q);
    pipe.project(q, tid, NT, rh.v);
  }
  __syncthreads();

  // sigma plane r from velocity planes r, r+1, r+2 over the (TX+2) x (TY+2) p-cells of the tile
  auto sigma_plane = [&](int r) {
    T *dst = ssig + (r % 3) * SSLOT;
    for (int l = tid; l < SPL; l += NT) {
      const int ly = l / SX, lx = l - ly * SX;  // p-cell (lx-1, ly-1) of the tile
      PlaneAcc<T, BX, UPL> U{pipe.uslot(r), pipe.uslot(r + 1), pipe.uslot(r + 2), (ly + 1) * BX + lx - 1 + XL};
      T sig[6];
      smag_sigma<T, 3>(U, c_eddy, rh.v, sig);
#pragma unroll
      for (int c = 0; c < 6; c++) dst[c * SPL + l] = sig[c];
    }
  };
  sigma_plane(0);
  sigma_plane(1);
  __syncthreads();  // velocity planes 0 and 1 are retired
  if (tid == 0) {
    if (NS <= qlast) pipe.issue(NS);
    if (NS + 1 <= qlast) pipe.issue(NS + 1);
  }

  const i64 ocell = (i64)(j0 + ty) * sy + i0 + tx;
  const int c_sig = (ty + 1) * SX + tx + 1;
  for (int k = kbeg; k < kend; k++) {
    const int r = k - kbeg + 1;  // sigma plane number of k
    // sigma(k+1) needs velocity planes r+1, r+2, r+3 (z = k .. k+2)
    pipe.wait(r + 3);
    pipe.project(r + 3, tid, NT, rh.v);
    __syncthreads();
    sigma_plane(r + 1);
    __syncthreads();
    // velocity plane r+1 (z = k) is retired: its slot takes plane r+1+NS
    if (tid == 0 && r + 1 + NS <= qlast) pipe.issue(r + 1 + NS);
    SigmaAcc<T, SX, SPL> Sg{ssig + ((r - 1) % 3) * SSLOT, ssig + (r % 3) * SSLOT, ssig + ((r + 1) % 3) * SSLOT, c_sig};
    const i64 og = (i64)k * sz + ocell;
#pragma unroll
    for (int a = 0; a < 3; a++) sout[og + a * sc] = divsigma_comp<T, 3>(Sg, a, rh.v);
  }
}

// -------------------------------------------------------------------------------------------------
// periodic ghost fill of a padded field: every ghost point within (gm minus, gp plus) layers copies its
// wrapped interior image (== the reference's axis-by-axis slab copies, boundary_conditions.jl:276-318).
// Regions (linear thread index): A = z ghost planes (skipped when fill_z == 0: filled over NCCL),
// B = y ghost rows of the interior planes, C = x ghost columns of the interior rows.
// -------------------------------------------------------------------------------------------------
template <typename T>
__global__ void k_fill_ghosts(T *__restrict__ org, i64 sy, i64 sz, i64 sc, int n0, int n1, int n2, int gm, int gp,
                              int fill_z, i64 cntA, i64 cntB, i64 cntC) {
  const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  T *f = org + (i64)blockIdx.y * sc;
  const int gw = gm + gp;
  const int ex = n0 + gw, ey = n1 + gw;
  int x, y, z;
  if (t < cntA) {
    // z ghost planes over the extended x-y range
    const int zz = (int)(t / ((i64)ex * ey));
    const int r = (int)(t - (i64)zz * ex * ey);
    y = r / ex - gm;
    x = r - (r / ex) * ex - gm;
    z = zz < gm ? zz - gm : n2 + (zz - gm);
  } else if (t < cntA + cntB) {
    const i64 u = t - cntA;
    z = (int)(u / ((i64)gw * ex));
    const int r = (int)(u - (i64)z * gw * ex);
    const int yy = r / ex;
    x = r - yy * ex - gm;
    y = yy < gm ? yy - gm : n1 + (yy - gm);
  } else if (t < cntA + cntB + cntC) {
    const i64 u = t - cntA - cntB;
    z = (int)(u / ((i64)n1 * gw));
    const int r = (int)(u - (i64)z * n1 * gw);
    y = r / gw;
    const int xx = r - y * gw;
    x = xx < gm ? xx - gm : n0 + (xx - gm);
  } else {
    return;
  }
  int xs = x, ys = y, zs = z;
  if (xs < 0) xs += n0; else if (xs >= n0) xs -= n0;
  if (ys < 0) ys += n1; else if (ys >= n1) ys -= n1;
  if (fill_z) { if (zs < 0) zs += n2; else if (zs >= n2) zs -= n2; }
  f[(i64)z * sz + (i64)y * sy + x] = f[(i64)zs * sz + (i64)ys * sy + xs];
}

}  // namespace insb200
