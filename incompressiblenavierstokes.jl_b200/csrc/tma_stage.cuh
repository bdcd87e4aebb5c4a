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
// Wait for phase `parity` of an mbarrier.  try_wait suspends the thread for a hardware-chosen time slice before it
// returns false, so the loop is not a hot spin.  -DINSB200_DEBUG bounds it: a wrong expect_tx byte count (the classic
// TMA bug) traps after about a second with the barrier's address instead of hanging the SM until the watchdog fires.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
#ifdef INSB200_DEBUG
  const long long t0 = clock64();
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
        "selp.u32 %0, 1, 0, P1;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (!done && clock64() - t0 > 2000000000LL) {
      printf("insb200: mbarrier %p (block %d,%d,%d thread %d) never completed phase %u -- wrong expect_tx?\n", (void *)bar, blockIdx.x,
             blockIdx.y, blockIdx.z, threadIdx.x + threadIdx.y * blockDim.x, parity);
      __trap();
    }
  }
#else
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
#endif
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
// NS velocity slots hold planes q-1, q, q+1 (consumed), q+2 (landing) [and q+3 in flight, NS = 5]; planes are numbered
// q = 0, 1, 2, ... from the first plane of the block.  With PGRAD the slot's barrier also covers the
// pressure plane q+1 (and q for the very first request), which the in-place projection of plane q needs;
// NS+1 pressure slots because the prologue holds planes 0..NS at once.  All shared-memory addressing is
// by element offset from one base pointer, and the slot offsets are rotated in registers (no modulo in
// the plane loop).
template <typename T, int TX, int BYU, bool PGRAD, int NS_, int NPS_ = NS_ + 1>
struct PlanePipe {
  static constexpr int NS = NS_;
  static constexpr int V = 16 / (int)sizeof(T);   // elements per 16 B
  static constexpr int XL = V;                    // left x margin of the box (>= 2 halo cells)
  static constexpr int BX = TX + 2 * V;           // box width: i0-XL .. i0+TX+V-1
  static constexpr int BYP = BYU + 1;
  static constexpr int NPS = NPS_;
  static constexpr int UPL = BYU * BX;            // elements of one component plane
  static constexpr int USLOT = ((3 * UPL * (int)sizeof(T) + 127) / 128) * 128 / (int)sizeof(T);
  static constexpr int PSLOT = ((BYP * BX * (int)sizeof(T) + 127) / 128) * 128 / (int)sizeof(T);
  static constexpr int POFF = NS * USLOT;         // pressure slots follow the velocity slots
  static constexpr int ELEMS = NS * USLOT + (PGRAD ? NPS * PSLOT : 0);
  static constexpr uint32_t UBYTES = 3u * UPL * sizeof(T), PBYTES = (uint32_t)(BYP * BX * sizeof(T));

  T *s;            // base of the slots
  uint64_t *full;  // NS barriers
  const CUtensorMap *tmu, *tmp;
  int x0, y0, z0;  // tensor coordinates of the box origin of plane q = 0

  __device__ __forceinline__ void init(T *base, uint64_t *bars, const CUtensorMap *tu, const CUtensorMap *tp, int x0_,
                                       int y0_, int z0_, int tid) {
    s = base;
    full = bars;
    tmu = tu; tmp = tp;
    x0 = x0_; y0 = y0_; z0 = z0_;
    if (tid == 0) {
#pragma unroll
      for (int i = 0; i < NS; i++) mbar_init(&full[i], 1);
      fence_barrier_init();
    }
  }
  static __device__ __forceinline__ int uoff(int q) { return (q % NS) * USLOT; }
  static __device__ __forceinline__ int poff(int q) { return POFF + (q % NPS) * PSLOT; }

  // one thread: request plane q (all three velocity components; the pressure planes it needs)
  __device__ __forceinline__ void issue(int q) const {
    uint64_t *bar = &full[q % NS];
    fence_proxy_async();  // earlier generic-proxy accesses of the slot are ordered before the bulk write
    if (PGRAD) {
      mbar_expect_tx(bar, UBYTES + (q == 0 ? 2u : 1u) * PBYTES);
      if (q == 0) tma_load_4d(s + poff(0), tmp, bar, x0, y0, z0, 0);
      tma_load_4d(s + poff(q + 1), tmp, bar, x0, y0, z0 + q + 1, 0);
    } else {
      mbar_expect_tx(bar, UBYTES);
    }
    tma_load_4d(s + uoff(q), tmu, bar, x0, y0, z0 + q, 0);
  }
  __device__ __forceinline__ void wait(int q) const { mbar_wait(&full[q % NS], (uint32_t)((q / NS) & 1)); }
  __device__ __forceinline__ void wait_slot(int slot, uint32_t parity) const { mbar_wait(&full[slot], parity); }

  // all threads: u(q) = u'(q) - grad p over the whole box (the last column / row of the box is never
  // consumed, so the neighbour reads p[l+1], p[l+BX] may run into the next row / the extra p row).
  // uo = velocity slot offset, po / pn = slot offsets of pressure planes q and q+1.
  __device__ __forceinline__ void project(int uo, int po, int pn, int tid, int nt, const T (&rh)[3]) const {
    if (!PGRAD) return;
    T *u = s + uo;
    const T *p = s + po, *p1 = s + pn;
    for (int l = tid; l < UPL; l += nt) {
      const T p0 = p[l];
      u[l] = u[l] - (p[l + 1] - p0) * rh[0];
      u[UPL + l] = u[UPL + l] - (p[l + BX] - p0) * rh[1];
      u[2 * UPL + l] = u[2 * UPL + l] - (p1[l] - p0) * rh[2];
    }
  }
  __device__ __forceinline__ void project_q(int q, int tid, int nt, const T (&rh)[3]) const {
    project(uoff(q), poff(q), poff(q + 1), tid, nt, rh);
  }
};

// velocity accessor on three resident plane slots: element offsets (slot + own cell) from one base
template <typename T, int BX, int UPL>
struct PlaneAcc {
  const T *s;
  int om, o0, op;  // planes k-1, k, k+1
  __device__ __forceinline__ T operator()(int c, int dx, int dy, int dz) const {
    const int o = dz < 0 ? om : (dz > 0 ? op : o0);
    return s[o + c * UPL + dy * BX + dx];
  }
};

// Arguments of k_stage3d_tma.  Every global operand is given as one base pointer PER COMPONENT (interior
// origin of that component), so an access is `base[a] + byte offset of the cell`: the bases live in the
// constant bank and the only per-thread address state is one 64-bit byte offset advanced by one plane
// per iteration.
template <typename T>
struct TmaStageArgs {
  const T *u0[3];     // ustart
  const T *sadd[3];   // closure term s (SADD)
  const T *force[3];  // stored body force (nullable: force[0] == nullptr)
  T *kout[3];         // k_i (GEN: nullable)
  T *unew[3];         // u'
  T *div;             // divergence(u') * Omega
  const T *kin[INSB200_MAX_STAGES][3];  // GEN: earlier stage derivatives k_j, per component
  T coef[INSB200_MAX_STAGES];
  int nk;
  T coef_self;
  i64 sy, sz, sc;     // strides in elements
};

template <typename T, bool PGRAD, int TX, int TY, int NS>
struct StageCfg {
  typedef PlanePipe<T, TX, TY + 3, PGRAD, NS> Pipe;
  static constexpr int NT = TX * TY;
  static constexpr size_t SMEM = 128 + (size_t)Pipe::ELEMS * sizeof(T) + (size_t)(2 * TY * (TX + 1) + 2 * (TY + 1) * TX) * sizeof(T) + 8 * NS + 16;
};

template <typename T>
__device__ __forceinline__ T ld_off(const T *base, size_t ob) {
  return __ldg(reinterpret_cast<const T *>(reinterpret_cast<const char *>(base) + ob));
}
template <typename T>
__device__ __forceinline__ void st_off(T *base, size_t ob, T v) {
  *reinterpret_cast<T *>(reinterpret_cast<char *>(base) + ob) = v;
}

// PGRAD: projection of the previous stage applied on load; !PGRAD: first stage, ustart is the tile itself.
// SADD: closure term added.  GEN: general tableau row (runtime list of earlier k_j, k_i optional);
// !GEN: the common row with a single entry (k_i stored, no earlier k_j) -- RK44 stages 1..3.
template <typename T, bool PGRAD, bool SADD, bool GEN, int TX, int TY, int NS>
__global__ void __launch_bounds__(TX *TY, (sizeof(T) == 4 ? 2 : 1))
    k_stage3d_tma(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_p,
                  const __grid_constant__ TmaStageArgs<T> A, Geo g, T visc, Rh<T> rh, T vol, int zc, int ox, int oy, int oz) {
  typedef StageCfg<T, PGRAD, TX, TY, NS> Cfg;
  typedef typename Cfg::Pipe Pipe;
  constexpr int BX = Pipe::BX, UPL = Pipe::UPL, NT = Cfg::NT, XL = Pipe::XL;
  constexpr int POFF = Pipe::POFF, PSLOT = Pipe::PSLOT, PEND = POFF + Pipe::NPS * PSLOT;
  constexpr bool U0TILE = !PGRAD;
  constexpr int NKP = GEN ? 3 : 1;  // earlier k_j prefetched in registers (RK44's last row has 3)
  extern __shared__ unsigned char smem_raw[];
  unsigned char *sb = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  T *su = reinterpret_cast<T *>(sb);
  T *sdx = su + Pipe::ELEMS;                          // [2][TY][TX+1]  u'_x incl. column -1
  T *sdy = sdx + 2 * TY * (TX + 1);                   // [2][TY+1][TX]  u'_y incl. row -1
  uint64_t *bars = reinterpret_cast<uint64_t *>((reinterpret_cast<uintptr_t>(sdy + 2 * (TY + 1) * TX) + 7) & ~(uintptr_t)7);

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TX + tx;
  const int i0 = blockIdx.x * TX, j0 = blockIdx.y * TY;
  const int kbeg = blockIdx.z * zc, kend = min(kbeg + zc, g.n[2]);
  const int n0 = g.n[0], n1 = g.n[1], n2 = g.n[2];
  const bool has_force = A.force[0] != nullptr;
  const bool has_kout = !GEN || A.kout[0] != nullptr;
  const int nk = GEN ? A.nk : 0;
  const T coef_self = A.coef_self;
  const size_t szb = (size_t)A.sz * sizeof(T), scb = (size_t)A.sc * sizeof(T);

  // plane q <-> z = kbeg - 2 + q
  Pipe pipe;
  pipe.init(su, bars, &tm_u, &tm_p, ox + i0 - XL, oy + j0 - 2, oz + kbeg - 2, tid);
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
    pipe.project_q(q, tid, NT, rh.v);
  }
  __syncthreads();

  const int gi = i0 + tx, gj = j0 + ty;
  const size_t ocell = ((size_t)gj * A.sy + gi) * sizeof(T);
  // the extra point of some threads on the minus halo: warp 1 (first TY lanes): (-1, l) component x;
  // warp 2: (l, -1) component y  (two different warps on two different schedulers)
  const bool hx = tid >= 32 && tid < 32 + TY, hy = tid >= 64 && tid < 64 + TX;
  const bool hany = hx || hy;
  const int hlx = hx ? -1 : tid - 64, hly = hx ? tid - 32 : -1;
  const int hgi = wrapi(i0 + hlx, n0, 1), hgj = wrapi(j0 + hly, n1, 1);
  const size_t ohalo = ((size_t)hgj * A.sy + hgi) * sizeof(T);
  const int ha = hx ? 0 : 1;
  const int c_own = (ty + 2) * BX + tx + XL;
  const int c_halo = hany ? (hly + 2) * BX + hlx + XL : c_own;
  // halo threads: bases of their one component
  const T *hb_u0 = hx ? A.u0[0] : A.u0[1];
  const T *hb_s = hx ? A.sadd[0] : A.sadd[1];
  const T *hb_f = hx ? A.force[0] : A.force[1];

  // f (convection-diffusion) -> k_i and u' of component a at the cell with byte offset ob
  // (the first NKP earlier derivatives k_j arrive prefetched in l_k; further ones are read in place)
  auto combine = [&](T f, const int a, const size_t ob, const T uc, const T l_u0, const T l_s, const T l_f,
                     const T (&l_k)[NKP], T &f_out) -> T {
    if (has_force) f = f + l_f;
    if (SADD) f = f + l_s;
    f_out = f;
    T acc = U0TILE ? uc : l_u0;
    if (GEN) {
#pragma unroll
      for (int q = 0; q < NKP; q++)
        if (q < nk) acc = acc + A.coef[q] * l_k[q];
      for (int q = NKP; q < nk; q++) acc = acc + A.coef[q] * ld_off(a == 0 ? A.kin[q][0] : (a == 1 ? A.kin[q][1] : A.kin[q][2]), ob);
    }
    return acc + coef_self * f;
  };

  // u'_z of plane kbeg-1 (for the z-difference of the divergence on the first plane of the chunk)
  T uz_prev = T(0);
  const bool have_prev = kbeg > 0 || g.wrap[2];  // otherwise plane 0 is completed by k_divergence_planes
  if (have_prev) {
    const int kz = wrapi(kbeg - 1, n2, g.wrap[2]);
    const size_t ob = (size_t)kz * szb + ocell;
    PlaneAcc<T, BX, UPL> U{su, Pipe::uoff(0) + c_own, Pipe::uoff(1) + c_own, Pipe::uoff(2) + c_own};
    const T l_u0 = U0TILE ? T(0) : ld_off(A.u0[2], ob);
    const T l_s = SADD ? ld_off(A.sadd[2], ob) : T(0);
    const T l_f = has_force ? ld_off(A.force[2], ob) : T(0);
    T l_k[NKP];
#pragma unroll
    for (int q = 0; q < NKP; q++) l_k[q] = (GEN && q < nk) ? ld_off(A.kin[q][2], ob) : T(0);
    T f;
    uz_prev = combine(convdiff_comp<T, 3, true, true>(U, 2, visc, rh.v), 2, ob, U(2, 0, 0, 0), l_u0, l_s, l_f, l_k, f);
  }
  __syncthreads();  // plane 0 is retired: its slot takes plane NS
  if (tid == 0 && NS <= qlast) pipe.issue(NS);

  // slot offsets of planes q-1, q, q+1, q+2 [, q+3] at q = 2, rotated every iteration
  int so[NS];
#pragma unroll
  for (int i = 0; i < NS; i++) so[i] = Pipe::uoff(1 + i);
  int ws = 4 % NS;                       // barrier slot and parity of plane q+2
  uint32_t wp = (uint32_t)((4 / NS) & 1);
  int pa = Pipe::poff(4), pb = Pipe::poff(5);  // pressure planes q+2, q+3
  size_t ob = (size_t)kbeg * szb + ocell, obh = (size_t)kbeg * szb + ohalo;

  // global operands (ustart, closure term, force) are software-pipelined one plane ahead in registers:
  // requested as one batch before the barrier of iteration k-1, consumed in iteration k
  T n_u0[3], n_s[3], n_f[3], n_k[3][NKP], nh_u0 = T(0), nh_s = T(0), nh_f = T(0), nh_k[NKP];
#pragma unroll
  for (int q = 0; q < NKP; q++) nh_k[q] = T(0);
  auto fetch_operands = [&](const size_t o1, const size_t o1h) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      n_u0[a] = U0TILE ? T(0) : ld_off(A.u0[a], o1);
      n_s[a] = SADD ? ld_off(A.sadd[a], o1) : T(0);
      n_f[a] = has_force ? ld_off(A.force[a], o1) : T(0);
#pragma unroll
      for (int q = 0; q < NKP; q++) n_k[a][q] = (GEN && q < nk) ? ld_off(A.kin[q][a], o1) : T(0);
    }
    if (hany) {
      if (!U0TILE) nh_u0 = ld_off(hb_u0, o1h);
      if (SADD) nh_s = ld_off(hb_s, o1h);
      if (has_force) nh_f = ld_off(hb_f, o1h);
#pragma unroll
      for (int q = 0; q < NKP; q++)
        if (GEN && q < nk) nh_k[q] = ld_off(hx ? A.kin[q][0] : A.kin[q][1], o1h);
    }
  };
  fetch_operands(ob, obh);
  for (int k = kbeg; k < kend; k++) {
    const int q = k - kbeg + 2;  // plane number of k
    T l_u0[3], l_s[3], l_f[3], l_k[3][NKP], h_k[NKP];
#pragma unroll
    for (int a = 0; a < 3; a++) {
      l_u0[a] = n_u0[a]; l_s[a] = n_s[a]; l_f[a] = n_f[a];
#pragma unroll
      for (int q = 0; q < NKP; q++) l_k[a][q] = n_k[a][q];
    }
#pragma unroll
    for (int q = 0; q < NKP; q++) h_k[q] = nh_k[q];
    const T h_u0 = nh_u0, h_s = nh_s, h_f = nh_f;
    PlaneAcc<T, BX, UPL> U{su, so[0] + c_own, so[1] + c_own, so[2] + c_own};
    T un[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
      T f;
      un[a] = combine(convdiff_comp<T, 3, true, true>(U, a, visc, rh.v), a, ob, U(a, 0, 0, 0), l_u0[a], l_s[a], l_f[a], l_k[a], f);
      if (has_kout) st_off(A.kout[a], ob, f);
      st_off(A.unew[a], ob, un[a]);
    }
    T *bx = sdx + (k & 1) * TY * (TX + 1), *by = sdy + (k & 1) * (TY + 1) * TX;
    bx[ty * (TX + 1) + tx + 1] = un[0];
    by[(ty + 1) * TX + tx] = un[1];
    if (hany) {
      PlaneAcc<T, BX, UPL> Uh{su, so[0] + c_halo, so[1] + c_halo, so[2] + c_halo};
      T f;
      const T fc = hx ? convdiff_comp<T, 3, true, true>(Uh, 0, visc, rh.v) : convdiff_comp<T, 3, true, true>(Uh, 1, visc, rh.v);
      const T v = combine(fc, ha, obh, Uh(ha, 0, 0, 0), h_u0, h_s, h_f, h_k, f);
      if (hx) bx[hly * (TX + 1)] = v;
      else by[hlx] = v;
    }
    // plane k+2 for the next iteration
    if (k + 1 < kend) {
      fetch_operands(ob + szb, obh + szb);
      pipe.wait_slot(ws, wp);
      pipe.project(so[3], pa, pb, tid, NT, rh.v);
    }
    __syncthreads();
    // slot of plane k-1 is free now: request plane k-1+NS
    if (tid == 0 && q - 1 + NS <= qlast) pipe.issue(q - 1 + NS);
    if (k > kbeg || have_prev) {
      T d = T(0);
      d += (un[0] - bx[ty * (TX + 1) + tx]) * rh.v[0];
      d += (un[1] - by[ty * TX + tx]) * rh.v[1];
      d += (un[2] - uz_prev) * rh.v[2];
      st_off(A.div, ob, d * vol);
    }
    uz_prev = un[2];
    const int t = so[0];
#pragma unroll
    for (int i = 0; i + 1 < NS; i++) so[i] = so[i + 1];
    so[NS - 1] = t;
    ws = ws + 1 == NS ? 0 : ws + 1;
    wp ^= (ws == 0) ? 1u : 0u;
    pa = pb;
    pb = pb + PSLOT == PEND ? POFF : pb + PSLOT;
    ob += szb;
    obh += szb;
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
  const T *s;
  int om, o0, op;  // sigma planes k-1, k, k+1 (slot + own cell)
  __device__ __forceinline__ T operator()(int c, int dx, int dy, int dz) const {
    const int o = dz < 0 ? om : (dz > 0 ? op : o0);
    return s[o + c * SPL + dy * SX + dx];
  }
};

template <typename T, bool PGRAD, int TX, int TY>
struct SmagCfg {
  typedef PlanePipe<T, TX, TY + 4, PGRAD, 4> Pipe;
  static constexpr int NT = TX * TY;
  static constexpr int SX = TX + 2, SY = TY + 2, SPL = SX * SY, SSLOT = 6 * SPL;
  static constexpr size_t SMEM = 128 + (size_t)(Pipe::ELEMS + 3 * SSLOT) * sizeof(T) + 8 * Pipe::NS + 16;
};

template <typename T, bool PGRAD, int TX, int TY>
__global__ void __launch_bounds__(TX *TY, (sizeof(T) == 4 ? 2 : 1))
    k_smag3d_tma(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_p, T *__restrict__ sout,
                 i64 sy_, i64 sz_, i64 sc_, Geo g, T c_eddy, Rh<T> rh, int zc, int ox, int oy, int oz) {
  typedef SmagCfg<T, PGRAD, TX, TY> Cfg;
  typedef typename Cfg::Pipe Pipe;
  constexpr int BX = Pipe::BX, UPL = Pipe::UPL, NT = Cfg::NT, XL = Pipe::XL, NS = Pipe::NS;
  constexpr int POFF = Pipe::POFF, PSLOT = Pipe::PSLOT, PEND = POFF + Pipe::NPS * PSLOT;
  constexpr int SX = Cfg::SX, SPL = Cfg::SPL, SSLOT = Cfg::SSLOT;
  constexpr int SOFF = Pipe::ELEMS;  // sigma ring follows the plane slots
  extern __shared__ unsigned char smem_raw[];
  unsigned char *sb = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  T *su = reinterpret_cast<T *>(sb);
  uint64_t *bars = reinterpret_cast<uint64_t *>((reinterpret_cast<uintptr_t>(su + SOFF + 3 * SSLOT) + 7) & ~(uintptr_t)7);

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TX + tx;
  const int i0 = blockIdx.x * TX, j0 = blockIdx.y * TY;
  const int kbeg = blockIdx.z * zc, kend = min(kbeg + zc, g.n[2]);
  const unsigned sy = (unsigned)sy_, sz = (unsigned)sz_, sc = (unsigned)sc_;

  Pipe pipe;
  pipe.init(su, bars, &tm_u, &tm_p, ox + i0 - XL, oy + j0 - 2, oz + kbeg - 2, tid);
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
    pipe.project_q(q, tid, NT, rh.v);
  }
  __syncthreads();

  // the (up to two) p-cells of this thread in the (TX+2) x (TY+2) sigma tile: l = tid and l = tid + NT
  const int l0 = tid, l1 = tid + NT;
  const int l0y = l0 / SX, l0x = l0 - l0y * SX;
  const int l1y = l1 / SX, l1x = l1 - l1y * SX;
  const int cu0 = (l0y + 1) * BX + l0x - 1 + XL, cu1 = (l1y + 1) * BX + l1x - 1 + XL;  // p-cell (lx-1, ly-1)
  const bool two = l1 < SPL;
  // sigma plane into ring slot offset `so` from velocity slots um, u0, up
  auto sigma_plane = [&](int so, int um, int u0, int up) {
    T sig[6];
    {
      PlaneAcc<T, BX, UPL> U{su, um + cu0, u0 + cu0, up + cu0};
      smag_sigma<T, 3>(U, c_eddy, rh.v, sig);
#pragma unroll
      for (int c = 0; c < 6; c++) su[so + c * SPL + l0] = sig[c];
    }
    if (two) {
      PlaneAcc<T, BX, UPL> U{su, um + cu1, u0 + cu1, up + cu1};
      smag_sigma<T, 3>(U, c_eddy, rh.v, sig);
#pragma unroll
      for (int c = 0; c < 6; c++) su[so + c * SPL + l1] = sig[c];
    }
  };
  sigma_plane(SOFF, Pipe::uoff(0), Pipe::uoff(1), Pipe::uoff(2));
  sigma_plane(SOFF + SSLOT, Pipe::uoff(1), Pipe::uoff(2), Pipe::uoff(3));
  __syncthreads();  // velocity planes 0 and 1 are retired
  if (tid == 0) {
    if (NS <= qlast) pipe.issue(NS);
    if (NS + 1 <= qlast) pipe.issue(NS + 1);
  }

  const int c_sig = (ty + 1) * SX + tx + 1;
  // at r = 1: velocity planes r+1, r+2 resident (slots 2, 3), r+3 landing in slot 0, r+4 in slot 1
  int u_a = Pipe::uoff(2), u_b = Pipe::uoff(3), u_c = Pipe::uoff(0), u_d = Pipe::uoff(1);
  int pa = Pipe::poff(4), pb = Pipe::poff(5);          // pressure planes r+3, r+4
  int g_m = SOFF, g_0 = SOFF + SSLOT, g_p = SOFF + 2 * SSLOT;  // sigma planes r-1, r, r+1
  unsigned o = (unsigned)kbeg * sz + (unsigned)(j0 + ty) * sy + i0 + tx;
  for (int k = kbeg; k < kend; k++) {
    const int r = k - kbeg + 1;  // sigma plane number of k
    // sigma(k+1) needs velocity planes r+1, r+2, r+3 (z = k .. k+2)
    pipe.wait(r + 3);
    pipe.project(u_c, pa, pb, tid, NT, rh.v);
    __syncthreads();
    sigma_plane(g_p, u_a, u_b, u_c);
    __syncthreads();
    // velocity plane r+1 (z = k) is retired: its slot takes plane r+1+NS
    if (tid == 0 && r + 1 + NS <= qlast) pipe.issue(r + 1 + NS);
    SigmaAcc<T, SX, SPL> Sg{su, g_m + c_sig, g_0 + c_sig, g_p + c_sig};
#pragma unroll
    for (int a = 0; a < 3; a++) sout[o + a * sc] = divsigma_comp<T, 3>(Sg, a, rh.v);
    const int t = u_a;
    u_a = u_b; u_b = u_c; u_c = u_d; u_d = t;
    const int gt = g_m;
    g_m = g_0; g_0 = g_p; g_p = gt;
    pa = pb;
    pb = pb + PSLOT == PEND ? POFF : pb + PSLOT;
    o += sz;
  }
}

// -------------------------------------------------------------------------------------------------
// k_smag3d_v2: the same operator with the z-march carried in REGISTERS.  Every thread keeps the part of
// its stencil that lies on the lower planes from the iterations before (13 velocity values per p-cell,
// 11 tensor values per u-cell), so per plane a p-cell reads 17 instead of 30 velocity words and a u-cell
// 19 instead of 30 tensor words from shared memory -- the v1 kernel is bound by shared-memory wavefronts
// and issue slots, not by HBM.  One block of TX x (TY+4) threads per SM: thread l < (TX+2)(TY+2) owns
// p-cell l of the sigma tile (no second pass), threads with ty < TY own a u-cell.  The loop is skewed so
// that ONE barrier per plane suffices: iteration `it` projects velocity plane it+3, forms sigma(it) from
// the velocity planes it+1 (registers + mid extras) and it+2 (top), and takes the divergence of sigma
// planes it-3 .. it-1 for the output plane z = kbeg + it - 3; the three parts touch disjoint memory.
// Arithmetic is smag_sigma / divsigma_comp unchanged (register-backed accessors): bit-identical to v1.
// -------------------------------------------------------------------------------------------------
// Register files of the march.  Planes rotate through three register sets; PH (= iteration mod 3, a
// compile-time constant of the unrolled loop) says which set is bottom / mid / top, so that advancing one
// plane moves no data.
template <typename T>
struct RegU {
  T x[3][2], xe[4];  // ux: {0, -x} per plane set; mid-plane extras {+y, +y-x, -y, -y-x}
  T y[3][2], ye[4];  // uy: {0, -y} per plane set; mid-plane extras {+x, +x-y, -x, -x-y}
  T z[3][5];         // uz: {0, +x, -x, +y, -y} per plane set (mid and bottom are live)
};
template <typename T, int PH>
struct RegUView {  // dz = -1, 0, +1 -> set (PH + dz + 1) % 3
  const RegU<T> &r;
  __device__ __forceinline__ T operator()(int c, int dx, int dy, int dz) const {
    const int s = (PH + dz + 1) % 3;
    if (c == 0) return dy == 0 ? r.x[s][-dx] : r.xe[(dy > 0 ? 0 : 2) - dx];
    if (c == 1) return dx == 0 ? r.y[s][-dy] : r.ye[(dx > 0 ? 0 : 2) - dy];
    return r.z[s][dx > 0 ? 1 : (dx < 0 ? 2 : (dy > 0 ? 3 : (dy < 0 ? 4 : 0)))];
  }
};

template <typename T>
struct RegSg {
  T xx[2], yy[2], zz[3];  // {0, +x}, {0, +y}; sigma_zz(0) per plane set (mid and top are live)
  T xy[8];                // (0,0) (1,0) (0,1) (1,1) (0,-1) (1,-1) (-1,0) (-1,1)
  T xz[3][3], yz[3][3];   // sigma_xz {0, +x, -x}, sigma_yz {0, +y, -y} per plane set
};
template <typename T, int PH>
struct RegSgView {
  const RegSg<T> &r;
  __device__ __forceinline__ T operator()(int c, int dx, int dy, int dz) const {
    const int s = (PH + dz + 1) % 3;
    if (c == 0) return r.xx[dx];
    if (c == 3) return r.yy[dy];
    if (c == 5) return r.zz[s];
    if (c == 1) return r.xy[dy == 0 ? (dx == 0 ? 0 : (dx > 0 ? 1 : 6)) : (dy > 0 ? (dx == 0 ? 2 : (dx > 0 ? 3 : 7)) : (dx == 0 ? 4 : 5))];
    if (c == 2) return r.xz[s][dx == 0 ? 0 : (dx > 0 ? 1 : 2)];
    return r.yz[s][dy == 0 ? 0 : (dy > 0 ? 1 : 2)];
  }
};

template <int J>
struct IntC {
  static constexpr int value = J;
};

template <typename T, bool PGRAD, int TX, int TY>
struct Smag2Cfg {
  typedef PlanePipe<T, TX, TY + 4, PGRAD, 6, 6> Pipe;  // 6 velocity and 6 pressure slots: period of the unrolled loop
  static constexpr int NT = TX * (TY + 4);
  static constexpr int SX = TX + 2, SY = TY + 2, SPL = SX * SY, SSLOT = 6 * SPL;
  static_assert(SPL <= NT, "one p-cell of the sigma tile per thread");
  static constexpr size_t SMEM = 128 + (size_t)(Pipe::ELEMS + 3 * SSLOT) * sizeof(T) + 8 * Pipe::NS + 16;
};

template <typename T, bool PGRAD, int TX, int TY>
__global__ void __launch_bounds__(TX *(TY + 4), (TY <= 8 ? 2 : 1))
    k_smag3d_v2(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_p, T *__restrict__ sout,
                i64 sy_, i64 sz_, i64 sc_, Geo g, T c_eddy, Rh<T> rh, int zc, int ox, int oy, int oz) {
  typedef Smag2Cfg<T, PGRAD, TX, TY> Cfg;
  typedef typename Cfg::Pipe Pipe;
  constexpr int BX = Pipe::BX, UPL = Pipe::UPL, NT = Cfg::NT, XL = Pipe::XL, NS = Pipe::NS, USLOT = Pipe::USLOT;
  constexpr int POFF = Pipe::POFF, PSLOT = Pipe::PSLOT;
  constexpr int SX = Cfg::SX, SPL = Cfg::SPL, SSLOT = Cfg::SSLOT;
  constexpr int SOFF = Pipe::ELEMS;  // sigma ring follows the plane slots
  static_assert(NS == 6 && Pipe::NPS == 6, "slot periods must divide the unroll factor");
  extern __shared__ unsigned char smem_raw[];
  unsigned char *sb = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  T *su = reinterpret_cast<T *>(sb);
  uint64_t *bars = reinterpret_cast<uint64_t *>((reinterpret_cast<uintptr_t>(su + SOFF + 3 * SSLOT) + 7) & ~(uintptr_t)7);

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TX + tx;
  const int i0 = blockIdx.x * TX, j0 = blockIdx.y * TY;
  const int kbeg = blockIdx.z * zc, nk = min(zc, g.n[2] - kbeg);
  const unsigned sz = (unsigned)sz_, sc = (unsigned)sc_;

  Pipe pipe;
  pipe.init(su, bars, &tm_u, &tm_p, ox + i0 - XL, oy + j0 - 2, oz + kbeg - 2, tid);
  __syncthreads();
  const int qlast = nk + 3;  // velocity planes q = 0 .. nk+3 <-> z = kbeg-2 .. kbeg+nk+1
  if (tid == 0) {
#pragma unroll
    for (int q = 0; q < NS - 1; q++)  // the pressure ring holds planes 0 .. NS-1 at this point
      if (q <= qlast) pipe.issue(q);
  }
  pipe.wait(0);
  pipe.project_q(0, tid, NT, rh.v);
  __syncthreads();

  const bool sig_thr = tid < SPL, div_thr = ty < TY;
  const int ly = tid / SX, lx = tid - ly * SX;
  const int cu = sig_thr ? (ly + 1) * BX + lx - 1 + XL : BX + XL;  // p-cell (lx-1, ly-1); box row 0 is y = -2
  const int c_sig = (ty + 1) * SX + tx + 1;
  unsigned o = (unsigned)kbeg * sz + (unsigned)(j0 + ty) * (unsigned)sy_ + i0 + tx;

  RegU<T> U;
  RegSg<T> Sg;
#pragma unroll
  for (int s = 0; s < 3; s++) {
    U.x[s][0] = U.x[s][1] = U.y[s][0] = U.y[s][1] = Sg.zz[s] = T(0);
#pragma unroll
    for (int i = 0; i < 5; i++) U.z[s][i] = T(0);
#pragma unroll
    for (int i = 0; i < 3; i++) Sg.xz[s][i] = Sg.yz[s][i] = T(0);
  }

  // One plane step; J = (it + 2) mod 6 is a compile-time constant, so every slot offset below is an
  // immediate: velocity plane q lives in slot q mod 6, pressure plane q in slot q mod 6, sigma plane r in
  // ring slot (r + 2) mod 3.
  auto step = [&](auto jc, const int it, const unsigned par) {
    constexpr int J = decltype(jc)::value, PH = J % 3;
    constexpr int u_mid = ((J + 5) % 6) * USLOT, u_top = (J % 6) * USLOT, u_prj = ((J + 1) % 6) * USLOT;
    constexpr int pa = POFF + ((J + 1) % 6) * PSLOT, pb = POFF + ((J + 2) % 6) * PSLOT;
    constexpr int g_w = SOFF + PH * SSLOT, g_t = SOFF + ((PH + 2) % 3) * SSLOT, g_m = SOFF + ((PH + 1) % 3) * SSLOT;
    constexpr int TOP = (PH + 2) % 3, MID = (PH + 1) % 3;
    // ---- A: velocity plane it+3 lands and is projected in place
    if (it + 3 <= qlast) {
      pipe.wait_slot((J + 1) % 6, J == 5 ? par ^ 1u : par);
      pipe.project(u_prj, pa, pb, tid, NT, rh.v);
    }
    // ---- B: sigma(it) at this thread's p-cell
    if (sig_thr && it <= nk + 1) {
      const T *ut = su + u_top + cu;
      U.x[TOP][0] = ut[0];
      U.x[TOP][1] = ut[-1];
      U.y[TOP][0] = ut[UPL];
      U.y[TOP][1] = ut[UPL - BX];
      if (it >= -1) {
        const T *um = su + u_mid + cu;
        U.xe[0] = um[BX];
        U.xe[1] = um[BX - 1];
        U.xe[2] = um[-BX];
        U.xe[3] = um[-BX - 1];
        U.ye[0] = um[UPL + 1];
        U.ye[1] = um[UPL + 1 - BX];
        U.ye[2] = um[UPL - 1];
        U.ye[3] = um[UPL - 1 - BX];
        U.z[MID][0] = um[2 * UPL];
        U.z[MID][1] = um[2 * UPL + 1];
        U.z[MID][2] = um[2 * UPL - 1];
        U.z[MID][3] = um[2 * UPL + BX];
        U.z[MID][4] = um[2 * UPL - BX];
        if (it >= 0) {
          T sig[6];
          RegUView<T, PH> Uv{U};
          smag_sigma<T, 3>(Uv, c_eddy, rh.v, sig);
#pragma unroll
          for (int c = 0; c < 6; c++) su[g_w + c * SPL + tid] = sig[c];
        }
      }
    }
    // ---- C: s(z = kbeg + it - 3) = div sigma from sigma planes it-3 (registers), it-2, it-1
    if (div_thr && it >= 1) {
      const T *st = su + g_t + c_sig;
      Sg.xz[TOP][0] = st[2 * SPL];
      Sg.xz[TOP][1] = st[2 * SPL + 1];
      Sg.xz[TOP][2] = st[2 * SPL - 1];
      Sg.yz[TOP][0] = st[4 * SPL];
      Sg.yz[TOP][1] = st[4 * SPL + SX];
      Sg.yz[TOP][2] = st[4 * SPL - SX];
      Sg.zz[TOP] = st[5 * SPL];
      if (it >= 3) {
        const T *sm = su + g_m + c_sig;
        Sg.xx[0] = sm[0];
        Sg.xx[1] = sm[1];
        Sg.yy[0] = sm[3 * SPL];
        Sg.yy[1] = sm[3 * SPL + SX];
        Sg.xy[0] = sm[SPL];
        Sg.xy[1] = sm[SPL + 1];
        Sg.xy[2] = sm[SPL + SX];
        Sg.xy[3] = sm[SPL + SX + 1];
        Sg.xy[4] = sm[SPL - SX];
        Sg.xy[5] = sm[SPL - SX + 1];
        Sg.xy[6] = sm[SPL - 1];
        Sg.xy[7] = sm[SPL + SX - 1];
        RegSgView<T, PH> Sv{Sg};
#pragma unroll
        for (int a = 0; a < 3; a++) sout[o + a * sc] = divsigma_comp<T, 3>(Sv, a, rh.v);
        o += sz;
      }
    }
    __syncthreads();
    // velocity plane it+1 is retired: its slot takes plane it+1+NS
    if (tid == 0 && it + 1 + NS <= qlast) pipe.issue(it + 1 + NS);
  };

  unsigned par = 0;  // phase parity of the barriers of planes 6m+1 .. 6m+5 (plane 6m+6 has the opposite one)
  for (int it = -2; it <= nk + 2; it += 6) {
    step(IntC<0>{}, it, par);
    if (it + 1 <= nk + 2) step(IntC<1>{}, it + 1, par);
    if (it + 2 <= nk + 2) step(IntC<2>{}, it + 2, par);
    if (it + 3 <= nk + 2) step(IntC<3>{}, it + 3, par);
    if (it + 4 <= nk + 2) step(IntC<4>{}, it + 4, par);
    if (it + 5 <= nk + 2) step(IntC<5>{}, it + 5, par);
    par ^= 1u;
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
