// tma_stage_x2.cuh -- k_stage3d_tma with TWO cells per thread on packed FP32 (f32x2.cuh), Float32 only.
//
// Same tile (TX x TY cells, z-marching), same TMA plane pipeline, same shared-memory layout and the same arithmetic
// as k_stage3d_tma (tma_stage.cuh) -- the stencil bodies are the SAME templates (convdiff_comp<T = f2>), so every
// lane performs the scalar kernel's operations in the scalar kernel's order: results are bit-identical.  A block has
// TX x TY/2 threads; thread (tx, ty) owns the cells (tx, ty) and (tx, ty + TY/2), the lo and hi lane of every packed
// value.  Shared-memory and global accesses stay 32-bit per cell (both lanes of a pair load straight into an aligned
// register pair); what halves is the FADD/FMUL count, the address arithmetic, the loop control and the barrier
// traffic per cell.
#pragma once
#include "f32x2.cuh"
#include "tma_stage.cuh"

namespace insb200 {

// velocity accessor on three resident plane slots for a PAIR of cells `po` elements apart
template <typename f2, int BX, int UPL>
struct PlaneAcc2 {
  const float *s;
  int om, o0, op;  // planes k-1, k, k+1 (slot + lo cell)
  int po;          // element offset of the hi cell
  __device__ __forceinline__ f2 operator()(int c, int dx, int dy, int dz) const {
    const int o = (dz < 0 ? om : (dz > 0 ? op : o0)) + c * UPL + dy * BX + dx;
    return f2(s[o], s[o + po]);
  }
};

template <typename f2>
__device__ __forceinline__ f2 ld2_off(const float *base, size_t ob, size_t pob) { return f2(ld_off(base, ob), ld_off(base, ob + pob)); }
template <typename f2>
__device__ __forceinline__ void st2_off(float *base, size_t ob, size_t pob, f2 v) {
  st_off(base, ob, v.lo());
  st_off(base, ob + pob, v.hi());
}

template <typename f2, bool PGRAD, bool SADD, bool GEN, int TX, int TY, int NS>
__global__ void __launch_bounds__(TX *TY / 2, (GEN ? 2 : 3))
    k_stage3d_tma_x2(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_p,
                     const __grid_constant__ TmaStageArgs<float> A, Geo g, float visc_, Rh<float> rh_, float vol_, int zc, int ox,
                     int oy, int oz) {
  typedef float T;
  typedef StageCfg<T, PGRAD, TX, TY, NS> Cfg;
  typedef typename Cfg::Pipe Pipe;
  constexpr int BX = Pipe::BX, UPL = Pipe::UPL, XL = Pipe::XL;
  constexpr int HY = TY / 2, NT = TX * HY;
  constexpr int POFF = Pipe::POFF, PSLOT = Pipe::PSLOT, PEND = POFF + Pipe::NPS * PSLOT;
  constexpr bool U0TILE = !PGRAD;
  constexpr int NKP = GEN ? 3 : 1;
  constexpr int PO = HY * BX;  // hi cell of a pair in the plane slots
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
  const f2 coef_self(A.coef_self), visc(visc_), vol(vol_);
  const f2 rh[3] = {f2(rh_.v[0]), f2(rh_.v[1]), f2(rh_.v[2])};
  f2 cf[NKP];
#pragma unroll
  for (int q = 0; q < NKP; q++) cf[q] = f2((GEN && q < nk) ? A.coef[q] : 0.0f);
  const size_t szb = (size_t)A.sz * sizeof(T);
  const size_t pob = (size_t)HY * A.sy * sizeof(T);   // hi cell of a pair in global memory

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
#pragma unroll
  for (int q = 0; q < 4; q++) {
    pipe.wait(q);
    pipe.project_q(q, tid, NT, rh_.v);
  }
  __syncthreads();

  const int gi = i0 + tx, gj = j0 + ty;
  const size_t ocell = ((size_t)gj * A.sy + gi) * sizeof(T);
  // the extra points on the minus halo: warp 1, first HY lanes: pairs (-1, l), (-1, l + HY) of component x;
  // warp 2: (l, -1) of component y (no partner row: both lanes carry the same cell, the lo lane is stored)
  const bool hx = tid >= 32 && tid < 32 + HY, hy = tid >= 64 && tid < 64 + TX;
  const bool hany = hx || hy;
  const int hlx = hx ? -1 : tid - 64, hly = hx ? tid - 32 : -1;
  const int hgi = wrapi(i0 + hlx, n0, 1), hgj = wrapi(j0 + hly, n1, 1);
  const size_t ohalo = ((size_t)hgj * A.sy + hgi) * sizeof(T);
  const size_t pobh = hx ? pob : 0;
  const int ha = hx ? 0 : 1;
  const int c_own = (ty + 2) * BX + tx + XL;
  const int c_halo = hany ? (hly + 2) * BX + hlx + XL : c_own;
  const int po_halo = hx ? PO : 0;
  const T *hb_u0 = hx ? A.u0[0] : A.u0[1];
  const T *hb_s = hx ? A.sadd[0] : A.sadd[1];
  const T *hb_f = hx ? A.force[0] : A.force[1];

  // f (convection-diffusion) -> k_i and u' of component a for the pair at byte offsets (ob, ob + pb)
  auto combine = [&](f2 f, const int a, const size_t ob, const size_t pb, const f2 uc, const f2 l_u0, const f2 l_s, const f2 l_f,
                     const f2 (&l_k)[NKP], f2 &f_out) -> f2 {
    if (has_force) f = f + l_f;
    if (SADD) f = f + l_s;
    f_out = f;
    f2 acc = U0TILE ? uc : l_u0;
    if (GEN) {
#pragma unroll
      for (int q = 0; q < NKP; q++)
        if (q < nk) acc = acc + cf[q] * l_k[q];
      for (int q = NKP; q < nk; q++)
        acc = acc + f2(A.coef[q]) * ld2_off<f2>(a == 0 ? A.kin[q][0] : (a == 1 ? A.kin[q][1] : A.kin[q][2]), ob, pb);
    }
    return acc + coef_self * f;
  };

  const f2 zero2(0.0f);
  // u'_z of plane kbeg-1 (for the z-difference of the divergence on the first plane of the chunk)
  f2 uz_prev = zero2;
  const bool have_prev = kbeg > 0 || g.wrap[2];  // otherwise plane 0 is completed by k_divergence_planes
  if (have_prev) {
    const int kz = wrapi(kbeg - 1, n2, g.wrap[2]);
    const size_t ob = (size_t)kz * szb + ocell;
    PlaneAcc2<f2, BX, UPL> U{su, Pipe::uoff(0) + c_own, Pipe::uoff(1) + c_own, Pipe::uoff(2) + c_own, PO};
    const f2 l_u0 = U0TILE ? zero2 : ld2_off<f2>(A.u0[2], ob, pob);
    const f2 l_s = SADD ? ld2_off<f2>(A.sadd[2], ob, pob) : zero2;
    const f2 l_f = has_force ? ld2_off<f2>(A.force[2], ob, pob) : zero2;
    f2 l_k[NKP];
#pragma unroll
    for (int q = 0; q < NKP; q++) l_k[q] = (GEN && q < nk) ? ld2_off<f2>(A.kin[q][2], ob, pob) : zero2;
    f2 f;
    uz_prev = combine(convdiff_comp<f2, 3, true, true>(U, 2, visc, rh), 2, ob, pob, U(2, 0, 0, 0), l_u0, l_s, l_f, l_k, f);
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

  // global operands (ustart, closure term, force, earlier k_j) software-pipelined one plane ahead in registers
  f2 n_u0[3], n_s[3], n_f[3], n_k[3][NKP], nh_u0 = zero2, nh_s = zero2, nh_f = zero2, nh_k[NKP];
#pragma unroll
  for (int q = 0; q < NKP; q++) nh_k[q] = zero2;
  auto fetch_operands = [&](const size_t o1, const size_t o1h) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      n_u0[a] = U0TILE ? zero2 : ld2_off<f2>(A.u0[a], o1, pob);
      n_s[a] = SADD ? ld2_off<f2>(A.sadd[a], o1, pob) : zero2;
      n_f[a] = has_force ? ld2_off<f2>(A.force[a], o1, pob) : zero2;
#pragma unroll
      for (int q = 0; q < NKP; q++) n_k[a][q] = (GEN && q < nk) ? ld2_off<f2>(A.kin[q][a], o1, pob) : zero2;
    }
    if (hany) {
      if (!U0TILE) nh_u0 = ld2_off<f2>(hb_u0, o1h, pobh);
      if (SADD) nh_s = ld2_off<f2>(hb_s, o1h, pobh);
      if (has_force) nh_f = ld2_off<f2>(hb_f, o1h, pobh);
#pragma unroll
      for (int q = 0; q < NKP; q++)
        if (GEN && q < nk) nh_k[q] = ld2_off<f2>(hx ? A.kin[q][0] : A.kin[q][1], o1h, pobh);
    }
  };
  fetch_operands(ob, obh);
  for (int k = kbeg; k < kend; k++) {
    const int q = k - kbeg + 2;  // plane number of k
    f2 l_u0[3], l_s[3], l_f[3], l_k[3][NKP], h_k[NKP];
#pragma unroll
    for (int a = 0; a < 3; a++) {
      l_u0[a] = n_u0[a]; l_s[a] = n_s[a]; l_f[a] = n_f[a];
#pragma unroll
      for (int qq = 0; qq < NKP; qq++) l_k[a][qq] = n_k[a][qq];
    }
#pragma unroll
    for (int qq = 0; qq < NKP; qq++) h_k[qq] = nh_k[qq];
    const f2 h_u0 = nh_u0, h_s = nh_s, h_f = nh_f;
    PlaneAcc2<f2, BX, UPL> U{su, so[0] + c_own, so[1] + c_own, so[2] + c_own, PO};
    f2 un[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
      f2 f;
      un[a] = combine(convdiff_comp<f2, 3, true, true>(U, a, visc, rh), a, ob, pob, U(a, 0, 0, 0), l_u0[a], l_s[a], l_f[a], l_k[a], f);
      if (has_kout) st2_off(A.kout[a], ob, pob, f);
      st2_off(A.unew[a], ob, pob, un[a]);
    }
    T *bx = sdx + (k & 1) * TY * (TX + 1), *by = sdy + (k & 1) * (TY + 1) * TX;
    bx[ty * (TX + 1) + tx + 1] = un[0].lo();
    bx[(ty + HY) * (TX + 1) + tx + 1] = un[0].hi();
    by[(ty + 1) * TX + tx] = un[1].lo();
    by[(ty + HY + 1) * TX + tx] = un[1].hi();
    if (hany) {
      PlaneAcc2<f2, BX, UPL> Uh{su, so[0] + c_halo, so[1] + c_halo, so[2] + c_halo, po_halo};
      f2 f;
      const f2 fc = hx ? convdiff_comp<f2, 3, true, true>(Uh, 0, visc, rh) : convdiff_comp<f2, 3, true, true>(Uh, 1, visc, rh);
      const f2 v = combine(fc, ha, obh, pobh, Uh(ha, 0, 0, 0), h_u0, h_s, h_f, h_k, f);
      if (hx) {
        bx[hly * (TX + 1)] = v.lo();
        bx[(hly + HY) * (TX + 1)] = v.hi();
      } else {
        by[hlx] = v.lo();
      }
    }
    // plane k+2 for the next iteration
    if (k + 1 < kend) {
      fetch_operands(ob + szb, obh + szb);
      pipe.wait_slot(ws, wp);
      pipe.project(so[3], pa, pb, tid, NT, rh_.v);
    }
    __syncthreads();
    // slot of plane k-1 is free now: request plane k-1+NS
    if (tid == 0 && q - 1 + NS <= qlast) pipe.issue(q - 1 + NS);
    if (k > kbeg || have_prev) {
      f2 d = zero2;
      d += (un[0] - f2(bx[ty * (TX + 1) + tx], bx[(ty + HY) * (TX + 1) + tx])) * rh[0];
      d += (un[1] - f2(by[ty * TX + tx], by[(ty + HY) * TX + tx])) * rh[1];
      d += (un[2] - uz_prev) * rh[2];
      st2_off(A.div, ob, pob, d * vol);
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

}  // namespace insb200
