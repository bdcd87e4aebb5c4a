// fast_stage.cuh -- k_stage3d_fast: ONE kernel per RK stage of the 3-D step on the padded layout (sm_100a), with the
// Smagorinsky closure folded in.  It replaces k_smag3d_v2 + k_stage3d_tma (tma_stage.cuh) as the default; those stay
// selectable with insb200_set_param("exact", 1) for the bit-for-bit operator tests.
//
// What one launch computes (step_explicit_runge_kutta.jl:17-50 up to the Poisson solve, operators.jl:1294-1305):
//   u    = u' - grad p          projection of the PREVIOUS stage, applied to each plane as it lands in shared memory
//   s    = div sigma(u)         sigma = 2 nu_t S at the p-cells lives ONLY in a 3-plane shared-memory ring (SMAG)
//   k_i  = convdiff(u) [+ force] [+ s]
//   u'   = ustart + sum_j (dt a_ij) k_j
//   div  = divergence(u') * Omega                    (the Poisson right-hand side)
// HBM words per cell and launch: R u (3), R p (1), R ustart (3), W k_i (3), W u' (3), W div (1) = 14 -- the closure
// adds none (as SURVEY 8d budgets it); the two-kernel form paid W s + R s + R u + R p = 10 more and projected twice.
//
// The old stage kernel was bound by instruction issue (428 warp instructions per warp and plane at 84 % issue-slot
// utilisation), not by HBM.  This kernel spends about a third of that on the same operator:
//   * FLUX FORM with the z-march in registers: the convective + diffusive flux through the upper z-face of every
//     component is formed once and carried to the next plane (it is the lower flux there); own-column values of the
//     plane above are loaded once and become the centre values of the next iteration: 17 instead of 44 shared-memory
//     reads per cell;
//   * products shared between components ((u_x+u_x(+y)) (u_y+u_y(+x)) is the upper flux of BOTH (x,y) and (y,x)),
//     diffusion as (up + um - 2 u), constants visc/h^2 and 1/(4h) folded, FMA contraction on (this translation unit
//     is compiled WITHOUT -fmad=false);
//   * the velocity gradient of the closure with its four one-sided differences telescoped, the corner averages of
//     divoftensor! telescoped the same way, both carried along z in registers.
// The arithmetic is the reference's operator in a different association order: results agree with the oracle to a few
// ulp per operator (tests/test_fast_formulas.py restates it in NumPy) and to <= 1e-5 (Float32) after RK steps, which is
// north_star's parity bar; it is NOT bit-identical (the exact kernels are).
//
// Block = tile of 32 x 16 cells, z-chunk of zc planes.  SMAG: 640 threads, one block per SM (156 KB of shared
// memory): thread l < 34 x 18 owns p-cell l of the sigma tile, threads with ty < 16 own a cell, warps 16 / 17 own the
// minus halo (u'_x at x = -1, u'_y at y = -1, needed by the divergence).  !SMAG: 512 threads, two blocks per SM, the
// halo points ride on warps 1 / 2.  ONE __syncthreads per plane: iteration r projects velocity plane r+4 (r+2 without
// closure), forms sigma(r+2) and does the stencil of plane r -- the three parts touch disjoint shared memory.
//
// Every routine that touches data is __host__ __device__: tools/fast_stage_emul.cu runs the same code on the host
// (threads serialised between barriers, TMA boxes copied by hand) against a plain triple loop -- the kernel's index
// arithmetic is checked on the CPU (tests/test_fast_stage_emul.py) before it ever reaches the GPU.
#pragma once
#include "tiled.cuh"
#include "tma_stage.cuh"

namespace insb200 {

#if defined(__CUDA_ARCH__)
#define FS_LDG(p) __ldg(p)
#else
#define FS_LDG(p) (*(p))
#endif
#define FS_HD __host__ __device__ __forceinline__

// constants of one launch (constant bank): 1/h, visc/h^2, 1/(4h), cell volume, 2 theta^2 d^2
template <typename T>
struct FastK {
  T rh[3], vrr[3], q[3], vol, e2c;
};

template <typename T, bool SMAG>
struct FastCfg {
  static constexpr int TX = 32, TY = 16;
  static constexpr int NTY = SMAG ? 20 : 16, NT = TX * NTY;
  static constexpr int V = 16 / (int)sizeof(T), XL = V, BX = TX + 2 * V;
  static constexpr int BY = SMAG ? TY + 4 : TY + 3, BYP = BY + 1;  // box rows: y = j0-2 .. j0+TY(+1)
  static constexpr int UPL = BY * BX;
  static constexpr int USLOT = ((3 * UPL * (int)sizeof(T) + 127) / 128) * 128 / (int)sizeof(T);
  static constexpr int PSLOT = ((BYP * BX * (int)sizeof(T) + 127) / 128) * 128 / (int)sizeof(T);
  static constexpr int NS = 8, NPS = 8, NG = 4;  // ring sizes: velocity planes, pressure planes, sigma planes (powers of two, see FsOffs)
  static constexpr int POFF = NS * USLOT;
  static constexpr int SX = TX + 2, SY = TY + 2, SPL = SX * SY, SSLOT = 6 * SPL;  // sigma tile: p-cells -1 .. TX / TY
  static constexpr int SOFF = POFF + NPS * PSLOT;
  static constexpr int DOFF = SOFF + (SMAG ? NG * SSLOT : 0);  // divergence staging: [2][TY][TX+1] + [2][TY+1][TX]
  static constexpr int DXN = TY * (TX + 1), DYN = (TY + 1) * TX;
  static constexpr int ELEMS = DOFF + 2 * DXN + 2 * DYN;
  static constexpr size_t SMEM = 128 + (size_t)ELEMS * sizeof(T) + 8 * NS + 16;
  static constexpr uint32_t UBYTES = 3u * UPL * sizeof(T), PBYTES = (uint32_t)(BYP * BX * sizeof(T));
  static constexpr int LEAD = SMAG ? 4 : 2;  // iteration r projects velocity plane r + LEAD
  static constexpr int R0 = -2 - LEAD;       // first iteration: projects plane -2
  static_assert(!SMAG || SPL <= NT, "one p-cell of the sigma tile per thread");
};

// ---- per-thread state of the march (registers on the device) ------------------------------------------------------
template <typename T>
struct FsCell {      // a cell's stencil: own-column values of the current plane and the lower z-face fluxes
  T cx0, cxm, cy0, cym, cz0;  // u_x(0), u_x(-x), u_y(0), u_y(-y), u_z(0) at plane r
  T fzx, fzy, fzz;            // fluxes through the lower z-face (diffusive - convective, scaled)
  T uzp;                      // u'_z of plane r-1
};
template <typename T>
struct FsCellSg {    // a cell's part of div sigma carried along z
  T sxz0, sxzm, syz0, syzm;   // sigma_xz(0)+sigma_xz(+x) at planes r, r-1; sigma_yz(0)+sigma_yz(+y) likewise
  T dxz0, dyz0, szz0;         // sigma_xz(+x)-sigma_xz(-x), sigma_yz(+y)-sigma_yz(-y), sigma_zz at plane r
};
template <typename T>
struct FsSig {       // a p-cell's part of grad u carried along z
  T ux0, uxm, uy0, uym;       // u_x(0)+u_x(-x) at planes rB, rB-1; u_y(0)+u_y(-y) likewise
  T uz0, uzm;                 // u_z at planes rB, rB-1
  T dxx0, dyy0;               // u_x(0)-u_x(-x), u_y(0)-u_y(-y) at plane rB
  T zxp, zxm, zyp, zym;       // u_z(+x), u_z(-x), u_z(+y), u_z(-y) at plane rB-1
};
template <typename T>
struct FsHalo {      // a halo point (one component): lower z-face flux and the carried closure sums
  T fz, s0, sm;
};

// sqrt for the eddy viscosity: x * rsqrt(x) (2 instructions on the device instead of the 9 of the IEEE sequence;
// relative error 2^-22, far inside the parity tolerance)
FS_HD float fs_sqrt(float x) {
#if defined(__CUDA_ARCH__)
  return x * rsqrtf(fmaxf(x, 1e-37f));
#else
  return sqrtf(x);
#endif
}
FS_HD double fs_sqrt(double x) { return sqrt(x); }

template <typename T>
FS_HD void fs_zero(T *p, int n) {
  for (int i = 0; i < n; i++) p[i] = T(0);
}

// ---- phase A: u(plane) = u'(plane) - grad p over the whole box ---------------------------------------------------
// 16 bytes (V elements) per thread and access: box rows are multiples of 16 B, so a plane is UPL / V vectors and the
// neighbours p(+y), p(+z) of a vector are vectors again; p(+x) needs one more scalar.  (The last column / row of the
// box is never consumed, so p[l+1] / p[l+BX] may run into the next row / the extra row of the pressure box.)
template <typename T>
struct alignas(16) FsVec {
  T v[16 / sizeof(T)];
};
// Thread `pt` of `np` projecting threads takes the vectors pt, pt + np, ...
template <typename T, typename Cfg>
FS_HD void fs_project(T *u, const T *p, const T *p1, int pt, int np, const T (&rh)[3]) {
  constexpr int V = Cfg::V, NV = Cfg::UPL / V;
  static_assert(Cfg::UPL % V == 0 && Cfg::BX % V == 0, "box rows are whole vectors");
  typedef FsVec<T> Vec;
  for (int l = pt; l < NV; l += np) {
    const Vec p0 = reinterpret_cast<const Vec *>(p)[l];
    const Vec py = reinterpret_cast<const Vec *>(p)[l + Cfg::BX / V];
    const Vec pz = reinterpret_cast<const Vec *>(p1)[l];
    const T pn = p[l * V + V];
    Vec ux = reinterpret_cast<Vec *>(u)[l], uy = reinterpret_cast<Vec *>(u + Cfg::UPL)[l], uz = reinterpret_cast<Vec *>(u + 2 * Cfg::UPL)[l];
#pragma unroll
    for (int i = 0; i < V; i++) {
      const T px = i + 1 < V ? p0.v[i + 1 < V ? i + 1 : 0] : pn;
      ux.v[i] = ux.v[i] - (px - p0.v[i]) * rh[0];
      uy.v[i] = uy.v[i] - (py.v[i] - p0.v[i]) * rh[1];
      uz.v[i] = uz.v[i] - (pz.v[i] - p0.v[i]) * rh[2];
    }
    reinterpret_cast<Vec *>(u)[l] = ux;
    reinterpret_cast<Vec *>(u + Cfg::UPL)[l] = uy;
    reinterpret_cast<Vec *>(u + 2 * Cfg::UPL)[l] = uz;
  }
}

// ---- phase B: sigma(rB) at this thread's p-cell -------------------------------------------------------------------
// M / Tq: this p-cell in the velocity planes rB / rB+1; out: this p-cell in the sigma slot of plane rB.
// rB = -3: top loads only; rB = -2: no tensor yet (the carried values of plane -2 are being set up)
// ST (steady state): every plane index is >= 0, the start-up conditions fold away at compile time
template <typename T, typename Cfg, bool ST>
FS_HD void fs_sigma(FsSig<T> &s, const T *M, const T *Tq, T *out, int rB, const FastK<T> &K) {
  constexpr int BX = Cfg::BX, UPL = Cfg::UPL, SPL = Cfg::SPL;
  const T t0 = Tq[0], t1 = Tq[-1], t2 = Tq[UPL], t3 = Tq[UPL - BX];
  const T uxt = t0 + t1, uyt = t2 + t3, uzt = Tq[2 * UPL];
  const T dxxt = t0 - t1, dyyt = t2 - t3;
  if (ST || rB >= -2) {
    const T zxp = M[2 * UPL + 1], zxm = M[2 * UPL - 1], zyp = M[2 * UPL + BX], zym = M[2 * UPL - BX];
    if (ST || rB >= -1) {
      const T gxy = (M[BX] + M[BX - 1]) - (M[-BX] + M[-BX - 1]);                              // d u_x / dy * 4 h_y
      const T gyx = (M[UPL + 1] + M[UPL + 1 - BX]) - (M[UPL - 1] + M[UPL - 1 - BX]);          // d u_y / dx * 4 h_x
      const T gxz = uxt - s.uxm, gyz = uyt - s.uym;                                           // * 4 h_z
      const T gzx = (zxp + s.zxp) - (zxm + s.zxm), gzy = (zyp + s.zyp) - (zym + s.zym);       // * 4 h_x, 4 h_y
      const T sxx = s.dxx0 * K.rh[0], syy = s.dyy0 * K.rh[1], szz = (s.uz0 - s.uzm) * K.rh[2];
      const T sxy2 = gxy * K.q[1] + gyx * K.q[0];   // 2 S_xy
      const T sxz2 = gxz * K.q[2] + gzx * K.q[0];
      const T syz2 = gyz * K.q[2] + gzy * K.q[1];
      const T ss = (sxx * sxx + syy * syy + szz * szz) + T(0.5) * (sxy2 * sxy2 + sxz2 * sxz2 + syz2 * syz2);
      const T e2 = K.e2c * fs_sqrt(T(2) * ss);      // 2 nu_t
      const T eh = T(0.5) * e2;
      out[0] = e2 * sxx;
      out[SPL] = eh * sxy2;
      out[2 * SPL] = eh * sxz2;
      out[3 * SPL] = e2 * syy;
      out[4 * SPL] = eh * syz2;
      out[5 * SPL] = e2 * szz;
    }
    s.zxp = zxp; s.zxm = zxm; s.zyp = zyp; s.zym = zym;
  }
  s.uxm = s.ux0; s.ux0 = uxt;
  s.uym = s.uy0; s.uy0 = uyt;
  s.uzm = s.uz0; s.uz0 = uzt;
  s.dxx0 = dxxt; s.dyy0 = dyyt;
}

// ---- phase C, closure part: s = div sigma at this thread's cell ---------------------------------------------------
// Gm / Gt: this cell in the sigma planes r / r+1.  r = -2: top loads only; r = -1: s_z only.
template <typename T, typename Cfg, bool ST>
FS_HD void fs_divsigma(FsCellSg<T> &c, const T *Gm, const T *Gt, int r, const FastK<T> &K, T (&s)[3]) {
  constexpr int SX = Cfg::SX, SPL = Cfg::SPL;
  const T xz0 = Gt[2 * SPL], xzp = Gt[2 * SPL + 1], xzm = Gt[2 * SPL - 1];
  const T yz0 = Gt[4 * SPL], yzp = Gt[4 * SPL + SX], yzm = Gt[4 * SPL - SX];
  const T zzt = Gt[5 * SPL];
  const T sxzt = xz0 + xzp, syzt = yz0 + yzp, dxzt = xzp - xzm, dyzt = yzp - yzm;
  s[0] = s[1] = s[2] = T(0);
  if (ST || r >= -1) {
    s[2] = (zzt - c.szz0) * K.rh[2] + (c.dxz0 + dxzt) * K.q[0] + (c.dyz0 + dyzt) * K.q[1];
    if (ST || r >= 0) {
      const T xypp = Gm[SPL + SX + 1];
      s[0] = (Gm[1] - Gm[0]) * K.rh[0] + ((Gm[SPL + SX] + xypp) - (Gm[SPL - SX] + Gm[SPL - SX + 1])) * K.q[1] +
             (sxzt - c.sxzm) * K.q[2];
      s[1] = (Gm[3 * SPL + SX] - Gm[3 * SPL]) * K.rh[1] + ((Gm[SPL + 1] + xypp) - (Gm[SPL - 1] + Gm[SPL + SX - 1])) * K.q[0] +
             (syzt - c.syzm) * K.q[2];
    }
  }
  c.sxzm = c.sxz0; c.sxz0 = sxzt;
  c.syzm = c.syz0; c.syz0 = syzt;
  c.dxz0 = dxzt; c.dyz0 = dyzt; c.szz0 = zzt;
}

// one component (a = 0: x at a point of the x-halo, a = 1: y at a point of the y-halo) of div sigma
template <typename T, typename Cfg, bool ST>
FS_HD T fs_divsigma_halo(FsHalo<T> &h, const int a, const T *Gm, const T *Gt, int r, const FastK<T> &K) {
  constexpr int SX = Cfg::SX, SPL = Cfg::SPL;
  const int d = a == 0 ? 1 : SX;            // +e_a
  const int e = a == 0 ? SX : 1;            // the other in-plane direction
  const int caz = a == 0 ? 2 * SPL : 4 * SPL, caa = a == 0 ? 0 : 3 * SPL;
  const T st = Gt[caz] + Gt[caz + d];
  T s = T(0);
  if (ST || r >= 0)
    s = (Gm[caa + d] - Gm[caa]) * K.rh[a] + ((Gm[SPL + e] + Gm[SPL + e + d]) - (Gm[SPL - e] + Gm[SPL - e + d])) * K.q[1 - a] +
        (st - h.sm) * K.q[2];
  h.sm = h.s0;
  h.s0 = st;
  return s;
}

// ---- phase C, convection-diffusion of this thread's cell ----------------------------------------------------------
// M / Tp: this cell in the velocity planes r / r+1.  r = -3: top loads only; r = -2: z-fluxes only; r = -1: F_z only.
// uc receives the cell's own velocity at plane r (ustart of the first stage).
template <typename T, typename Cfg, bool ST>
FS_HD void fs_convdiff(FsCell<T> &c, const T *M, const T *Tp, int r, const FastK<T> &K, T (&F)[3], T (&uc)[3]) {
  constexpr int BX = Cfg::BX, UPL = Cfg::UPL;
  const T tx0 = Tp[0], txm = Tp[-1], ty0 = Tp[UPL], tym = Tp[UPL - BX], tz0 = Tp[2 * UPL];
  F[0] = F[1] = F[2] = T(0);
  uc[0] = c.cx0; uc[1] = c.cy0; uc[2] = c.cz0;
  if (ST || r >= -2) {
    const T zxp = M[2 * UPL + 1], zyp = M[2 * UPL + BX];
    const T szx = c.cz0 + zxp, szy = c.cz0 + zyp;
    const T pzx = szx * (c.cx0 + tx0), pzy = szy * (c.cy0 + ty0), c1 = c.cz0 + tz0;
    const T gx = K.vrr[2] * (tx0 - c.cx0) - K.q[2] * pzx;
    const T gy = K.vrr[2] * (ty0 - c.cy0) - K.q[2] * pzy;
    const T gz = K.vrr[2] * (tz0 - c.cz0) - K.q[2] * (c1 * c1);
    if (ST || r >= -1) {
      const T zxm = M[2 * UPL - 1], zym = M[2 * UPL - BX];
      const T two_z = c.cz0 + c.cz0;
      const T conv_zx = pzx - (zxm + c.cz0) * (c.cxm + txm);
      const T conv_zy = pzy - (zym + c.cz0) * (c.cym + tym);
      F[2] = (gz - c.fzz) + (K.vrr[0] * ((zxp + zxm) - two_z) - K.q[0] * conv_zx) + (K.vrr[1] * ((zyp + zym) - two_z) - K.q[1] * conv_zy);
      if (ST || r >= 0) {
        const T xxp = M[1], xyp = M[BX], xym = M[-BX], xmy = M[BX - 1];
        const T yyp = M[UPL + BX], yxp = M[UPL + 1], yxm = M[UPL - 1], ypm = M[UPL + 1 - BX];
        const T a1 = c.cx0 + xxp, a2 = c.cxm + c.cx0, b1 = c.cy0 + yyp, b2 = c.cym + c.cy0;
        const T pxy = (c.cx0 + xyp) * (c.cy0 + yxp);
        const T conv_xx = a1 * a1 - a2 * a2;
        const T conv_xy = pxy - (xym + c.cx0) * (c.cym + ypm);
        const T conv_yx = pxy - (yxm + c.cy0) * (c.cxm + xmy);
        const T conv_yy = b1 * b1 - b2 * b2;
        const T two_x = c.cx0 + c.cx0, two_y = c.cy0 + c.cy0;
        F[0] = (gx - c.fzx) + (K.vrr[0] * ((xxp + c.cxm) - two_x) - K.q[0] * conv_xx) + (K.vrr[1] * ((xyp + xym) - two_x) - K.q[1] * conv_xy);
        F[1] = (gy - c.fzy) + (K.vrr[0] * ((yxp + yxm) - two_y) - K.q[0] * conv_yx) + (K.vrr[1] * ((yyp + c.cym) - two_y) - K.q[1] * conv_yy);
      }
    }
    c.fzx = gx; c.fzy = gy; c.fzz = gz;
  }
  c.cx0 = tx0; c.cxm = txm; c.cy0 = ty0; c.cym = tym; c.cz0 = tz0;
}

// one component (a = 0 or 1) of the convection-diffusion at a halo point; only the z-flux is carried
template <typename T, typename Cfg, bool ST>
FS_HD T fs_convdiff_halo(FsHalo<T> &h, const int a, const T *M, const T *Tp, int r, const FastK<T> &K, T &uc) {
  constexpr int BX = Cfg::BX, UPL = Cfg::UPL;
  const int o = a * UPL, ob = (1 - a) * UPL;       // own component, the other in-plane component
  const int d = a == 0 ? 1 : BX, e = a == 0 ? BX : 1;  // +e_a, +e_b (b = the other in-plane direction)
  const T u0 = M[o], ut = Tp[o];
  uc = u0;
  T F = T(0);
  if (ST || r >= -2) {
    const T g = K.vrr[2] * (ut - u0) - K.q[2] * ((M[2 * UPL] + M[2 * UPL + d]) * (u0 + ut));
    if (ST || r >= 0) {
      const T up = M[o + d], um = M[o - d], vp = M[o + e], vm = M[o - e];
      const T a1 = u0 + up, a2 = um + u0;
      const T conv_aa = a1 * a1 - a2 * a2;
      const T conv_ab = (u0 + vp) * (M[ob] + M[ob + d]) - (vm + u0) * (M[ob - e] + M[ob - e + d]);
      const T two = u0 + u0;
      F = (g - h.fz) + (K.vrr[a] * ((up + um) - two) - K.q[a] * conv_aa) + (K.vrr[1 - a] * ((vp + vm) - two) - K.q[1 - a] * conv_ab);
    }
    h.fz = g;
  }
  return F;
}

// ---- one thread of the block ---------------------------------------------------------------------------------------
// All per-thread variables of the march and the two halves of an iteration: pre() = phases A, B, C up to the block
// barrier, post() = the divergence after it.  The device kernel keeps one FsThread in registers; the host emulator
// (tools/fast_stage_emul.cu) keeps an array of them and runs the threads one after the other between barriers.
template <typename T>
FS_HD T fs_ld(const T *base, size_t ob) {
  return FS_LDG(reinterpret_cast<const T *>(reinterpret_cast<const char *>(base) + ob));
}
template <typename T>
FS_HD void fs_st(T *base, size_t ob, T v) {
  *reinterpret_cast<T *>(reinterpret_cast<char *>(base) + ob) = v;
}
FS_HD int fs_wrap(int i, int n, int wrap) {
  if (!wrap) return i;
  return i < 0 ? i + n : (i >= n ? i - n : i);
}

// Block-uniform shared-memory offsets (elements) of the planes one iteration touches.  Velocity plane q lives in slot
// q mod 8, pressure plane q in slot q mod 8, sigma plane rB in slot rB mod 4, the staging rows alternate.  With r the
// iteration (stencil plane), q = r + 2:
//   u_mid / u_top  velocity planes r, r+1 (stencil)      u_bm / u_bt  velocity planes r+2, r+3 (sigma)
//   u_a            velocity plane r + LEAD (projected)    p_a / p_b    its pressure planes q_A, q_A + 1
//   g_w            sigma plane r+2 (written)              g_m / g_t    sigma planes r, r+1 (read)
// make(r): general form (start-up and tail iterations).  steady<J>(H, HP): r = 4 g + J inside the 4x unrolled steady
// loop -- every offset is an IMMEDIATE plus or minus H = (g & 1) * 4 * USLOT (HP likewise for the pressure ring),
// because (c + 4 g) mod 8 = c xor 4 (g & 1): the loop carries two uniform registers instead of ten slot counters.
template <typename Cfg>
struct FsOffs {
  int u_mid, u_top, u_bm, u_bt, u_a, p_a, p_b, g_w, g_m, g_t, stg;
  static FS_HD FsOffs make(int r) {
    FsOffs o;
    const int q = r + 2 + 8;  // >= 0 for every r >= R0
    o.u_mid = (q & 7) * Cfg::USLOT;
    o.u_top = ((q + 1) & 7) * Cfg::USLOT;
    o.u_bm = ((q + 2) & 7) * Cfg::USLOT;
    o.u_bt = ((q + 3) & 7) * Cfg::USLOT;
    o.u_a = ((q + Cfg::LEAD) & 7) * Cfg::USLOT;
    o.p_a = Cfg::POFF + ((q + Cfg::LEAD) & 7) * Cfg::PSLOT;
    o.p_b = Cfg::POFF + ((q + Cfg::LEAD + 1) & 7) * Cfg::PSLOT;
    o.g_w = Cfg::SOFF + ((r + 2 + 8) & 3) * Cfg::SSLOT;
    o.g_m = Cfg::SOFF + ((r + 8) & 3) * Cfg::SSLOT;
    o.g_t = Cfg::SOFF + ((r + 1 + 8) & 3) * Cfg::SSLOT;
    o.stg = Cfg::DOFF + (r & 1) * Cfg::DXN;
    return o;
  }
  template <int J>
  static FS_HD FsOffs steady(const int H, const int HP) {
    FsOffs o;
    constexpr int c0 = (2 + J) & 7, c1 = (3 + J) & 7, c2 = (4 + J) & 7, c3 = (5 + J) & 7, ca = (2 + J + Cfg::LEAD) & 7, cb = (3 + J + Cfg::LEAD) & 7;
    o.u_mid = c0 * Cfg::USLOT + (c0 < 4 ? H : -H);
    o.u_top = c1 * Cfg::USLOT + (c1 < 4 ? H : -H);
    o.u_bm = c2 * Cfg::USLOT + (c2 < 4 ? H : -H);
    o.u_bt = c3 * Cfg::USLOT + (c3 < 4 ? H : -H);
    o.u_a = ca * Cfg::USLOT + (ca < 4 ? H : -H);
    o.p_a = Cfg::POFF + ca * Cfg::PSLOT + (ca < 4 ? HP : -HP);
    o.p_b = Cfg::POFF + cb * Cfg::PSLOT + (cb < 4 ? HP : -HP);
    o.g_w = Cfg::SOFF + ((J + 2) & 3) * Cfg::SSLOT;
    o.g_m = Cfg::SOFF + (J & 3) * Cfg::SSLOT;
    o.g_t = Cfg::SOFF + ((J + 1) & 3) * Cfg::SSLOT;
    o.stg = Cfg::DOFF + (J & 1) * Cfg::DXN;
    return o;
  }
};

template <typename T, bool PGRAD, bool SMAG, bool GEN>
struct FsThread {
  typedef FastCfg<T, SMAG> Cfg;
  static constexpr int TX = Cfg::TX, TY = Cfg::TY, NT = Cfg::NT, BX = Cfg::BX, UPL = Cfg::UPL, XL = Cfg::XL;
  static constexpr int SX = Cfg::SX, SY = Cfg::SY, SPL = Cfg::SPL, DOFF = Cfg::DOFF;
  static constexpr int LEAD = Cfg::LEAD;
  static constexpr bool U0TILE = !PGRAD;
  static constexpr int NKP = GEN ? 3 : 1;  // earlier k_j prefetched in registers (RK44's last row has 3)

  // block-wide
  T *su;
  const TmaStageArgs<T> *A;
  const FastK<T> *K;
  int kbeg, nk, n2, wrap2, qlast;
  bool has_force, has_kout, have_prev;  // force and "k_i not stored" only exist in the GEN kernels (the host picks them)
  int nkin;
  size_t szb;
  // roles and offsets
  int tid;
  bool cell_thr, hx, hy, hany, sig_thr;
  int ha, c_own, c_halo, c_sigu, c_sigw, c_sig, c_sigh;
  int d_own_x, d_own_y, d_halo;  // staging rows: this cell's u'_x / u'_y entries; the halo thread's entry
  size_t ob, obh;  // byte offsets of this thread's cell / halo point in plane max(r, 0)
  // march state
  FsCell<T> cs;
  FsCellSg<T> cg;
  FsSig<T> sg;
  FsHalo<T> hs;
  T un[3];
  T n_u0[3], n_f[3], n_k[3][NKP], nh_u0, nh_f, nh_k[NKP];

  FS_HD size_t plane_off(int r) const { return (size_t)(r < 0 ? fs_wrap(kbeg + r, n2, wrap2) : kbeg + r) * szb; }

  // `warp`: the warp index of this thread, handed in as a WARP-UNIFORM value (device: read from lane 0 by a shuffle),
  // so that every role test below is a uniform branch
  FS_HD void init(T *su_, const TmaStageArgs<T> *A_, const FastK<T> *K_, const Geo &g, int zc, int tid_, int warp, int bx_, int by_, int bz_) {
    su = su_; A = A_; K = K_;
    tid = tid_;
    const int lane = tid & 31;
    const int ty = warp, tx = lane;  // TX == 32: a warp is one row of the tile
    const int i0 = bx_ * TX, j0 = by_ * TY;
    kbeg = bz_ * zc;
    nk = zc < g.n[2] - kbeg ? zc : g.n[2] - kbeg;
    n2 = g.n[2]; wrap2 = g.wrap[2];
    qlast = nk + (SMAG ? 3 : 2);
    has_force = GEN && A->force[0] != nullptr;
    has_kout = !GEN || A->kout[0] != nullptr;
    nkin = GEN ? A->nk : 0;
    szb = (size_t)A->sz * sizeof(T);
    have_prev = kbeg > 0 || g.wrap[2];  // otherwise plane 0's divergence is completed by k_divergence_planes
    cell_thr = warp < TY;
    const int hw = SMAG ? 16 : 1;  // warp of the x-halo; the y-halo is the next one
    hx = warp == hw && lane < TY;
    hy = warp == hw + 1;
    hany = hx || hy;
    const int hlx = hx ? -1 : lane, hly = hx ? lane : -1;
    ha = hx ? 0 : 1;
    c_own = (ty + 2) * BX + tx + XL;
    c_halo = hany ? (hly + 2) * BX + hlx + XL : c_own;
    const size_t ocell = ((size_t)(j0 + ty) * A->sy + (i0 + tx)) * sizeof(T);
    const int hgi = fs_wrap(i0 + hlx, g.n[0], 1), hgj = fs_wrap(j0 + hly, g.n[1], 1);
    const size_t ohalo = hany ? ((size_t)hgj * A->sy + hgi) * sizeof(T) : ocell;
    // sigma tile (SX x SY p-cells, x = -1 .. TX, y = -1 .. TY): warps 0 .. SY-1 own the columns 0 .. 31 of one row each
    // (conflict-free rows), the two remaining columns of all rows go to the next 2 SY threads
    int slx = lane, sly = warp;
    if (warp >= SY) { const int e = tid - 32 * SY; sly = e >> 1; slx = 32 + (e & 1); }
    sig_thr = SMAG && tid < SPL;
    c_sigu = sig_thr ? (sly + 1) * BX + slx - 1 + XL : BX + XL;  // p-cell (slx-1, sly-1) in the velocity box
    c_sigw = sig_thr ? sly * SX + slx : 0;                        // ... and in the sigma tile
    c_sig = (ty + 1) * SX + tx + 1;                               // this thread's cell in the sigma tile
    c_sigh = hany ? (hly + 1) * SX + hlx + 1 : c_sig;
    d_own_x = ty * (TX + 1) + tx + 1;
    d_own_y = 2 * Cfg::DXN + (ty + 1) * TX + tx;
    d_halo = hx ? hly * (TX + 1) : 2 * Cfg::DXN + hlx;
    fs_zero(reinterpret_cast<T *>(&cs), (int)(sizeof(cs) / sizeof(T)));
    fs_zero(reinterpret_cast<T *>(&cg), (int)(sizeof(cg) / sizeof(T)));
    fs_zero(reinterpret_cast<T *>(&sg), (int)(sizeof(sg) / sizeof(T)));
    fs_zero(reinterpret_cast<T *>(&hs), (int)(sizeof(hs) / sizeof(T)));
    for (int a = 0; a < 3; a++) {
      un[a] = n_u0[a] = n_f[a] = T(0);
      for (int q = 0; q < NKP; q++) n_k[a][q] = T(0);
    }
    nh_u0 = nh_f = T(0);
    for (int q = 0; q < NKP; q++) nh_k[q] = T(0);
    ob = plane_off(0) + ocell;
    obh = plane_off(0) + ohalo;
    if (have_prev) {
      if (cell_thr) fetch_operands<true, !SMAG>(plane_off(-1) + ocell, plane_off(-1) + ohalo);
      else fetch_operands<false, SMAG>(plane_off(-1) + ocell, plane_off(-1) + ohalo);
    }
  }

  // global operands (ustart, force, earlier k_j) of one plane into registers
  template <bool CELL, bool HALO>
  FS_HD void fetch_operands(const size_t o1, const size_t o1h) {
    if (CELL) {
#pragma unroll
      for (int a = 0; a < 3; a++) {
        if (!U0TILE) n_u0[a] = fs_ld(A->u0[a], o1);
        if (GEN && has_force) n_f[a] = fs_ld(A->force[a], o1);
#pragma unroll
        for (int q = 0; q < NKP; q++)
          if (GEN && q < nkin) n_k[a][q] = fs_ld(A->kin[q][a], o1);
      }
    }
    if (HALO && hany) {
      if (!U0TILE) nh_u0 = fs_ld(hx ? A->u0[0] : A->u0[1], o1h);
      if (GEN && has_force) nh_f = fs_ld(hx ? A->force[0] : A->force[1], o1h);
#pragma unroll
      for (int q = 0; q < NKP; q++)
        if (GEN && q < nkin) nh_k[q] = fs_ld(hx ? A->kin[q][0] : A->kin[q][1], o1h);
    }
  }

  // k_i and u' of one component from f = convection-diffusion + closure (at most NKP earlier k_j: the host takes the
  // exact kernels for longer tableau rows)
  FS_HD T combine(T f, const T uc, const T l_u0, const T l_f, const T (&l_k)[NKP], T &f_out) const {
    if (GEN && has_force) f = f + l_f;
    f_out = f;
    T acc = U0TILE ? uc : l_u0;
    if (GEN) {
#pragma unroll
      for (int q = 0; q < NKP; q++)
        if (q < nkin) acc = acc + A->coef[q] * l_k[q];
    }
    return acc + A->coef_self * f;
  }

  // is the velocity plane of phase A (q = r + LEAD + 2) part of this block's march?
  FS_HD bool project_active(int r) const { return r + LEAD + 2 <= qlast; }

  // phases A (after the plane's barrier has completed), B, C.  ST: steady state (r >= 0).  CELL: this warp owns a row
  // of cells (compile-time role: the cell warps and the other warps run separate copies of the loop, so that no role
  // test -- and no register shuffling around a skipped block -- remains inside it).  The halo points ride on the cell
  // warps 1 / 2 without closure and on the warps 16 / 17 with it.
  template <bool ST, bool CELL>
  FS_HD void pre(const int r, const FsOffs<Cfg> &S) {
    constexpr bool HALO = SMAG ? !CELL : CELL;
    // ---- A: velocity plane r + LEAD is projected in place
    // (load balance: with the closure the four helper warps 16 .. 19 project, the cell warps have the stencil; without
    // it the vectors start at warp 3, because warps 1 and 2 carry the halo points)
    if (PGRAD && !(SMAG && CELL) && project_active(r))
      fs_project<T, Cfg>(su + S.u_a, su + S.p_a, su + S.p_b, SMAG ? tid - 32 * TY : (tid >= 96 ? tid - 96 : tid + NT - 96),
                         SMAG ? NT - 32 * TY : NT, K->rh);
    // ---- B: sigma(r + 2) at this thread's p-cell.  In the steady state the planes beyond the chunk are "computed" as
    // well (from stale slots, into ring slots nobody reads any more): cheaper than a test around the carried registers
    if (SMAG) {
      const int rB = r + 2;
      if ((CELL || sig_thr) && (ST || (rB >= -3 && rB <= nk)))
        fs_sigma<T, Cfg, ST>(sg, su + S.u_bm + c_sigu, su + S.u_bt + c_sigu, su + S.g_w + c_sigw, rB, *K);
    }
    // ---- C: stencil of plane r
    if (ST || r >= -3) {
      T *stg = su + S.stg;                                                  // staging rows of this parity: u'_x block ...
      T *stgy = stg + ((S.stg - DOFF) / Cfg::DXN) * (Cfg::DYN - Cfg::DXN);  // ... and (+ 2 DXN, see d_own_y) the u'_y block
      if (CELL) {
        T F[3], uc[3], s[3] = {T(0), T(0), T(0)};
        fs_convdiff<T, Cfg, ST>(cs, su + S.u_mid + c_own, su + S.u_top + c_own, r, *K, F, uc);
        if (SMAG && (ST || r >= -2)) fs_divsigma<T, Cfg, ST>(cg, su + S.g_m + c_sig, su + S.g_t + c_sig, r, *K, s);
        if (ST || r >= 0) {
#pragma unroll
          for (int a = 0; a < 3; a++) {
            T f;
            un[a] = combine(F[a] + s[a], uc[a], n_u0[a], n_f[a], n_k[a], f);
            if (has_kout) fs_st(A->kout[a], ob, f);
            fs_st(A->unew[a], ob, un[a]);
          }
          stg[d_own_x] = un[0];
          stgy[d_own_y] = un[1];
        } else if (r == -1 && have_prev) {  // u'_z of the plane below the chunk, for the divergence of its first plane
          T f;
          cs.uzp = combine(F[2] + s[2], uc[2], n_u0[2], n_f[2], n_k[2], f);
        }
      }
      if (HALO && hany && (ST || r >= -2)) {
        T uc;
        T F = fs_convdiff_halo<T, Cfg, ST>(hs, ha, su + S.u_mid + c_halo, su + S.u_top + c_halo, r, *K, uc);
        if (SMAG) F = F + fs_divsigma_halo<T, Cfg, ST>(hs, ha, su + S.g_m + c_sigh, su + S.g_t + c_sigh, r, *K);
        if (ST || r >= 0) {
          T f;
          const T v = combine(F, uc, nh_u0, nh_f, nh_k, f);
          (hx ? stg : stgy)[d_halo] = v;
        }
      }
      // operands of plane r + 1 for the next iteration
      if ((ST || r >= -1) && r + 1 < nk) {
        const size_t step = (ST || r >= 0) ? szb : 0;  // ob / obh stay on plane 0 until r = 0 (post())
        fetch_operands<CELL, HALO>(ob + step, obh + step);
      }
    }
  }

  // after the block barrier: divergence of u' (the neighbours' u'_x, u'_y are in the staging rows)
  template <bool ST, bool CELL>
  FS_HD void post(const int r, const FsOffs<Cfg> &S) {
    if (CELL && (ST || r >= 0)) {
      const T *stg = su + S.stg;
      const T *stgy = stg + ((S.stg - DOFF) / Cfg::DXN) * (Cfg::DYN - Cfg::DXN);
      if (r > 0 || have_prev) {
        T d = (un[0] - stg[d_own_x - 1]) * K->rh[0];
        d += (un[1] - stgy[d_own_y - TX]) * K->rh[1];
        d += (un[2] - cs.uzp) * K->rh[2];
        fs_st(A->div, ob, d * K->vol);
      }
      cs.uzp = un[2];
    }
    if (ST || r >= 0) { ob += szb; obh += szb; }
  }
};

// ---- the kernel ---------------------------------------------------------------------------------------------------
#if defined(__CUDACC__)
template <typename T, bool PGRAD, bool SMAG, bool GEN>
__global__ void __launch_bounds__(FastCfg<T, SMAG>::NT, ((SMAG || GEN || sizeof(T) == 8) ? 1 : 2))
    k_stage3d_fast(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_p,
                   const __grid_constant__ TmaStageArgs<T> A, const __grid_constant__ FastK<T> K, Geo g, int zc, int ox, int oy,
                   int oz) {
  typedef FastCfg<T, SMAG> Cfg;
  constexpr int NS = Cfg::NS, NPS = Cfg::NPS;
  extern __shared__ unsigned char smem_raw[];
  unsigned char *sb = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  T *su = reinterpret_cast<T *>(sb);
  uint64_t *bars = reinterpret_cast<uint64_t *>((reinterpret_cast<uintptr_t>(su + Cfg::ELEMS) + 7) & ~(uintptr_t)7);
  const int tid = threadIdx.y * Cfg::TX + threadIdx.x;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // warp-uniform by construction: role tests become uniform branches

  FsThread<T, PGRAD, SMAG, GEN> t;
  t.init(su, &A, &K, g, zc, tid, warp, blockIdx.x, blockIdx.y, blockIdx.z);

  // velocity plane q <-> z = kbeg - 2 + q; q = 0 .. qlast are loaded, NS slots, one mbarrier per slot
  const int x0 = ox + blockIdx.x * Cfg::TX - Cfg::XL, y0 = oy + blockIdx.y * Cfg::TY - 2, z0 = oz + t.kbeg - 2;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < NS; i++) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  __syncthreads();
  auto issue = [&](int q) {  // one thread: velocity plane q (3 components) + pressure plane q+1 (and plane 0 with q = 0)
    uint64_t *bar = &bars[q % NS];
    fence_proxy_async();  // earlier generic-proxy accesses of the slot are ordered before the bulk write
    if (PGRAD) {
      mbar_expect_tx(bar, Cfg::UBYTES + (q == 0 ? 2u : 1u) * Cfg::PBYTES);
      if (q == 0) tma_load_4d(su + Cfg::POFF, &tm_p, bar, x0, y0, z0, 0);
      tma_load_4d(su + Cfg::POFF + ((q + 1) % NPS) * Cfg::PSLOT, &tm_p, bar, x0, y0, z0 + q + 1, 0);
    } else {
      mbar_expect_tx(bar, Cfg::UBYTES);
    }
    tma_load_4d(su + (q % NS) * Cfg::USLOT, &tm_u, bar, x0, y0, z0 + q, 0);
  };
  if (tid == 0) {  // NS - 1 planes: the pressure ring holds planes 0 .. NS-1 at this point
#pragma unroll
    for (int q = 0; q < NS - 1; q++)
      if (q <= t.qlast) issue(q);
  }
  // One iteration: wait for plane r + LEAD, phases A / B / C, block barrier, refill (thread 0), divergence.  After the
  // barrier velocity plane r (q = r + 2) is retired and its slot takes plane q + NS (the very first iteration frees
  // pressure slot 0 instead, which plane NS - 1 was waiting for).  The cell warps and the other warps (closure only:
  // sigma tile, halo points) run their own copy of the loop and meet at the same hardware barrier.
  auto march = [&](auto cellc) {
    constexpr bool CELL = decltype(cellc)::value != 0;
    auto barrier_of = [&](int r) -> uint64_t * { return &bars[(r + Cfg::LEAD + 2) & (NS - 1)]; };
    auto parity_of = [&](int r) -> uint32_t { return (uint32_t)(((r + Cfg::LEAD + 2) >> 3) & 1); };
    int r = Cfg::R0;
    for (; r < 0; r++) {  // start-up: the pipeline fills, carried values are set up (general code, a few iterations)
      const FsOffs<Cfg> S = FsOffs<Cfg>::make(r);
      if (t.project_active(r)) mbar_wait(barrier_of(r), parity_of(r));  // plane r + LEAD has landed
      t.template pre<false, CELL>(r, S);
      asm volatile("bar.sync 0;" ::: "memory");
      if (CELL && tid == 0) {
        if (r == Cfg::R0 && NS - 1 <= t.qlast) issue(NS - 1);
        if (r >= -2 && r + 2 + NS <= t.qlast) issue(r + 2 + NS);
      }
      t.template post<false, CELL>(r, S);
    }
    // steady state, 4 planes per trip: r = 4 g + J, all shared-memory offsets are immediates +- H (FsOffs::steady)
    int H = 0, HP = 0;
    auto one = [&](auto jc) {
      constexpr int J = decltype(jc)::value;
      const FsOffs<Cfg> S = FsOffs<Cfg>::template steady<J>(H, HP);
      if (t.project_active(r)) mbar_wait(barrier_of(r), parity_of(r));
      t.template pre<true, CELL>(r, S);
      asm volatile("bar.sync 0;" ::: "memory");
      if (CELL && tid == 0 && r + 2 + NS <= t.qlast) issue(r + 2 + NS);
      t.template post<true, CELL>(r, S);
      r++;
    };
    for (; r + 4 <= t.nk;) {
      one(IntC<0>{});
      one(IntC<1>{});
      one(IntC<2>{});
      one(IntC<3>{});
      H = 4 * Cfg::USLOT - H;
      HP = 4 * Cfg::PSLOT - HP;
    }
    for (; r < t.nk; r++) {  // tail (fewer than 4 planes left): general offsets
      const FsOffs<Cfg> S = FsOffs<Cfg>::make(r);
      if (t.project_active(r)) mbar_wait(barrier_of(r), parity_of(r));
      t.template pre<true, CELL>(r, S);
      asm volatile("bar.sync 0;" ::: "memory");
      if (CELL && tid == 0 && r + 2 + NS <= t.qlast) issue(r + 2 + NS);
      t.template post<true, CELL>(r, S);
    }
  };
  if (!SMAG || warp < Cfg::TY) march(IntC<1>{});
  else march(IntC<0>{});
}
#endif  // __CUDACC__

}  // namespace insb200
