// zsolve.cuh -- last-axis part of the spectral pressure solve in ONE pass over the half spectrum:
//   forward FFT along z  ->  p^ = -p^ / ((ax+ay)+az) * (1/N), zero mode -> 0  ->  inverse FFT along z
// (src/pressure.jl:323-344: mul!(phat, plan, pI); phat = -phat/(ax+ay+az); phat[1] = 0; ldiv!(pI, plan, phat)).
// cuFFT does the batched in-plane (x, y) transforms; this kernel replaces cuFFT's z pass of the forward
// transform, the symbol-scaling kernel and the z pass of the inverse transform (3 read+write passes over
// the spectrum -> 1).
//
// A block owns COLS consecutive columns c = ky*nxw + kx of the spectrum (contiguous in memory for every z)
// and all nz points of each: the tile [nz][COLS] sits in shared memory.  Forward = in-place
// decimation-in-frequency passes of radix 16/8/4/2 (natural order in, digit-reversed order out); the symbol
// is applied at digit-reversed positions; inverse = the adjoint passes in reverse order (digit-reversed in,
// natural out).  No reordering pass, no second buffer.  Twiddles exp(-2 pi i t / nz) come from a table
// computed in Float64 on the host.
#pragma once
#include "ctx.cuh"
#include "tma_stage.cuh"

namespace insb200 {

__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *tm, uint64_t *bar, int x, int y) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                   smem_u32(dst)),
               "l"(reinterpret_cast<uint64_t>(tm)), "r"(x), "r"(y), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap *tm, const void *src, int x, int y) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(reinterpret_cast<uint64_t>(tm)),
               "r"(x), "r"(y), "r"(smem_u32(src))
               : "memory");
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *tm, uint64_t *bar, int x, int y, int z) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                   smem_u32(dst)),
               "l"(reinterpret_cast<uint64_t>(tm)), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap *tm, const void *src, int x, int y, int z) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(reinterpret_cast<uint64_t>(tm)),
               "r"(x), "r"(y), "r"(z), "r"(smem_u32(src))
               : "memory");
}
__device__ __forceinline__ void tma_store_commit_wait() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

template <typename T> struct Cx { T x, y; };
template <typename T> __device__ __forceinline__ Cx<T> cadd(Cx<T> a, Cx<T> b) { return {a.x + b.x, a.y + b.y}; }
template <typename T> __device__ __forceinline__ Cx<T> csub(Cx<T> a, Cx<T> b) { return {a.x - b.x, a.y - b.y}; }
// a * b, and a * conj(b)
template <typename T, bool CONJ> __device__ __forceinline__ Cx<T> cmul(Cx<T> a, Cx<T> b) {
  if (CONJ) return {a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y};
  return {a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x};
}
// multiply by -i (forward) or +i (inverse)
template <typename T, bool INV> __device__ __forceinline__ Cx<T> mul_mi(Cx<T> a) {
  if (INV) return {-a.y, a.x};
  return {a.y, -a.x};
}

// in-register R-point DFT, forward: y_q = sum_m x_m exp(-2 pi i m q / R); INV: conjugate kernel
template <typename T, bool INV> __device__ __forceinline__ void dft2(Cx<T> (&v)[2]) {
  const Cx<T> a = v[0], b = v[1];
  v[0] = cadd(a, b);
  v[1] = csub(a, b);
}
template <typename T, bool INV> __device__ __forceinline__ void dft4(Cx<T> (&v)[4]) {
  const Cx<T> a = cadd(v[0], v[2]), b = csub(v[0], v[2]), c = cadd(v[1], v[3]), d = mul_mi<T, INV>(csub(v[1], v[3]));
  v[0] = cadd(a, c);
  v[1] = cadd(b, d);
  v[2] = csub(a, c);
  v[3] = csub(b, d);
}
template <typename T, bool INV> __device__ __forceinline__ void dft8(Cx<T> (&v)[8]) {
  // two radix-4 on even / odd inputs, then the radix-2 combination with w8^q
  Cx<T> e[4] = {v[0], v[2], v[4], v[6]}, o[4] = {v[1], v[3], v[5], v[7]};
  dft4<T, INV>(e);
  dft4<T, INV>(o);
  const T h = T(0.70710678118654752440);
  // w8^1 = (1 - i)/sqrt2 (forward), (1 + i)/sqrt2 (inverse); w8^2 = -i / +i; w8^3 = (-1 - i)/sqrt2 / (-1 + i)/sqrt2
  const Cx<T> o1 = INV ? Cx<T>{(o[1].x - o[1].y) * h, (o[1].x + o[1].y) * h} : Cx<T>{(o[1].x + o[1].y) * h, (o[1].y - o[1].x) * h};
  const Cx<T> o2 = mul_mi<T, INV>(o[2]);
  const Cx<T> o3 = INV ? Cx<T>{(-o[3].x - o[3].y) * h, (o[3].x - o[3].y) * h} : Cx<T>{(o[3].y - o[3].x) * h, (-o[3].x - o[3].y) * h};
  v[0] = cadd(e[0], o[0]); v[4] = csub(e[0], o[0]);
  v[1] = cadd(e[1], o1);   v[5] = csub(e[1], o1);
  v[2] = cadd(e[2], o2);   v[6] = csub(e[2], o2);
  v[3] = cadd(e[3], o3);   v[7] = csub(e[3], o3);
}
// 16 points as 4 x 4 (n = 4a + b, k = q1 + 4 q2):  w16^{nk} = w4^{a q1} * w16^{b q1} * w4^{b q2}
template <typename T, bool INV> __device__ __forceinline__ void dft16(Cx<T> (&v)[16]) {
  const T c1 = T(0.92387953251128675613), s1 = T(0.38268343236508977173), h = T(0.70710678118654752440);
  // w16^k = exp(-2 pi i k / 16) for k = b * q1 (cmul<.., INV> multiplies by the conjugate for the inverse)
  const Cx<T> w[10] = {{T(1), T(0)}, {c1, -s1}, {h, -h}, {s1, -c1}, {T(0), T(-1)}, {-s1, -c1}, {-h, -h}, {-c1, -s1}, {T(-1), T(0)}, {-c1, s1}};
  Cx<T> t[4][4];  // t[b][q1]
#pragma unroll
  for (int b = 0; b < 4; b++) {
    Cx<T> x[4] = {v[b], v[4 + b], v[8 + b], v[12 + b]};
    dft4<T, INV>(x);
#pragma unroll
    for (int q1 = 0; q1 < 4; q1++) t[b][q1] = (b == 0 || q1 == 0) ? x[q1] : cmul<T, INV>(x[q1], w[b * q1]);
  }
#pragma unroll
  for (int q1 = 0; q1 < 4; q1++) {
    Cx<T> y[4] = {t[0][q1], t[1][q1], t[2][q1], t[3][q1]};
    dft4<T, INV>(y);
#pragma unroll
    for (int q2 = 0; q2 < 4; q2++) v[q1 + 4 * q2] = y[q2];
  }
}
template <typename T, int R, bool INV> __device__ __forceinline__ void dftR(Cx<T> (&v)[R]) {
  if constexpr (R == 2) dft2<T, INV>(v);
  else if constexpr (R == 4) dft4<T, INV>(v);
  else if constexpr (R == 8) dft8<T, INV>(v);
  else dft16<T, INV>(v);
}

// one in-place pass over the tile s[nz][COLS] for blocks of length L = R << lsub (sub-stride 1 << lsub).
//   forward (DIF):  y_q = DFT_R(x)_q * w_L^{j q}          inverse (DIT): y_m = IDFT_R(x_q * conj(w_L^{j q}))_m
// All sizes are powers of two: index arithmetic is shifts and masks.
template <typename T, int R, bool INV, int COLS>
__device__ __forceinline__ void zpass(Cx<T> *s, const Cx<T> *tw, int nz, int lsub, int tid, int nt) {
  constexpr int LR = R == 16 ? 4 : (R == 8 ? 3 : (R == 4 ? 2 : 1)), LC = COLS == 16 ? 4 : (COLS == 8 ? 3 : 2);
  const int sub = 1 << lsub, nw = (nz >> LR) << LC;
  int ltstep = 0;  // log2(nz / L)
  while ((R << lsub << ltstep) < nz) ltstep++;
#pragma unroll 2
  for (int w = tid; w < nw; w += nt) {
    const int col = w & (COLS - 1), jb = w >> LC;
    const int b = jb >> lsub, j = jb & (sub - 1);
    Cx<T> *p = s + ((((b << LR) << lsub) + j) << LC) + col;
    const int st = sub << LC;
    Cx<T> v[R];
#pragma unroll
    for (int m = 0; m < R; m++) v[m] = p[m * st];
    const int jt = j << ltstep;
    if (INV) {
#pragma unroll
      for (int q = 1; q < R; q++) v[q] = cmul<T, true>(v[q], tw[jt * q]);
      dftR<T, R, true>(v);
    } else {
      dftR<T, R, false>(v);
#pragma unroll
      for (int q = 1; q < R; q++) v[q] = cmul<T, false>(v[q], tw[jt * q]);
    }
#pragma unroll
    for (int m = 0; m < R; m++) p[m * st] = v[m];
  }
}

template <typename T> __device__ __forceinline__ T rcp_(T x) { return T(1) / x; }
template <> __device__ __forceinline__ float rcp_<float>(float x) { return __frcp_rn(x); }

// The last forward pass (sub-stride 1: its twiddles are all 1), the symbol and the first inverse pass touch the
// same R consecutive positions of a column: one trip through registers.  Position b*R+q holds the mode
// kz = digitrev(b*R+q), whose az is tabulated in azr; the scaling is v * 1/(((ax+ay)+az) * (-N)).
// `axy`, `zero_col` belong to the thread's column, which does not change over the loop (nt % COLS == 0).
template <typename T, int R, int COLS>
__device__ __forceinline__ void zpass_mid(Cx<T> *s, const T *azr, int nz, int tid, int nt, T axy, T negN, bool zero_col) {
  constexpr int LR = R == 16 ? 4 : (R == 8 ? 3 : (R == 4 ? 2 : 1)), LC = COLS == 16 ? 4 : (COLS == 8 ? 3 : 2);
  const int nw = (nz >> LR) << LC;
#pragma unroll 2
  for (int w = tid; w < nw; w += nt) {
    const int col = w & (COLS - 1), b = w >> LC;
    Cx<T> *p = s + ((b << LR) << LC) + col;
    Cx<T> v[R];
#pragma unroll
    for (int m = 0; m < R; m++) v[m] = p[m * COLS];
    dftR<T, R, false>(v);
    const T *a = azr + (b << LR);
#pragma unroll
    for (int q = 0; q < R; q++) {
      T r = rcp_<T>(((axy + a[q])) * negN);
      if (q == 0 && zero_col && b == 0) r = T(0);  // the constant mode
      v[q] = Cx<T>{v[q].x * r, v[q].y * r};
    }
    dftR<T, R, true>(v);
#pragma unroll
    for (int m = 0; m < R; m++) p[m * COLS] = v[m];
  }
}

struct ZRadix {
  int n;        // number of passes
  int r[12];    // log2 of the radices of the forward passes in order (4 [Float32 only], 3, 2 or 1)
};

// twiddles and az-at-digit-reversed-positions tables of a block (all threads; followed by a barrier at the caller)
template <typename T>
__device__ __forceinline__ void zsolve_tables(Cx<T> *tw, T *azr, const Cx<T> *__restrict__ twg, const T *__restrict__ az, int nz,
                                              int lnz, const ZRadix &rad, int tid, int nt) {
  for (int t = tid; t < nz; t += nt) {
    tw[t] = twg[t];
    // position t = sum_i q_i * (nz / (R_1..R_i))  <->  kz = sum_i q_i * (R_1..R_{i-1})
    int kz = 0, lspan = lnz, lw = 0;
    for (int p = 0; p < rad.n; p++) {
      lspan -= rad.r[p];
      kz += ((t >> lspan) & ((1 << rad.r[p]) - 1)) << lw;
      lw += rad.r[p];
    }
    azr[t] = az[kz];
  }
}

// forward passes, symbol, inverse passes on the resident tile s[nz][COLS]; ends with a barrier
template <typename T, int COLS>
__device__ __forceinline__ void zsolve_tile(Cx<T> *s, const Cx<T> *tw, const T *azr, int nz, int lnz, const ZRadix &rad, int tid,
                                            int nt, T axy, T negN, bool zero_col) {
  // forward: natural -> digit-reversed; all passes but the last
  int lL = lnz;  // log2 of the current block length
  for (int p = 0; p + 1 < rad.n; p++) {
    const int lr = rad.r[p];
    if (sizeof(T) == 4 && lr == 4) { if constexpr (sizeof(T) == 4) zpass<T, 16, false, COLS>(s, tw, nz, lL - 4, tid, nt); }
    else if (lr == 3) zpass<T, 8, false, COLS>(s, tw, nz, lL - 3, tid, nt);
    else if (lr == 2) zpass<T, 4, false, COLS>(s, tw, nz, lL - 2, tid, nt);
    else zpass<T, 2, false, COLS>(s, tw, nz, lL - 1, tid, nt);
    lL -= lr;
    __syncthreads();
  }
  {
    const int lr = rad.r[rad.n - 1];  // == lL
    if (sizeof(T) == 4 && lr == 4) { if constexpr (sizeof(T) == 4) zpass_mid<T, 16, COLS>(s, azr, nz, tid, nt, axy, negN, zero_col); }
    else if (lr == 3) zpass_mid<T, 8, COLS>(s, azr, nz, tid, nt, axy, negN, zero_col);
    else if (lr == 2) zpass_mid<T, 4, COLS>(s, azr, nz, tid, nt, axy, negN, zero_col);
    else zpass_mid<T, 2, COLS>(s, azr, nz, tid, nt, axy, negN, zero_col);
    __syncthreads();
  }
  // inverse: digit-reversed -> natural (adjoint passes in reverse order)
  for (int p = rad.n - 2; p >= 0; p--) {
    const int lr = rad.r[p];
    if (sizeof(T) == 4 && lr == 4) { if constexpr (sizeof(T) == 4) zpass<T, 16, true, COLS>(s, tw, nz, lL, tid, nt); }
    else if (lr == 3) zpass<T, 8, true, COLS>(s, tw, nz, lL, tid, nt);
    else if (lr == 2) zpass<T, 4, true, COLS>(s, tw, nz, lL, tid, nt);
    else zpass<T, 2, true, COLS>(s, tw, nz, lL, tid, nt);
    lL += lr;
    __syncthreads();
  }
}

// S: spectrum, layout [z][ncols] (column index fastest: c = ky*nxw + kx_local).  nz = 2^m.
// TAG only names the instantiation: insb200.cu (built with -fmad=false: TAG 0, the `exact` mode) and fast.cu (FMA
// contraction on: TAG 1, the default) both instantiate this template, and two kernels with one mangled name but
// different code would collide when the library is linked.
template <typename T, int COLS, int TAG = 0>
__global__ void __launch_bounds__(512) k_zsolve(const __grid_constant__ CUtensorMap tm, const Cx<T> *__restrict__ twg,
                                                const T *__restrict__ ax, const T *__restrict__ ay,
                                                const T *__restrict__ az, int nz, long long ncols, int nxw, int kx0,
                                                T negN, ZRadix rad) {
  // The tile [nz][COLS] is moved by TMA in boxes of (COLS complex) x (up to 256 rows): columns past ncols are
  // zero-filled on load and clipped on store, and no thread spends instructions on the copy.
  extern __shared__ unsigned char zs_raw[];
  unsigned char *zb = zs_raw + ((128u - (smem_u32(zs_raw) & 127u)) & 127u);
  Cx<T> *s = reinterpret_cast<Cx<T> *>(zb);       // [nz][COLS]
  Cx<T> *tw = s + (size_t)nz * COLS;              // [nz]
  T *azr = reinterpret_cast<T *>(tw + nz);        // [nz]  az at digit-reversed positions
  uint64_t *bar = reinterpret_cast<uint64_t *>(azr + nz);
  const int tid = threadIdx.x, nt = blockDim.x;
  const long long c0 = (long long)blockIdx.x * COLS;
  const int rows = nz < 256 ? nz : 256;
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
    mbar_expect_tx(bar, (uint32_t)((size_t)nz * COLS * sizeof(Cx<T>)));
    for (int z = 0; z < nz; z += rows) tma_load_2d(s + (size_t)z * COLS, &tm, bar, (int)(2 * c0), z);
  }
  int lnz = 0;
  while ((1 << lnz) < nz) lnz++;
  zsolve_tables<T>(tw, azr, twg, az, nz, lnz, rad, tid, nt);
  // this thread's column (columns past ncols are zero-filled by the load and clipped by the store)
  const long long c = min(c0 + (tid & (COLS - 1)), ncols - 1);
  const int ky = (int)(c / nxw), kx = (int)(c - (long long)ky * nxw) + kx0;
  const T axy = ax[kx] + ay[ky];
  const bool zero_col = kx == 0 && ky == 0;
  __syncthreads();
  mbar_wait(bar, 0);
  zsolve_tile<T, COLS>(s, tw, azr, nz, lnz, rad, tid, nt, axy, negN, zero_col);
  fence_proxy_async();  // the tile written by all threads becomes visible to the bulk store
  __syncthreads();
  if (tid == 0) {
    for (int z = 0; z < nz; z += rows) tma_store_2d(&tm, s + (size_t)z * COLS, (int)(2 * c0), z);
    tma_store_commit_wait();
  }
}

#ifdef INSB200_EXPERIMENTAL  // negative result kept for reference: slower than scatter kernels + k_zsolve
// -------------------------------------------------------------------------------------------------
// k_zsolve_peer: the same solve for a slab-decomposed grid, FUSED with both transposes of the distributed
// FFT.  Rank r owns the kx range [kx0, kx0+kxl) for all ky and all z; the z range [q*nzl, (q+1)*nzl) of a
// column lives in rank q's in-slab spectrum S_q[zl][ky][kx] (row pitch nxp).  The block's tile
// [nz][CY x CX columns] is gathered by TMA straight from the P peers' buffers over NVLink (one 3-D box per
// peer and 256 planes), solved in shared memory and scattered back the same way: no packed copy of the
// spectrum exists, and the NVLink transfers of one tile overlap the arithmetic of the others.
// maps.m[q]: tensor (2*kxl words, ny, nzl) over S_q starting at column kx0, box (2*CX, CY, min(nzl,256)).
// local_load = 1 ("push only"): the forward transpose was done by k_fft_scatter_fwd (peer STORES, which NVLink
// moves faster than the 64-byte-row peer LOADS of a tile gather); the tile is then read from the local
// transposed copy through `lmap` (tensor (2*kxl, ny, nz), same box) and only the way back is fused.
// -------------------------------------------------------------------------------------------------
struct PeerMaps {
  CUtensorMap m[8];
};

template <typename T, int COLS>
__global__ void __launch_bounds__(512) k_zsolve_peer(const __grid_constant__ PeerMaps maps, const __grid_constant__ CUtensorMap lmap,
                                                     int local_load, const Cx<T> *__restrict__ twg,
                                                     const T *__restrict__ ax, const T *__restrict__ ay,
                                                     const T *__restrict__ az, int nz, int nzl, int P, int lcx, int tiles_x,
                                                     int kxl, int kx0, int ny, T negN, ZRadix rad) {
  extern __shared__ unsigned char zs_raw[];
  unsigned char *zb = zs_raw + ((128u - (smem_u32(zs_raw) & 127u)) & 127u);
  Cx<T> *s = reinterpret_cast<Cx<T> *>(zb);       // [nz][COLS], column = dy*CX + dx
  Cx<T> *tw = s + (size_t)nz * COLS;              // [nz]
  T *azr = reinterpret_cast<T *>(tw + nz);        // [nz]
  uint64_t *bar = reinterpret_cast<uint64_t *>(azr + nz);
  const int tid = threadIdx.x, nt = blockDim.x;
  const int CX = 1 << lcx, CY = COLS >> lcx;
  const int ty_ = blockIdx.x / tiles_x, tx_ = blockIdx.x - ty_ * tiles_x;
  const int cx0 = tx_ * CX, ky0 = ty_ * CY;
  const int rows = nzl < 256 ? nzl : 256;
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
    mbar_expect_tx(bar, (uint32_t)((size_t)nz * COLS * sizeof(Cx<T>)));
    if (local_load) {  // the transposed copy [z global][ky][kx local] made by k_fft_scatter_fwd (box rows = `rows` as well)
      for (int z = 0; z < nz; z += rows) tma_load_3d(s + (size_t)z * COLS, &lmap, bar, 2 * cx0, ky0, z);
    } else {
      for (int q = 0; q < P; q++)
        for (int z = 0; z < nzl; z += rows) tma_load_3d(s + (size_t)(q * nzl + z) * COLS, &maps.m[q], bar, 2 * cx0, ky0, z);
    }
  }
  int lnz = 0;
  while ((1 << lnz) < nz) lnz++;
  zsolve_tables<T>(tw, azr, twg, az, nz, lnz, rad, tid, nt);
  const int col = tid & (COLS - 1), dy = col >> lcx, dx = col & (CX - 1);
  const int kx = kx0 + min(cx0 + dx, kxl - 1), ky = min(ky0 + dy, ny - 1);  // clipped columns: any valid table entry
  const T axy = ax[kx] + ay[ky];
  const bool zero_col = kx == 0 && ky == 0;
  __syncthreads();
  mbar_wait(bar, 0);
  zsolve_tile<T, COLS>(s, tw, azr, nz, lnz, rad, tid, nt, axy, negN, zero_col);
  fence_proxy_async();
  __syncthreads();
  if (tid == 0) {
    for (int q = 0; q < P; q++)
      for (int z = 0; z < nzl; z += rows) tma_store_3d(&maps.m[q], s + (size_t)(q * nzl + z) * COLS, 2 * cx0, ky0, z);
    tma_store_commit_wait();
  }
}

#endif  // INSB200_EXPERIMENTAL

}  // namespace insb200
