// comm.cuh -- device side of the multi-GPU exchange over peer memory (NVLink / NVSwitch, buffers mapped with CUDA IPC):
//   k_peer_barrier     all ranks' earlier peer stores are complete and visible: flag words in every rank's flag buffer
//                      (st.release.sys / ld.acquire.sys) instead of a 4-byte ncclAllReduce per exchange
//   k_fft_scatter2_*   the two transposes of the distributed FFT: pack + transfer in ONE kernel, one warp per spectrum
//                      row, 16-byte accesses, every rank staggered onto a different peer
//   k_halo_push        ghost planes of the decomposed axis: one kernel stores all planes of all components straight
//                      into the ring neighbours' ghost planes
// The reference is single-device: nothing here replaces reference code, it is what the slab decomposition adds.
#pragma once
#include "ctx.cuh"

namespace insb200 {

struct PeerPtrs {
  void *p[16];
};

// ---- barrier ------------------------------------------------------------------------------------------------------
// Thread q < nranks signals peer q (flag word [rank] of q's buffer := epoch) and waits for peer q's signal in its own
// buffer.  Everything every rank stored into peer memory BEFORE its launch (earlier kernels / copies of the same
// stream) is visible to everything that follows this kernel on the waiting rank's stream.  Epochs only grow (all
// ranks launch the same sequence of barriers), so a flag is never reset; the compare is wrap-safe.  The spin is
// bounded (about 4 s): a rank that died must not hang the others' GPUs -- they give up, set flags[63] and the host
// reports it at the next synchronisation.
static __global__ void __launch_bounds__(32) k_peer_barrier(PeerPtrs F, int rank, int nranks, unsigned epoch, int slot0) {
  const int q = threadIdx.x;
  if (q >= nranks) return;
  __threadfence_system();
  unsigned *mine = reinterpret_cast<unsigned *>(F.p[rank]) + slot0;
  unsigned *theirs = reinterpret_cast<unsigned *>(F.p[q]) + slot0;
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs + rank), "r"(epoch) : "memory");
  const long long t0 = clock64();
  unsigned v;
  while (true) {
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine + q) : "memory");
    if ((int)(v - epoch) >= 0) break;
    if (clock64() - t0 > 8000000000LL) {
      reinterpret_cast<unsigned *>(F.p[rank])[63] = 0xdeadu;
      break;
    }
    __nanosleep(64);
  }
  __threadfence_system();
}

// ---- distributed-FFT transposes -----------------------------------------------------------------------------------
// In-slab spectrum of rank r: S[row][kx], row = zl * ny + ky (rows in all), row pitch ps, kx in [0, nxh).
// Peer q owns kx in [q kxc, q kxc + w_q), w_q = kxc except for the LAST rank, which also takes the remainder
// (n = 1024 on 8 ranks: 64 columns each, the last rank 65 with the Nyquist column): rows of 64 complex words are 512 B,
// aligned, and one warp moves a row with one 16-byte access per lane.
// Transposed spectrum of rank q: W_q[(src * rows + row)][k], k < w_q, dense -- all z of its kx range.
//   forward : S[row][q kxc + k]  ->  W_q[rank * rows + row][k]
//   backward: W[(q * rows + row)][k] (k < w_rank)  ->  S_q[row][kx0 + k]
// Launch: a PERSISTENT grid of `gridDim.x` blocks of kSc2Warps warps (set_param("scatter_blocks")).  Work item i (one
// per warp and trip) = row group i / P for destination slot i % P, destination q = (slot + rank + 1) mod P: consecutive
// warps serve different peers, every rank is staggered onto a different one (the schedule of a ring all-to-all; all ranks
// hitting one peer's ingress at once cost a factor of two at 8 GPUs, profiles/r1_*), and the local part of the transpose
// (q = rank) hides under the transfers.  Measured at 2 GPUs (profiles/r2_sweep_2gpu.txt): below 64 blocks the kernel
// starves NVLink, above 128 it takes SMs away from the cuFFT it overlaps with.  The default transposes are pitched
// peer copies on the copy engines (insb200.cu, transpose_ce); these kernels are the SM alternative.
constexpr int kSc2Warps = 8, kSc2Rows = 8;

template <typename C>
struct alignas(16) CVec {
  C v[16 / sizeof(C) > 0 ? 16 / sizeof(C) : 1];
};

template <typename C>
__device__ __forceinline__ void copy_rows(C *__restrict__ dst, i64 dpitch, const C *__restrict__ src, i64 spitch, int w, int nr, int lane) {
  constexpr int V = 16 / (int)sizeof(C) > 0 ? 16 / (int)sizeof(C) : 1;  // complex words per 16 bytes (2 in Float32, 1 in Float64)
  const bool vec = V > 1 && ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src) | (uintptr_t)(dpitch * sizeof(C)) |
                              (uintptr_t)(spitch * sizeof(C))) & 15) == 0;
  if (vec) {
    typedef CVec<C> Vt;
    const int wv = w / V;
    for (int k = lane; k < wv; k += 32) {
      Vt v[kSc2Rows];
#pragma unroll
      for (int j = 0; j < kSc2Rows; j++)
        if (j < nr) v[j] = reinterpret_cast<const Vt *>(src + (i64)j * spitch)[k];
#pragma unroll
      for (int j = 0; j < kSc2Rows; j++)
        if (j < nr) reinterpret_cast<Vt *>(dst + (i64)j * dpitch)[k] = v[j];
    }
    for (int k = wv * V + lane; k < w; k += 32)  // odd tail
      for (int j = 0; j < nr; j++) dst[(i64)j * dpitch + k] = src[(i64)j * spitch + k];
  } else {
    for (int k = lane; k < w; k += 32) {
      C v[kSc2Rows];
#pragma unroll
      for (int j = 0; j < kSc2Rows; j++)
        if (j < nr) v[j] = src[(i64)j * spitch + k];
#pragma unroll
      for (int j = 0; j < kSc2Rows; j++)
        if (j < nr) dst[(i64)j * dpitch + k] = v[j];
    }
  }
}

template <typename C>
__global__ void __launch_bounds__(32 * kSc2Warps)
    k_fft_scatter2_fwd(PeerPtrs W, const C *__restrict__ S, int nxh, int ps, i64 rows, int kxc, int rank, int nranks, i64 rbeg, i64 rend) {
  const i64 ngroups = (rend - rbeg + kSc2Rows - 1) / kSc2Rows, nitems = ngroups * nranks;
  const i64 nwarps = (i64)gridDim.x * kSc2Warps;
  for (i64 it = (i64)blockIdx.x * kSc2Warps + threadIdx.y; it < nitems; it += nwarps) {
    const i64 g = it / nranks;
    int q = (int)(it - g * nranks) + rank + 1;
    if (q >= nranks) q -= nranks;
    const int w = q == nranks - 1 ? nxh - q * kxc : kxc;
    if (w <= 0) continue;
    const i64 r0 = rbeg + g * kSc2Rows;
    const int nr = (int)(rend - r0 < kSc2Rows ? rend - r0 : kSc2Rows);
    copy_rows<C>(reinterpret_cast<C *>(W.p[q]) + ((i64)rank * rows + r0) * w, w, S + r0 * ps + (i64)q * kxc, ps, w, nr, threadIdx.x);
  }
}

template <typename C>
__global__ void __launch_bounds__(32 * kSc2Warps)
    k_fft_scatter2_bwd(PeerPtrs S, const C *__restrict__ W, int ps, i64 rows, int kxl, int kx0, int rank, int nranks, i64 rbeg, i64 rend) {
  if (kxl <= 0) return;
  const i64 ngroups = (rend - rbeg + kSc2Rows - 1) / kSc2Rows, nitems = ngroups * nranks;
  const i64 nwarps = (i64)gridDim.x * kSc2Warps;
  for (i64 it = (i64)blockIdx.x * kSc2Warps + threadIdx.y; it < nitems; it += nwarps) {
    const i64 g = it / nranks;
    int q = (int)(it - g * nranks) + rank + 1;
    if (q >= nranks) q -= nranks;
    const i64 r0 = rbeg + g * kSc2Rows;  // row of q's slab
    const int nr = (int)(rend - r0 < kSc2Rows ? rend - r0 : kSc2Rows);
    copy_rows<C>(reinterpret_cast<C *>(S.p[q]) + r0 * ps + kx0, ps, W + ((i64)q * rows + r0) * kxl, kxl, kxl, nr, threadIdx.x);
  }
}

// Local unpack of the packed transposes (transpose_mode 2): R[src][row][k] (k < w_src, blocks of rows * w_src words at
// offset rows * src * kxc -- what the peers' contiguous copies left in this rank's staging buffer) -> S[row][src kxc + k].
// The local PACK is k_fft_scatter2_fwd itself with all destination pointers aimed at the local staging buffer.
template <typename C>
__global__ void __launch_bounds__(32 * kSc2Warps)
    k_fft_unpack(C *__restrict__ S, const C *__restrict__ R, int nxh, int ps, i64 rows, int kxc, int nranks, i64 rbeg, i64 rend) {
  const i64 ngroups = (rend - rbeg + kSc2Rows - 1) / kSc2Rows, nitems = ngroups * nranks;
  const i64 nwarps = (i64)gridDim.x * kSc2Warps;
  for (i64 it = (i64)blockIdx.x * kSc2Warps + threadIdx.y; it < nitems; it += nwarps) {
    const i64 g = it / nranks;
    const int src = (int)(it - g * nranks);
    const int w = src == nranks - 1 ? nxh - src * kxc : kxc;
    if (w <= 0) continue;
    const i64 r0 = rbeg + g * kSc2Rows;
    const int nr = (int)(rend - r0 < kSc2Rows ? rend - r0 : kSc2Rows);
    copy_rows<C>(S + r0 * ps + (i64)src * kxc, ps, R + rows * (i64)src * kxc + r0 * w, w, w, nr, threadIdx.x);
  }
}

// ---- halo planes --------------------------------------------------------------------------------------------------
// Planes of a compact / padded buffer (nc components of stride sc, plane size pl elements, G ghost planes below the
// nl interior planes): the top wminus interior planes go into the NEXT rank's lower ghost planes, the bottom wplus
// interior planes into the PREVIOUS rank's upper ghost planes.  grid (blocks, nc, 2): z = 0 -> next, 1 -> prev.
template <typename T>
__global__ void __launch_bounds__(256) k_halo_push(T *__restrict__ next, T *__restrict__ prev, const T *__restrict__ base, i64 sc, i64 pl,
                                                   i64 nl, int G, int wminus, int wplus) {
  const bool to_next = blockIdx.z == 0;
  const int w = to_next ? wminus : wplus;
  if (w <= 0) return;
  const i64 n = (i64)w * pl;
  const T *src = base + (i64)blockIdx.y * sc + (to_next ? (G + nl - wminus) * pl : (i64)G * pl);
  T *dst = (to_next ? next + (i64)(G - wminus) * pl : prev + (G + nl) * pl) + (i64)blockIdx.y * sc;
  constexpr int V = 16 / (int)sizeof(T);
  const i64 stride = (i64)gridDim.x * blockDim.x;
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
    struct alignas(16) Vt { T v[V]; };
    const i64 nv = n / V;
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) reinterpret_cast<Vt *>(dst)[i] = reinterpret_cast<const Vt *>(src)[i];
    for (i64 i = nv * V + (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = src[i];
  } else {
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = src[i];
  }
}

}  // namespace insb200
