// insb200.cu -- C ABI (include/insb200.h) and host orchestration of the B200 time-stepping core.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false ... (see build.py)
//
// No CPU fallback exists anywhere in this file: every operator is a CUDA kernel launch (own
// kernels + cuFFT for the transforms + NCCL for halos / the FFT transpose).
#include <algorithm>
#include <limits>
#include <type_traits>

#include "ctx.cuh"
#include "stencil.cuh"
#include "tiled.cuh"
#include "tma_stage.cuh"
#ifdef INSB200_EXPERIMENTAL  // negative results kept for reference: packed-FP32 stage kernel, fused peer z-solve
#include "tma_stage_x2.cuh"
#endif
#include "stage2d.cuh"
#include "zsolve.cuh"
#include "zsolve_launch.cuh"
#include "comm.cuh"
#include "fast.cuh"

namespace insb200 {
thread_local std::string g_create_error;
static void clear_graphs(insb200_ctx *c);  // CUDA graphs of whole RK steps (below, next to the stepping loops)

static const int kG = 2;  // ghost planes on the decomposed axis of compact buffers (nranks > 1)

// ---- small host helpers ------------------------------------------------------------------------
static inline dim3 cell_block() { return dim3(64, 4, 1); }
static inline dim3 cell_grid(const int n[3]) {
  return dim3((n[0] + 63) / 64, (n[1] + 3) / 4, n[2]);
}

static Geo geo_ghosted(const insb200_ctx *c) {
  Geo g;
  for (int d = 0; d < 3; d++) { g.n[d] = c->nloc[d]; g.wrap[d] = 0; }
  return g;
}
static Geo geo_compact(const insb200_ctx *c) {
  Geo g;
  for (int d = 0; d < 3; d++) { g.n[d] = c->nloc[d]; g.wrap[d] = 1; }
  if (c->nranks > 1) g.wrap[c->last] = 0;
  return g;
}

// reference-layout view of a caller array
template <typename T>
static View<T> view_ghosted(const insb200_ctx *c, void *base) {
  View<T> v;
  const i64 N0 = c->nloc[0] + 2, N1 = c->nloc[1] + 2, N2 = c->D == 3 ? c->nloc[2] + 2 : 1;
  v.sy = N0;
  v.sz = c->D == 3 ? N0 * N1 : 0;
  v.sc = N0 * N1 * N2;
  v.p = (T *)base + 1 + v.sy + v.sz;
  return v;
}
template <typename T>
static View<const T> view_ghosted_c(const insb200_ctx *c, const void *base) {
  View<T> v = view_ghosted<T>(c, const_cast<void *>(base));
  return View<const T>{v.p, v.sy, v.sz, v.sc};
}
// compact view of an internal buffer (vector, scalar or tensor: sc is the same for all)
template <typename T>
static View<T> view_compact(const insb200_ctx *c, void *base) {
  View<T> v;
  v.sy = c->csy;
  v.sz = c->csz;
  v.sc = c->ccomp;
  v.p = (T *)base + c->corg;
  return v;
}
template <typename T>
static View<const T> view_compact_c(const insb200_ctx *c, const void *base) {
  View<T> v = view_compact<T>(c, const_cast<void *>(base));
  return View<const T>{v.p, v.sy, v.sz, v.sc};
}

template <typename T>
static Rh<T> make_rh(const insb200_ctx *c) {
  Rh<T> r;
  for (int d = 0; d < 3; d++) r.v[d] = T(1) / T(c->h[d]);
  return r;
}
template <typename T>
static T volume(const insb200_ctx *c) {  // (dx*dy)*dz in T  (operators.jl:92)
  T v = T(c->h[0]) * T(c->h[1]);
  if (c->D == 3) v = v * T(c->h[2]);
  return v;
}

static int dev_alloc(insb200_ctx *c, void **p, size_t bytes) {
  if (*p) return 0;
  bytes += 256;  // slack: TMA boxes / neighbour reads may touch a few elements past the last row
  INS_CUDA(c, cudaMalloc(p, bytes));
  INS_CUDA(c, cudaMemsetAsync(*p, 0, bytes, c->stream));
  c->ws_bytes += (i64)bytes;
  return 0;
}

static double sinpi01(double x) {  // x in [0,1]; reduced so that multiples of 1/2 are exact
  double r = x > 0.5 ? 1.0 - x : x;
  return std::sin(M_PI * r);
}

#define DISPATCH(c, FN, ...)                                                           \
  ((c)->dtype == INSB200_F32 ? ((c)->D == 2 ? FN<float, 2>(__VA_ARGS__) : FN<float, 3>(__VA_ARGS__)) \
                             : ((c)->D == 2 ? FN<double, 2>(__VA_ARGS__) : FN<double, 3>(__VA_ARGS__)))


#define CHECK_CTX(c) \
  if (!(c)) return 1; \
  if (cudaSetDevice((c)->device) != cudaSuccess) INS_FAIL((c), "cudaSetDevice failed")

// ---- barrier over peer memory -------------------------------------------------------------------------
// "every rank's peer stores issued so far on its stream are complete and visible to what follows on mine".  With the
// flag buffers mapped (setup_p2p) this is the 32-thread kernel k_peer_barrier (comm.cuh); otherwise -- and with
// set_param("fast_barrier", 0) -- a 4-byte ncclAllReduce.  All barriers of a context go to c->stream, in the same
// order on every rank, so one epoch counter serves them all.
static int peer_barrier(insb200_ctx *c) {
  if (c->fast_barrier && c->peerF[0]) {
    PeerPtrs F;
    for (int q = 0; q < 16; q++) F.p[q] = c->peerF[q];
    k_peer_barrier<<<1, 32, 0, c->stream>>>(F, c->rank, c->nranks, ++c->bar_epoch, 0);
    CHECK_LAUNCH(c);
    return 0;
  }
  INS_NCCL(c, ncclAllReduce(c->barbuf, c->barbuf, 1, ncclInt, ncclMax, c->comm, c->stream));
  return 0;
}

// ---- halo exchange along the decomposed axis (nranks > 1) ---------------------------------------
// compact buffer with c->G ghost planes each side; ncomp components of stride ccomp.
// wminus planes: my bottom ghosts <- prev's top interior planes; wplus: my top ghosts <- next's bottom.
template <typename T>
static int halo_compact(insb200_ctx *c, void *base, int ncomp, int wminus, int wplus) {
  if (c->nranks == 1 || (wminus <= 0 && wplus <= 0)) return 0;
  ScopedTimer tm(c, TM_HALO);
  const int prev = (c->rank + c->nranks - 1) % c->nranks, next = (c->rank + 1) % c->nranks;
  const ncclDataType_t dt = sizeof(T) == 4 ? ncclFloat : ncclDouble;
  const i64 pl = c->cplane, nl = c->nloc[c->last], G = c->G;
  const i64 Gp = c->padded ? c->gp : G;
  if (wminus > G || wplus > Gp || wminus > nl || wplus > nl) INS_FAIL(c, "halo width exceeds the ghost planes");
  auto hp = c->halo_peers.find(base);
  if (c->p2p && c->p2p_halo && hp != c->halo_peers.end()) {
    // registered buffer: ONE kernel stores all planes of all components straight into the ring neighbours' ghost
    // planes (buffers mapped with CUDA IPC), the flag barrier orders them before the consumers on every rank
    const i64 nmax = (i64)std::max(wminus, wplus) * pl;
    dim3 grd((unsigned)std::min<i64>((nmax / (16 / sizeof(T)) + 255) / 256, 4 * (i64)c->nsm), ncomp, 2);
    k_halo_push<T><<<grd, 256, 0, c->stream>>>((T *)hp->second.next, (T *)hp->second.prev, (const T *)base, c->ccomp, pl, nl, (int)G, wminus, wplus);
    CHECK_LAUNCH(c);
    return peer_barrier(c);
  }
  INS_NCCL(c, ncclGroupStart());
  for (int q = 0; q < ncomp; q++) {
    T *b = (T *)base + q * c->ccomp;
    if (wminus > 0) {  // top interior planes -> next ; recv bottom ghosts <- prev
      INS_NCCL(c, ncclSend(b + (G + nl - wminus) * pl, wminus * pl, dt, next, c->comm, c->stream));
      INS_NCCL(c, ncclRecv(b + (G - wminus) * pl, wminus * pl, dt, prev, c->comm, c->stream));
    }
    if (wplus > 0) {  // bottom interior planes -> prev ; recv top ghosts <- next
      INS_NCCL(c, ncclSend(b + G * pl, wplus * pl, dt, prev, c->comm, c->stream));
      INS_NCCL(c, ncclRecv(b + (G + nl) * pl, wplus * pl, dt, next, c->comm, c->stream));
    }
  }
  INS_NCCL(c, ncclGroupEnd());
  return 0;
}

}  // namespace insb200
using namespace insb200;

// =================================================================================================
// life cycle
// =================================================================================================
extern "C" int insb200_version(void) { return 100; }

extern "C" int insb200_comm_unique_id(void *id128) {
  static_assert(sizeof(ncclUniqueId) <= INSB200_NCCL_ID_BYTES, "id size");
  ncclUniqueId id;
  if (ncclGetUniqueId(&id) != ncclSuccess) { g_create_error = "ncclGetUniqueId failed"; return 1; }
  memset(id128, 0, INSB200_NCCL_ID_BYTES);
  memcpy(id128, &id, sizeof id);
  return 0;
}

template <typename T>
static int upload_symbol(insb200_ctx *c) {
  // ahat[a][k] = 4*Omega*sinpi(k/n_a)^2 / dx_a^2, evaluated in Float64 and stored as T
  // (src/pressure.jl:305-311: k/Np is Float64; Omega = prod(dx) and dx^2 are formed in T)
  T om = T(c->h[0]);
  for (int a = 1; a < c->D; a++) om = om * T(c->h[a]);
  for (int a = 0; a < c->D; a++) {
    const int n = c->nglob[a];
    const int kmax = a == 0 ? n / 2 + 1 : n;
    std::vector<T> v(kmax);
    const T dx = T(c->h[a]);
    const double num = (double)(T(4) * om), den = (double)(dx * dx);
    for (int k = 0; k < kmax; k++) {
      const double s = sinpi01((double)k / (double)n);
      v[k] = (T)(num * (s * s) / den);
    }
    INS_TRY(dev_alloc(c, &c->ahat[a], kmax * sizeof(T)));
    INS_CUDA(c, cudaMemcpyAsync(c->ahat[a], v.data(), kmax * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    INS_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return 0;
}

extern "C" int insb200_decomposition(int n_first, int n_last, int rank, int nranks, int *z0, int *nzl, int *kx0, int *kxl) {
  if (nranks < 1 || rank < 0 || rank >= nranks || n_first < 2 || n_last < 1 || n_last % nranks) return 1;
  const int nxh = n_first / 2 + 1, kxc = nxh / nranks;
  if (nzl) *nzl = n_last / nranks;
  if (z0) *z0 = rank * (n_last / nranks);
  if (kx0) *kx0 = rank * kxc;
  if (kxl) *kxl = rank == nranks - 1 ? nxh - rank * kxc : kxc;
  return 0;
}

extern "C" int insb200_create(insb200_ctx **out, int D, const int *n, int dtype, const double *h, double Re,
                              int device, int rank, int nranks, const void *nccl_id) {
  if (!out) { g_create_error = "ctx output pointer is NULL"; return 1; }
  *out = nullptr;
  if (D != 2 && D != 3) INS_FAIL(nullptr, "D must be 2 or 3");
  if (dtype != INSB200_F32 && dtype != INSB200_F64) INS_FAIL(nullptr, "dtype must be 0 (Float32) or 1 (Float64)");
  if (nranks < 1 || rank < 0 || rank >= nranks) INS_FAIL(nullptr, "bad rank / nranks");
  for (int a = 0; a < D; a++) {
    if (n[a] < 2 || (n[a] & 1)) INS_FAIL(nullptr, "spectral pressure solver requires an even number of volumes >= 2 per axis (utils.jl:10-12)");
    if (!(h[a] > 0)) INS_FAIL(nullptr, "grid spacing must be positive");
  }
  if (!(Re > 0)) INS_FAIL(nullptr, "Re must be positive");
  if (n[D - 1] % nranks) INS_FAIL(nullptr, "the last axis must be divisible by the number of ranks");
  if (nranks > 1 && n[D - 1] / nranks < 2) INS_FAIL(nullptr, "each rank needs at least 2 planes");
  if (nranks > 1 && !nccl_id) INS_FAIL(nullptr, "nccl_id required when nranks > 1");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    INS_FAIL(nullptr, "no CUDA device available: libinsb200 has no CPU fallback");
  if (device < 0 || device >= ndev) INS_FAIL(nullptr, "bad device ordinal");

  insb200_ctx *c = new insb200_ctx();
  c->D = D; c->dtype = dtype; c->Re = Re; c->device = device; c->rank = rank; c->nranks = nranks;
  c->last = D - 1;
  for (int a = 0; a < D; a++) { c->nglob[a] = n[a]; c->nloc[a] = n[a]; c->h[a] = h[a]; }
  if (insb200_decomposition(n[0], n[c->last], rank, nranks, &c->off_last, &c->nloc[c->last], &c->kx0, &c->kxl)) {
    delete c;
    INS_FAIL(nullptr, "bad decomposition");
  }
  if (D == 3) {
    // padded layout (ctx.cuh): 64 B of x margin each side, gm/gp ghost layers in y and z
    const int es = dtype == INSB200_F64 ? 8 : 4;
    c->padded = 1;
    c->gm = 2; c->gp = 3;
    c->org[0] = 64 / es; c->org[1] = c->gm; c->org[2] = c->gm;
    c->ext[0] = c->nloc[0] + 128 / es; c->ext[1] = c->nloc[1] + c->gm + c->gp; c->ext[2] = c->nloc[2] + c->gm + c->gp;
    c->G = c->gm;
    c->csy = c->ext[0];
    c->csz = (i64)c->ext[0] * c->ext[1];
    c->cplane = c->csz;
    c->cint = (i64)c->nloc[0] * c->nloc[1] * c->nloc[2];
    c->ccomp = c->csz * c->ext[2];
    c->corg = c->org[0] + c->org[1] * c->csy + c->org[2] * c->csz;
  } else {
    c->G = nranks > 1 ? kG : 0;
    c->csy = c->nloc[0];
    c->csz = 0;
    c->cplane = c->csy;
    c->cint = (i64)c->nloc[0] * c->nloc[1];
    c->ccomp = c->cint + 2 * (i64)c->G * c->cplane;
    c->corg = (i64)c->G * c->cplane;
  }
  c->nxh = n[0] / 2 + 1;
  // kx split of the transposed spectrum: nxh / nranks columns per rank, the LAST rank also takes the remainder
  // (n = 1024 on 8 ranks: 64 columns = 512-byte rows each, the last rank 65 with the Nyquist column)
  c->kxc = c->nxh / nranks;  // kx0, kxl: insb200_decomposition above
  c->nxp = c->nxh;

  auto bail = [&](int r) { std::string e = c->err; insb200_destroy(c); g_create_error = e; return r; };
  if (cudaSetDevice(device) != cudaSuccess) { c->err = "cudaSetDevice failed"; return bail(1); }
  if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess) { c->err = "stream creation failed"; return bail(1); }
  c->stream = c->own_stream;
  if (cudaDeviceGetAttribute(&c->nsm, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || c->nsm < 1) c->nsm = 148;
  if (nranks > 1) {
    ncclUniqueId id;
    memcpy(&id, nccl_id, sizeof id);
    ncclResult_t r = ncclCommInitRank(&c->comm, nranks, id, rank);
    if (r != ncclSuccess) { c->err = std::string("ncclCommInitRank: ") + ncclGetErrorString(r); return bail(1); }
  }
  int r = dtype == INSB200_F32 ? upload_symbol<float>(c) : upload_symbol<double>(c);
  if (r) return bail(r);
  // reduction scratch: partials for the largest launch + 4 outputs
  dim3 g = cell_grid(c->nloc);
  const i64 nb = (i64)g.x * g.y * g.z;
  if (dev_alloc(c, &c->red, (size_t)(4 * nb + 8) * sizeof(double))) return bail(1);
  if (cudaMallocHost((void **)&c->red_host, 8 * sizeof(double)) != cudaSuccess) { c->err = "cudaMallocHost failed"; return bail(1); }
  *out = c;
  return 0;
}

extern "C" int insb200_destroy(insb200_ctx *c) {
  if (!c) return 0;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  clear_graphs(c);
  if (c->plans) {
    if (c->plan_fwd) cufftDestroy(c->plan_fwd);
    if (c->plan_inv) cufftDestroy(c->plan_inv);
    if (c->plan_fwd_in) cufftDestroy(c->plan_fwd_in);
    if (c->plan_inv_in) cufftDestroy(c->plan_inv_in);
    if (c->plan_last) cufftDestroy(c->plan_last);
  }
  for (auto &st : c->ce_streams) if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
  for (auto &e : c->ce_events) if (e) cudaEventDestroy(e);
  if (c->p2p) {  // unmap the peers' buffers (every rank synchronised its stream above; destroy is collective)
    for (int q = 0; q < 16; q++) {
      if (q != c->rank && c->peerS[q]) cudaIpcCloseMemHandle(c->peerS[q]);
      if (q != c->rank && c->peerB[q]) cudaIpcCloseMemHandle(c->peerB[q]);
      if (q != c->rank && c->peerF[q]) cudaIpcCloseMemHandle(c->peerF[q]);
      if (q != c->rank && c->peerP[q]) cudaIpcCloseMemHandle(c->peerP[q]);
    }
  }
  for (auto &kv : c->halo_peers) {
    if (kv.second.next) cudaIpcCloseMemHandle(kv.second.next);
    if (kv.second.prev && kv.second.prev != kv.second.next) cudaIpcCloseMemHandle(kv.second.prev);
  }
  if (c->comm_stream) {
    cudaStreamSynchronize(c->comm_stream);
    for (auto &e : c->pipe_ev) if (e) cudaEventDestroy(e);
    for (auto &e : c->pack_ev) if (e) cudaEventDestroy(e);
    if (c->pipe_sync) cudaEventDestroy(c->pipe_sync);
    cudaStreamDestroy(c->comm_stream);
  }
  if (c->barbuf) cudaFree(c->barbuf);
  if (c->flagbuf) cudaFree(c->flagbuf);
  if (c->pstage) cudaFree(c->pstage);
  for (auto &p : c->ubuf) if (p) cudaFree(p);
  for (auto &p : c->kbuf) if (p) cudaFree(p);
  for (auto &p : c->ahat) if (p) cudaFree(p);
  void *ptrs[] = {c->force, c->force_g, c->sigma, c->sbuf, c->preal, c->preal2, c->pcplx, c->pcplx2, c->fwork, c->red, c->cufft_work, c->hostio, c->ztw, c->tws, c->zl_x};
  for (void *p : ptrs) if (p) cudaFree(p);
  if (c->red_host) cudaFreeHost(c->red_host);
  for (auto &r : c->recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  for (auto &e : c->pool) cudaEventDestroy(e);
  if (c->comm) ncclCommDestroy(c->comm);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  delete c;
  return 0;
}

extern "C" const char *insb200_last_error(const insb200_ctx *c) {
  return c ? c->err.c_str() : g_create_error.c_str();
}

extern "C" int insb200_set_stream(insb200_ctx *c, void *s) {
  CHECK_CTX(c);
  if (c->stream != (cudaStream_t)s) clear_graphs(c);
  c->stream = (cudaStream_t)s;  // NULL = the CUDA legacy default stream (not the ctx-owned one)
  if (c->plans) {
    cufftHandle hs[] = {c->plan_fwd, c->plan_inv, c->plan_fwd_in, c->plan_inv_in, c->plan_last};
    for (cufftHandle hd : hs) if (hd) INS_CUFFT(c, cufftSetStream(hd, c->stream));
  }
  return 0;
}

extern "C" int insb200_synchronize(insb200_ctx *c) {
  CHECK_CTX(c);
  INS_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int insb200_local_geometry(const insb200_ctx *c, int *nloc, int *off) {
  if (!c) return 1;
  for (int a = 0; a < c->D; a++) nloc[a] = c->nloc[a];
  if (off) *off = c->off_last;
  return 0;
}

// =================================================================================================
// lazily created workspaces and plans
// =================================================================================================
namespace insb200 {

static void close_p2p(insb200_ctx *c) {
  for (int q = 0; q < 16; q++) {
    if (q != c->rank) {
      if (c->peerS[q]) cudaIpcCloseMemHandle(c->peerS[q]);
      if (c->peerB[q]) cudaIpcCloseMemHandle(c->peerB[q]);
      if (c->peerF[q]) cudaIpcCloseMemHandle(c->peerF[q]);
      if (c->peerP[q]) cudaIpcCloseMemHandle(c->peerP[q]);
    }
    c->peerS[q] = c->peerB[q] = c->peerF[q] = c->peerP[q] = nullptr;
  }
  c->p2p = 0;
}

// Map every rank's two spectrum buffers into this process (CUDA IPC, peer access over NVLink) so the
// distributed-FFT transposes can store straight into the peers' memory.  Collective; on any failure on any
// rank all ranks keep the NCCL send/recv transpose (c->p2p stays 0).
static int setup_p2p(insb200_ctx *c) {
  if (!c->p2p_enable || c->nranks > 16) return 0;
  struct Rec {
    cudaIpcMemHandle_t h[4];
    int ok, has_stage, pad[14];
  };
  INS_TRY(dev_alloc(c, &c->barbuf, 256));
  INS_TRY(dev_alloc(c, &c->flagbuf, (size_t)2 << 20));  // zeroed: epochs start at 1.  2 MiB: an allocation of its own (IPC maps whole allocations)
  INS_CUDA(c, cudaStreamSynchronize(c->stream));
  std::vector<Rec> all(c->nranks);
  Rec mine;
  memset(&mine, 0, sizeof(mine));
  mine.ok = cudaIpcGetMemHandle(&mine.h[0], c->pcplx) == cudaSuccess && cudaIpcGetMemHandle(&mine.h[1], c->pcplx2) == cudaSuccess &&
            cudaIpcGetMemHandle(&mine.h[2], c->flagbuf) == cudaSuccess;
  mine.has_stage = c->pstage != nullptr;
  if (mine.ok && c->pstage) mine.ok = cudaIpcGetMemHandle(&mine.h[3], c->pstage) == cudaSuccess;
  if (!mine.ok) cudaGetLastError();
  void *dev = nullptr;
  INS_CUDA(c, cudaMalloc(&dev, sizeof(Rec) * c->nranks));
  INS_CUDA(c, cudaMemcpyAsync((char *)dev + sizeof(Rec) * c->rank, &mine, sizeof(Rec), cudaMemcpyHostToDevice, c->stream));
  INS_NCCL(c, ncclAllGather((char *)dev + sizeof(Rec) * c->rank, dev, sizeof(Rec), ncclChar, c->comm, c->stream));
  INS_CUDA(c, cudaMemcpyAsync(all.data(), dev, sizeof(Rec) * c->nranks, cudaMemcpyDeviceToHost, c->stream));
  INS_CUDA(c, cudaStreamSynchronize(c->stream));
  int ok = 1;
  for (int q = 0; q < c->nranks; q++) ok &= all[q].ok;
  if (ok) {
    for (int q = 0; q < c->nranks && ok; q++) {
      if (q == c->rank) { c->peerS[q] = c->pcplx; c->peerB[q] = c->pcplx2; c->peerF[q] = c->flagbuf; c->peerP[q] = c->pstage; continue; }
      if (c->pstage && all[q].has_stage && cudaIpcOpenMemHandle(&c->peerP[q], all[q].h[3], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { c->peerP[q] = nullptr; ok = 0; }
      if (cudaIpcOpenMemHandle(&c->peerS[q], all[q].h[0], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { c->peerS[q] = nullptr; ok = 0; }
      else if (cudaIpcOpenMemHandle(&c->peerB[q], all[q].h[1], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { c->peerB[q] = nullptr; ok = 0; }
      else if (cudaIpcOpenMemHandle(&c->peerF[q], all[q].h[2], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { c->peerF[q] = nullptr; ok = 0; }
    }
    if (!ok) cudaGetLastError();
  }
  // agree: min over ranks
  int *flag = (int *)c->barbuf;
  INS_CUDA(c, cudaMemcpyAsync(flag, &ok, sizeof(int), cudaMemcpyHostToDevice, c->stream));
  INS_NCCL(c, ncclAllReduce(flag, flag, 1, ncclInt, ncclMin, c->comm, c->stream));
  INS_CUDA(c, cudaMemcpyAsync(&ok, flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  INS_CUDA(c, cudaStreamSynchronize(c->stream));
  INS_CUDA(c, cudaFree(dev));
  if (!ok) {
    close_p2p(c);
    return 0;
  }
  c->p2p = 1;
  return 0;
}

// Map the ring neighbours' copies of a local buffer (same role, same size on every rank) so that halo planes can
// be copied straight into their ghost planes.  Collective (all ranks, same order); keeps NCCL halos on failure.
static int register_halo_buffer(insb200_ctx *c, void *base) {
  if (!c->p2p || !c->p2p_halo || !base || c->halo_peers.count(base)) return 0;
  struct Rec {
    cudaIpcMemHandle_t h;
    int ok, pad[15];
  };
  std::vector<Rec> all(c->nranks);
  Rec mine;
  memset(&mine, 0, sizeof(mine));
  mine.ok = cudaIpcGetMemHandle(&mine.h, base) == cudaSuccess;
  if (!mine.ok) cudaGetLastError();
  void *dev = nullptr;
  INS_CUDA(c, cudaMalloc(&dev, sizeof(Rec) * c->nranks));
  INS_CUDA(c, cudaMemcpyAsync((char *)dev + sizeof(Rec) * c->rank, &mine, sizeof(Rec), cudaMemcpyHostToDevice, c->stream));
  INS_NCCL(c, ncclAllGather((char *)dev + sizeof(Rec) * c->rank, dev, sizeof(Rec), ncclChar, c->comm, c->stream));
  INS_CUDA(c, cudaMemcpyAsync(all.data(), dev, sizeof(Rec) * c->nranks, cudaMemcpyDeviceToHost, c->stream));
  INS_CUDA(c, cudaStreamSynchronize(c->stream));
  int ok = 1;
  for (int q = 0; q < c->nranks; q++) ok &= all[q].ok;
  const int prev = (c->rank + c->nranks - 1) % c->nranks, next = (c->rank + 1) % c->nranks;
  insb200_ctx::NeighbourBuf nb;
  if (ok) {
    if (cudaIpcOpenMemHandle(&nb.next, all[next].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { nb.next = nullptr; ok = 0; }
    if (ok && prev == next) nb.prev = nb.next;
    else if (ok && cudaIpcOpenMemHandle(&nb.prev, all[prev].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { nb.prev = nullptr; ok = 0; }
    if (!ok) cudaGetLastError();
  }
  int *flag = (int *)c->barbuf;
  INS_CUDA(c, cudaMemcpyAsync(flag, &ok, sizeof(int), cudaMemcpyHostToDevice, c->stream));
  INS_NCCL(c, ncclAllReduce(flag, flag, 1, ncclInt, ncclMin, c->comm, c->stream));
  INS_CUDA(c, cudaMemcpyAsync(&ok, flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  INS_CUDA(c, cudaStreamSynchronize(c->stream));
  INS_CUDA(c, cudaFree(dev));
  if (!ok) {
    if (nb.next) cudaIpcCloseMemHandle(nb.next);
    if (nb.prev && nb.prev != nb.next) cudaIpcCloseMemHandle(nb.prev);
    c->p2p_halo = 0;  // (collectively) back to NCCL halos for every buffer
    return 0;
  }
  c->halo_peers[base] = nb;
  return 0;
}

static int ensure_fft(insb200_ctx *c) {
  if (c->plans) return 0;
  const size_t es = esize(c);
  const int D = c->D;
  const bool f32 = c->dtype == INSB200_F32;
  // real buffer: interior + ghost planes on the decomposed axis (the gradient reads p[I+e])
  INS_TRY(dev_alloc(c, &c->preal, (size_t)c->ccomp * es));
  if (c->nranks == 1) {
    const i64 nc = (i64)c->nxh * c->nloc[1] * c->nloc[2];
    INS_TRY(dev_alloc(c, &c->pcplx, (size_t)nc * 2 * es));
    if (D == 2) {
      INS_CUFFT(c, cufftPlan2d(&c->plan_fwd, c->nloc[1], c->nloc[0], f32 ? CUFFT_R2C : CUFFT_D2Z));
      INS_CUFFT(c, cufftPlan2d(&c->plan_inv, c->nloc[1], c->nloc[0], f32 ? CUFFT_C2R : CUFFT_Z2D));
    } else {
      // advanced data layout: the real side is the interior of the padded scalar buffer
      int nn[3] = {c->nloc[2], c->nloc[1], c->nloc[0]};
      int remb[3] = {c->ext[2], c->ext[1], c->ext[0]};
      int cemb[3] = {c->nloc[2], c->nloc[1], c->nxh};
      INS_CUFFT(c, cufftPlanMany(&c->plan_fwd, 3, nn, remb, 1, (int)c->ccomp, cemb, 1, cemb[0] * cemb[1] * cemb[2], f32 ? CUFFT_R2C : CUFFT_D2Z, 1));
      INS_CUFFT(c, cufftPlanMany(&c->plan_inv, 3, nn, cemb, 1, cemb[0] * cemb[1] * cemb[2], remb, 1, (int)c->ccomp, f32 ? CUFFT_C2R : CUFFT_Z2D, 1));
    }
    INS_CUFFT(c, cufftSetStream(c->plan_fwd, c->stream));
    INS_CUFFT(c, cufftSetStream(c->plan_inv, c->stream));
    if (D == 3) {  // batched in-plane transforms for the k_zsolve path
      int nn[2] = {c->nloc[1], c->nloc[0]};
      int remb[2] = {c->ext[1], c->ext[0]};
      int cemb[2] = {c->nloc[1], c->nxh};
      const int rdist = (int)c->csz, cdist = c->nloc[1] * c->nxh;
      c->fft_batch = (c->fft_chunk > 0 && c->nloc[2] % c->fft_chunk == 0) ? c->fft_chunk : c->nloc[2];
      c->fft_cdist = cdist;
      INS_CUFFT(c, cufftPlanMany(&c->plan_fwd_in, 2, nn, remb, 1, rdist, cemb, 1, cdist, f32 ? CUFFT_R2C : CUFFT_D2Z, c->fft_batch));
      INS_CUFFT(c, cufftPlanMany(&c->plan_inv_in, 2, nn, cemb, 1, cdist, remb, 1, rdist, f32 ? CUFFT_C2R : CUFFT_Z2D, c->fft_batch));
      INS_CUFFT(c, cufftSetStream(c->plan_fwd_in, c->stream));
      INS_CUFFT(c, cufftSetStream(c->plan_inv_in, c->stream));
    }
  } else {
    // rows of the in-slab spectrum: 3-D: ny*nzl, 2-D: nyl
    const i64 rows = D == 3 ? (i64)c->nloc[1] * c->nloc[2] : c->nloc[1];
    // row pitch of the in-slab spectrum: a multiple of 16 B so that the peers can address it by TMA
    // row pitch of the in-slab spectrum: a multiple of 128 B, so that the row pieces the transposes read and write
    // (64 complex words = 512 B at 1024^3 on 8 ranks) start on 128-byte lines
    if (D == 3) c->nxp = (c->nxh + (f32 ? 15 : 7)) & ~(f32 ? 15 : 7);
    const i64 kxmax = c->nxh - (i64)(c->nranks - 1) * c->kxc;                 // the last rank's (largest) column count
    const i64 slots = std::max((i64)c->nranks * kxmax, (i64)c->nxp) * rows;   // >= rows*nxp and >= nranks*rows*kxl
    INS_TRY(dev_alloc(c, &c->pcplx, (size_t)slots * 2 * es));
    // (pcplx2 is also the exchange buffer of the line solve, zline.cuh: room for two areas of up to 4 sub-chunks on tiny grids)
    const i64 slots2 = std::max(slots, 16 * (D == 3 ? (i64)c->nxp * c->nloc[1] : (i64)c->nxh) + 64);
    INS_TRY(dev_alloc(c, &c->pcplx2, (size_t)slots2 * 2 * es));
    c->cplx_slots = slots2;
    // form of the transposes: row pieces of at least 1 KB travel well as pitched peer copies
    // (2 GPUs at 512^3 per GPU: 17.6 against 18.9 ms/step); shorter ones (512 B at 4 and 8 GPUs) are packed first
    c->transpose_mode = c->transpose_mode_req >= 0 ? c->transpose_mode_req : ((size_t)c->kxc * 2 * es >= 1024 ? 1 : 2);
    if (c->transpose_mode < 2) c->transpose_ce = c->transpose_mode;
    if (D == 3 && c->p2p_enable && c->transpose_mode == 2) INS_TRY(dev_alloc(c, &c->pstage, (size_t)slots * 2 * es));
    if (D == 3) {
      int nn[2] = {c->nloc[1], c->nloc[0]};
      int remb[2] = {c->ext[1], c->ext[0]};
      int cemb[2] = {c->nloc[1], c->nxp};
      const int rdist = (int)c->csz, cdist = c->nloc[1] * c->nxp;
      c->fft_batch = (c->fft_chunk > 0 && c->nloc[2] % c->fft_chunk == 0) ? c->fft_chunk : c->nloc[2];
      // (measured: +2 % at 8 GPUs, -1 % at 2 GPUs, where the scatter competes with cuFFT for HBM: used from 4 ranks on)
      if (c->fft_chunk <= 0 && c->p2p_enable && c->nranks >= c->fft_pipe_min_ranks && c->fft_pipe > 1 && c->fft_pipe <= 16 && c->nloc[2] % c->fft_pipe == 0 && c->nloc[2] / c->fft_pipe >= 8)
        c->fft_batch = c->nloc[2] / c->fft_pipe;  // z-chunks of the transform / transpose pipeline
      c->fft_cdist = cdist;
      INS_CUFFT(c, cufftPlanMany(&c->plan_fwd_in, 2, nn, remb, 1, rdist, cemb, 1, cdist, f32 ? CUFFT_R2C : CUFFT_D2Z, c->fft_batch));
      INS_CUFFT(c, cufftPlanMany(&c->plan_inv_in, 2, nn, cemb, 1, cdist, remb, 1, rdist, f32 ? CUFFT_C2R : CUFFT_Z2D, c->fft_batch));
    } else {
      int nn[1] = {c->nloc[0]};
      INS_CUFFT(c, cufftPlanMany(&c->plan_fwd_in, 1, nn, nullptr, 1, 0, nullptr, 1, 0, f32 ? CUFFT_R2C : CUFFT_D2Z, c->nloc[1]));
      INS_CUFFT(c, cufftPlanMany(&c->plan_inv_in, 1, nn, nullptr, 1, 0, nullptr, 1, 0, f32 ? CUFFT_C2R : CUFFT_Z2D, c->nloc[1]));
    }
    INS_CUFFT(c, cufftSetStream(c->plan_fwd_in, c->stream));
    INS_CUFFT(c, cufftSetStream(c->plan_inv_in, c->stream));
    if (c->kxl > 0) {
      // transposed spectrum (kxl fastest, ky, z-global): transform along the last axis
      int nn[1] = {c->nglob[c->last]};
      const int stride = D == 3 ? c->kxl * c->nglob[1] : c->kxl;
      int emb[1] = {c->nglob[c->last]};
      INS_CUFFT(c, cufftPlanMany(&c->plan_last, 1, nn, emb, stride, 1, emb, stride, 1, f32 ? CUFFT_C2C : CUFFT_Z2Z, stride));
      INS_CUFFT(c, cufftSetStream(c->plan_last, c->stream));
    }
    INS_TRY(setup_p2p(c));
  }
  c->plans = true;
  return 0;
}

static int ensure_vec(insb200_ctx *c, void **p) { return dev_alloc(c, p, (size_t)c->ccomp * c->D * esize(c)); }
static int ensure_sigma(insb200_ctx *c) {
  return dev_alloc(c, &c->sigma, (size_t)c->ccomp * (c->D * (c->D + 1) / 2) * esize(c));
}

// ---- padded layout: ghost refresh and TMA descriptors ---------------------------------------------
// periodic ghosts of a padded buffer with nc components: x and y always, z by index wrap on one rank,
// by plane exchange with the ring neighbours otherwise (wm / wp planes on the minus / plus side)
template <typename T>
static int fill_ghosts(insb200_ctx *c, void *base, int nc, int wm, int wp) {
  if (!c->padded) return 0;
  {
    ScopedTimer tm(c, TM_GHOSTS);
    const int gm = c->gm, gp = c->gp, gw = gm + gp;
    const int n0 = c->nloc[0], n1 = c->nloc[1], n2 = c->nloc[2];
    const int fill_z = c->nranks == 1;
    const i64 cntA = fill_z ? (i64)gw * (n0 + gw) * (n1 + gw) : 0;
    const i64 cntB = (i64)n2 * gw * (n0 + gw), cntC = (i64)n2 * n1 * gw;
    const i64 tot = cntA + cntB + cntC;
    dim3 grd((unsigned)((tot + 255) / 256), nc, 1);
    k_fill_ghosts<T><<<grd, 256, 0, c->stream>>>((T *)base + c->corg, c->csy, c->csz, c->ccomp, n0, n1, n2, gm, gp, fill_z, cntA,
                                                cntB, cntC);
    CHECK_LAUNCH(c);
  }
  return halo_compact<T>(c, base, nc, wm, wp);
}



// descriptor of a padded buffer with nc components, box = (box_x, box_y, 1, nc)
int get_tmap(insb200_ctx *c, const void *base, int nc, int box_x, int box_y, CUtensorMap *out) {
  auto key = std::make_tuple(base, nc, box_y);
  auto it = c->tmaps.find(key);
  if (it != c->tmaps.end()) { *out = it->second; return 0; }
  static PFN_tmap_encode enc = nullptr;
  if (!enc) {
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    INS_CUDA(c, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
    if (!fn || qr != cudaDriverEntryPointSuccess) INS_FAIL(c, "cuTensorMapEncodeTiled is not available from the driver");
    enc = (PFN_tmap_encode)fn;
  }
  const cuuint64_t es = esize(c);
  cuuint64_t dims[4] = {(cuuint64_t)c->ext[0], (cuuint64_t)c->ext[1], (cuuint64_t)c->ext[2], (cuuint64_t)nc};
  cuuint64_t strides[3] = {(cuuint64_t)c->csy * es, (cuuint64_t)c->csz * es, (cuuint64_t)c->ccomp * es};
  cuuint32_t box[4] = {(cuuint32_t)box_x, (cuuint32_t)box_y, 1u, (cuuint32_t)nc};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUtensorMap tm;
  CUresult r = enc(&tm, c->dtype == INSB200_F64 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4,
                   const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) INS_FAIL(c, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
  c->tmaps[key] = tm;
  *out = tm;
  return 0;
}

// ---- layout conversion ----------------------------------------------------------------------------
// rows of the internal layout start on 16-byte (Float32) / 32-byte (Float64) boundaries when these hold
static bool copy4_ok(const insb200_ctx *c) {
  const size_t es = esize(c), a = 4 * es;
  return c->nloc[0] % 4 == 0 && (c->corg * es) % a == 0 && (c->csy * es) % a == 0 && (c->csz * es) % a == 0 && (c->ccomp * es) % a == 0;
}
static dim3 copy4_grid(const insb200_ctx *c) { return dim3((c->nloc[0] / 4 + 63) / 64, (c->nloc[1] + 3) / 4, c->nloc[2]); }

template <typename T, int D>
static int import_field(insb200_ctx *c, void *compact, const void *ghosted, int nc) {
  ScopedTimer tm(c, TM_IMPORT);
  Geo g = geo_ghosted(c);
  if (copy4_ok(c) && ((uintptr_t)compact % (4 * sizeof(T))) == 0) {
    k_copy_view4<T, true><<<copy4_grid(c), cell_block(), 0, c->stream>>>(view_compact<T>(c, compact), view_ghosted_c<T>(c, ghosted), g, nc);
    CHECK_LAUNCH(c);
    return 0;
  }
  k_copy_view<T><<<cell_grid(c->nloc), cell_block(), 0, c->stream>>>(view_compact<T>(c, compact), view_ghosted_c<T>(c, ghosted), g, nc);
  CHECK_LAUNCH(c);
  return 0;
}
template <typename T, int D>
static int export_field(insb200_ctx *c, void *ghosted, const void *compact, int nc) {
  ScopedTimer tm(c, TM_EXPORT);
  Geo g = geo_ghosted(c);
  if (copy4_ok(c) && ((uintptr_t)compact % (4 * sizeof(T))) == 0) {
    k_copy_view4<T, false><<<copy4_grid(c), cell_block(), 0, c->stream>>>(view_ghosted<T>(c, ghosted), view_compact_c<T>(c, compact), g, nc);
    CHECK_LAUNCH(c);
    return 0;
  }
  k_copy_view<T><<<cell_grid(c->nloc), cell_block(), 0, c->stream>>>(view_ghosted<T>(c, ghosted), view_compact_c<T>(c, compact), g, nc);
  CHECK_LAUNCH(c);
  return 0;
}

// periodic ghost fill of a reference-layout array with nc components (apply_bc_u! / apply_bc_p!)
template <typename T, int D>
static int bc_fill(insb200_ctx *c, void *base, int nc) {
  ScopedTimer tm(c, TM_BC);
  View<T> v = view_ghosted<T>(c, base);
  const int skip = c->nranks > 1 ? c->last : -1;
  for (int axis = 0; axis < D; axis++) {
    if (axis == skip) continue;
    const int o0 = axis == 0 ? 1 : 0;
    const int o1 = D == 3 ? (axis == 2 ? 1 : 2) : -1;
    dim3 blk(128, D == 3 ? 2 : 1, 1);
    dim3 grd((c->nloc[o0] + 2 + 127) / 128, D == 3 ? (c->nloc[o1] + 2 + 1) / 2 : 1, 2 * nc);
    k_bc_axis<T><<<grd, blk, 0, c->stream>>>(v.p, v.sy, v.sz, v.sc, c->nloc[0], c->nloc[1], c->nloc[2], D, nc, axis, skip);
    CHECK_LAUNCH(c);
  }
  if (c->nranks > 1) {
    // ghost planes of the decomposed axis: full planes (x/y ghosts included) from the ring neighbours
    const int prev = (c->rank + c->nranks - 1) % c->nranks, next = (c->rank + 1) % c->nranks;
    const ncclDataType_t dt = sizeof(T) == 4 ? ncclFloat : ncclDouble;
    const i64 pl = D == 3 ? (i64)(c->nloc[0] + 2) * (c->nloc[1] + 2) : (c->nloc[0] + 2);
    const i64 nl = c->nloc[c->last];
    INS_NCCL(c, ncclGroupStart());
    for (int q = 0; q < nc; q++) {
      T *b = (T *)base + q * v.sc;
      INS_NCCL(c, ncclSend(b + nl * pl, pl, dt, next, c->comm, c->stream));        // last interior plane
      INS_NCCL(c, ncclRecv(b, pl, dt, prev, c->comm, c->stream));                  // ghost plane 0
      INS_NCCL(c, ncclSend(b + pl, pl, dt, prev, c->comm, c->stream));             // first interior plane
      INS_NCCL(c, ncclRecv(b + (nl + 1) * pl, pl, dt, next, c->comm, c->stream));  // ghost plane N-1
    }
    INS_NCCL(c, ncclGroupEnd());
  }
  return 0;
}

// ---- fused last-axis transform + symbol + inverse (zsolve.cuh) --------------------------------------
// in-plane (x, y) transforms of the slab, `fft_batch` planes per cuFFT exec
template <typename T, typename C>
static int exec_fwd_in(insb200_ctx *c, T *pr, C *S) {
  ScopedTimer tm(c, TM_FFT_FWD);
  const int nb = (c->D == 3 && c->fft_batch > 0) ? c->nloc[2] / c->fft_batch : 1;
  for (int b = 0; b < nb; b++) {
    T *r = pr + (i64)b * c->fft_batch * c->csz;
    C *s = S + (i64)b * c->fft_batch * c->fft_cdist;
    if (sizeof(T) == 4) INS_CUFFT(c, cufftExecR2C(c->plan_fwd_in, (cufftReal *)r, (cufftComplex *)s));
    else INS_CUFFT(c, cufftExecD2Z(c->plan_fwd_in, (cufftDoubleReal *)r, (cufftDoubleComplex *)s));
  }
  c->fft_execs++;
  return 0;
}
template <typename T, typename C>
static int exec_inv_in(insb200_ctx *c, C *S, T *pr) {
  ScopedTimer tm(c, TM_FFT_INV);
  const int nb = (c->D == 3 && c->fft_batch > 0) ? c->nloc[2] / c->fft_batch : 1;
  for (int b = 0; b < nb; b++) {
    T *r = pr + (i64)b * c->fft_batch * c->csz;
    C *s = S + (i64)b * c->fft_batch * c->fft_cdist;
    if (sizeof(T) == 4) INS_CUFFT(c, cufftExecC2R(c->plan_inv_in, (cufftComplex *)s, (cufftReal *)r));
    else INS_CUFFT(c, cufftExecZ2D(c->plan_inv_in, (cufftDoubleComplex *)s, (cufftDoubleReal *)r));
  }
  c->fft_execs++;
  return 0;
}

static bool zsolve_supported(const insb200_ctx *c) {
  const int nz = c->nglob[2];
  return c->D == 3 && c->zsolve && nz >= 8 && nz <= 2048 && (nz & (nz - 1)) == 0;
}

// twiddle table exp(-2 pi i t / nz) (Float64 on the host, stored in T) and the radix plan of the last axis
template <typename T>
static int zsolve_prepare(insb200_ctx *c, ZRadix &rad) {
  const int nz = c->nglob[2];
  if (!c->ztw) {
    std::vector<Cx<T>> tw(nz);
    for (int t = 0; t < nz; t++) {
      const double a = -2.0 * M_PI * (double)t / (double)nz;
      tw[t] = Cx<T>{(T)std::cos(a), (T)std::sin(a)};
    }
    INS_TRY(dev_alloc(c, &c->ztw, nz * sizeof(Cx<T>)));
    INS_CUDA(c, cudaMemcpyAsync(c->ztw, tw.data(), nz * sizeof(Cx<T>), cudaMemcpyHostToDevice, c->stream));
    INS_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  rad.n = 0;
  int m = 0;
  while ((1 << m) < nz) m++;
  if (sizeof(T) == 4 && c->zradix16) {
    // fewest trips through shared memory with radices up to 16 (Float32: 16 complex values in registers):
    // ceil(m / 4) digits as even as possible, the larger ones last -- 1024 = 8 * 8 * 16 (5 passes instead of the 7 of
    // 8 * 8 * 8 * 2), 256 = 16 * 16, 512 = 8 * 8 * 8 as before
    const int nd = (m + 3) / 4, base = m / nd, rem = m % nd;
    for (int i = 0; i < nd; i++) rad.r[rad.n++] = base + (i >= nd - rem ? 1 : 0);
    return 0;
  }
  while (m >= 3) { rad.r[rad.n++] = 3; m -= 3; }  // log2 of the radix: 8, 8, .. then 4 or 2
  if (m > 0) rad.r[rad.n++] = m;
  return 0;
}

// S: [z][ncols] spectrum with all nz points of the last axis local; columns c = ky*nxw + (kx - kx0)
template <typename T>
static int launch_zsolve(insb200_ctx *c, void *S, i64 ncols, int nxw, int kx0, T invN) {
  ScopedTimer tm(c, TM_POISSON_SCALE);
  ZRadix rad;
  INS_TRY((zsolve_prepare<T>(c, rad)));
  // default: the instantiation of fast.cu (FMA contraction on -- an FFT butterfly has nothing to be bit-identical
  // with); set_param("exact", 1): the one of this translation unit (-fmad=false)
  if (!c->exact) return launch_zsolve_fast<T>(c, S, ncols, nxw, kx0, invN, rad);
  return launch_zsolve_cols<T, 0>(c, S, ncols, nxw, kx0, invN, rad);
}

#ifdef INSB200_EXPERIMENTAL
// fused transposes + last-axis solve over the peers' in-slab spectra (k_zsolve_peer)
template <typename T, int COLS>
static int launch_zsolve_peer_i(insb200_ctx *c, T invN, const ZRadix &rad, bool local_load) {
  const int nz = c->nglob[2], nzl = c->nloc[2], ny = c->nglob[1], P = c->nranks;
  // columns of a tile: CX consecutive kx times CY = COLS/CX consecutive ky; least padding of the kx range wins
  int CX = COLS;
  for (int cx = COLS; cx * (int)sizeof(Cx<T>) >= 32; cx >>= 1) {
    const int pad = (c->kxl + cx - 1) / cx * cx, best = (c->kxl + CX - 1) / CX * CX;
    if (pad < best) CX = cx;
  }
  const int CY = COLS / CX;
  int lcx = 0;
  while ((1 << lcx) < CX) lcx++;
  const int key = COLS * 64 + CX;
  auto it = c->peer_tmaps.find(key);
  if (it == c->peer_tmaps.end()) {
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    INS_CUDA(c, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
    if (!fn || qr != cudaDriverEntryPointSuccess) INS_FAIL(c, "cuTensorMapEncodeTiled is not available from the driver");
    std::vector<CUtensorMap> maps(P + 1);  // [P]: the local transposed copy pcplx2 = [nz][ny][kxl, pitch kxp]
    const int kxp = sizeof(T) == 4 ? (c->kxl + 1) & ~1 : c->kxl;
    for (int q = 0; q <= P; q++) {
      void *base = q < P ? (void *)((char *)c->peerS[q] + (size_t)c->kx0 * sizeof(Cx<T>)) : c->pcplx2;
      const int pitch = q < P ? c->nxp : kxp;
      cuuint64_t dims[3] = {(cuuint64_t)(2 * c->kxl), (cuuint64_t)ny, (cuuint64_t)(q < P ? nzl : nz)};
      cuuint64_t strides[2] = {(cuuint64_t)pitch * sizeof(Cx<T>), (cuuint64_t)pitch * ny * sizeof(Cx<T>)};
      cuuint32_t box[3] = {(cuuint32_t)(2 * CX), (cuuint32_t)CY, (cuuint32_t)std::min(nzl, 256)};
      cuuint32_t estr[3] = {1, 1, 1};
      CUresult r = ((PFN_tmap_encode)fn)(&maps[q], sizeof(T) == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3,
                                         base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) INS_FAIL(c, "cuTensorMapEncodeTiled (peer spectrum) failed with CUresult " + std::to_string((int)r));
    }
    it = c->peer_tmaps.emplace(key, std::move(maps)).first;
  }
  PeerMaps pm;
  memset(&pm, 0, sizeof(pm));
  for (int q = 0; q < P; q++) pm.m[q] = it->second[q];
  const size_t smem = 128 + (size_t)(nz * COLS + nz) * sizeof(Cx<T>) + (size_t)nz * sizeof(T) + 16;
  auto kern = k_zsolve_peer<T, COLS>;
  INS_DYN_SMEM(c, kern, smem);
  const int tiles_x = (c->kxl + CX - 1) / CX, tiles_y = (ny + CY - 1) / CY;
  kern<<<(unsigned)(tiles_x * tiles_y), c->zthreads, smem, c->stream>>>(pm, it->second[P], local_load ? 1 : 0, (const Cx<T> *)c->ztw, (const T *)c->ahat[0],
                                                                      (const T *)c->ahat[1], (const T *)c->ahat[2], nz, nzl, P, lcx,
                                                                      tiles_x, c->kxl, c->kx0, ny, T(-1) / invN, rad);
  CHECK_LAUNCH(c);
  return 0;
}

template <typename T>
static int launch_zsolve_peer(insb200_ctx *c, T invN, bool local_load) {
  ScopedTimer tm(c, TM_POISSON_SCALE);
  ZRadix rad;
  INS_TRY((zsolve_prepare<T>(c, rad)));
  const size_t col_bytes = (size_t)c->nglob[2] * sizeof(Cx<T>);
  if (c->zcols == 4) return launch_zsolve_peer_i<T, 4>(c, invN, rad, local_load);
  if (c->zcols == 8 && 8 * col_bytes <= 131072) return launch_zsolve_peer_i<T, 8>(c, invN, rad, local_load);
  if (16 * col_bytes <= 65536) return launch_zsolve_peer_i<T, 16>(c, invN, rad, local_load);
  if (8 * col_bytes <= 65536) return launch_zsolve_peer_i<T, 8>(c, invN, rad, local_load);
  return launch_zsolve_peer_i<T, 4>(c, invN, rad, local_load);
}

#endif  // INSB200_EXPERIMENTAL

// ---- the last axis without transposes (zline.cuh): in-slab transforms -> chunk sums -> barrier -> two recursions --
static bool zline_active(const insb200_ctx *c) {
  if (c->zline_req == 0) return false;
  // k_zline_zero keeps the constant mode's column in shared memory (16 B per point): automatic up to the 48 KB every
  // launch may use (measured: lines of up to 1024 points); up to 128 KB on request ("zline" = 1)
  if (c->nglob[c->last] > (c->zline_req == 1 ? 8192 : 3072)) return false;
  if (c->nranks == 1) return c->zline_req == 1 && c->D == 3;  // one rank: k_zsolve is the default (one pass instead of five)
  return c->p2p != 0;
}

template <typename T, int D, typename C>
static int poisson_line(insb200_ctx *c, T *pr, C *S) {
  ZlGeo g;
  g.nkx = c->nxh;
  g.pitch = (D == 3 && c->nranks > 1) ? c->nxp : c->nxh;
  g.nky = D == 3 ? c->nloc[1] : 1;
  g.pstride = (i64)g.pitch * g.nky;
  g.mloc = c->nloc[c->last];
  g.P = c->nranks;
  g.rank = c->rank;
  g.n = c->nglob[c->last];
  int ch = 1;
  if (c->zline_chunks > 0) {
    if (g.mloc % c->zline_chunks) INS_FAIL(c, "zline_chunks must divide the local extent of the last axis");
    ch = c->zline_chunks;
  } else {
    // enough threads (one per in-plane mode and sub-chunk) to fill the machine, and no more: every sub-chunk adds sums to
    // exchange and a term to the chains (8 GPUs, 1024^3, 540 k modes per plane: 16.06 / 16.29 / 16.82 ms/step with 1 / 2 / 4
    // sub-chunks, profiles/r2_sweep_19_8gpu_zline.txt)
    while (g.pstride * ch < 300000 && g.mloc % (2 * ch) == 0 && g.mloc / (2 * ch) >= 16) ch *= 2;
  }
  g.C = ch;
  g.mv = g.mloc / ch;
  // dz^2 / W in the arithmetic of upload_symbol (W and dz^2 formed in T)
  T om = T(c->h[0]);
  for (int a = 1; a < D; a++) om = om * T(c->h[a]);
  const T dl = T(c->h[c->last]);
  g.kappa = (double)(dl * dl) / (double)om;
  g.scale = g.kappa / ((double)c->nglob[0] * (D == 3 ? (double)c->nglob[1] : 1.0));
  const i64 area = zl_area_words(g);
  PeerPtrs X;
  for (int q = 0; q < 16; q++) X.p[q] = nullptr;
  if (c->nranks == 1) {
    INS_TRY(dev_alloc(c, &c->zl_x, (size_t)(2 * area) * sizeof(C)));
    X.p[0] = c->zl_x;
  } else {
    if (2 * area > c->cplx_slots) INS_FAIL(c, "line solve: the exchange areas do not fit the peer-mapped buffer (fewer zline_chunks)");
    for (int q = 0; q < c->nranks; q++) {
      if (!c->peerB[q]) INS_FAIL(c, "line solve: a peer's exchange buffer is not mapped");
      X.p[q] = c->peerB[q];
    }
  }
  g.xoff = (i64)(c->zl_count++ & 1u) * area;  // two areas alternate: a peer may still read the previous solve's sums
  INS_TRY((exec_fwd_in<T, C>(c, pr, S)));
  {
    ScopedTimer tm(c, TM_ZSUMS);
    INS_TRY((launch_zline_forward<T>(c, g, X, S)));
  }
  if (c->nranks > 1) {
    ScopedTimer tm(c, TM_A2A);
    INS_TRY(peer_barrier(c));
  }
  {
    ScopedTimer tm(c, TM_POISSON_SCALE);
    INS_TRY((launch_zline_backward<T>(c, g, X, S)));
  }
  INS_TRY((exec_inv_in<T, C>(c, S, pr)));
  return 0;
}

// ---- spectral Poisson solve on the compact real buffer (in: preal interior = rhs; out: solution) --
template <typename T, int D>
static int poisson_solve(insb200_ctx *c) {
  INS_TRY(ensure_fft(c));
  typedef typename std::conditional<sizeof(T) == 4, cufftComplex, cufftDoubleComplex>::type C;
  T *pr = (T *)c->preal + c->corg;
  C *S = (C *)c->pcplx;
  const double ntot = (double)c->nglob[0] * c->nglob[1] * (D == 3 ? c->nglob[2] : 1);
  const T invN = T(1) / T(ntot);
  if (zline_active(c)) return poisson_line<T, D, C>(c, pr, S);
  if (c->nranks == 1 && D == 3 && zsolve_supported(c)) {
    INS_TRY((exec_fwd_in<T, C>(c, pr, S)));
    INS_TRY((launch_zsolve<T>(c, S, (i64)c->nloc[1] * c->nxh, c->nxh, 0, invN)));
    INS_TRY((exec_inv_in<T, C>(c, S, pr)));
    return 0;
  }
  if (c->nranks == 1) {
    {
      ScopedTimer tm(c, TM_FFT_FWD);
      if (sizeof(T) == 4) INS_CUFFT(c, cufftExecR2C(c->plan_fwd, (cufftReal *)pr, (cufftComplex *)S));
      else INS_CUFFT(c, cufftExecD2Z(c->plan_fwd, (cufftDoubleReal *)pr, (cufftDoubleComplex *)S));
      c->fft_execs++;
    }
    {
      ScopedTimer tm(c, TM_POISSON_SCALE);
      dim3 blk(64, 4, 1), grd((c->nxh + 63) / 64, (c->nloc[1] + 3) / 4, c->nloc[2]);
      k_poisson_scale<T, C, D><<<grd, blk, 0, c->stream>>>(S, (const T *)c->ahat[0], (const T *)c->ahat[1], (const T *)c->ahat[2],
                                                         c->nxh, c->nloc[1], c->nloc[2], 0, invN);
      CHECK_LAUNCH(c);
    }
    {
      ScopedTimer tm(c, TM_FFT_INV);
      if (sizeof(T) == 4) INS_CUFFT(c, cufftExecC2R(c->plan_inv, (cufftComplex *)S, (cufftReal *)pr));
      else INS_CUFFT(c, cufftExecZ2D(c->plan_inv, (cufftDoubleComplex *)S, (cufftDoubleReal *)pr));
      c->fft_execs++;
    }
    return 0;
  }
  // ---- slab-decomposed transform: in-slab FFT -> transpose -> last-axis solve -> transpose back -> in-slab inverse FFT
  C *B = (C *)c->pcplx2;
  const int P = c->nranks;
  const i64 rows = D == 3 ? (i64)c->nloc[1] * c->nloc[2] : c->nloc[1];
  const ncclDataType_t dt = sizeof(T) == 4 ? ncclFloat : ncclDouble;
  auto chunk_w = [&](int q) { return q == P - 1 ? c->nxh - q * c->kxc : c->kxc; };
  dim3 pblk(64, 4, 1), pgrd((unsigned)((rows + 3) / 4), (c->nxh + 63) / 64, 1);
  dim3 sblk(32, kSc2Warps, 1);
  auto sgrid = [&](i64 nr) {  // persistent scatter grid: at most `scatter_blocks` blocks (comm.cuh)
    const i64 need = ((nr + kSc2Rows - 1) / kSc2Rows * P + kSc2Warps - 1) / kSc2Warps;
    return dim3((unsigned)std::max<i64>(1, std::min<i64>(need, c->scatter_blocks)), 1, 1);
  };
  auto lgrid = [&](i64 nr) {  // the same kernels moving LOCAL memory (pack / unpack of the packed transposes): fill the machine
    const i64 need = ((nr + kSc2Rows - 1) / kSc2Rows * P + kSc2Warps - 1) / kSc2Warps;
    return dim3((unsigned)std::max<i64>(1, std::min<i64>(need, 8 * (i64)c->nsm)), 1, 1);
  };
  const bool pipelined = c->p2p && D == 3 && zsolve_supported(c) && c->fft_batch > 0 && c->nloc[2] % c->fft_batch == 0;
  C *W = c->p2p ? B : S;  // the transposed spectrum (kxl fastest, ky, global z) of this rank
  PeerPtrs pS, pB;
  for (int q = 0; q < 16; q++) { pS.p[q] = c->peerS[q]; pB.p[q] = c->peerB[q]; }
  if (pipelined) {
    // ---- pipelined: the slab is transformed in NB z-chunks (NB = 1: the whole slab); the transpose of chunk b runs
    // while cuFFT works on chunk b+1 (forward), and the C2R of chunk b while chunk b+1 travels back.
    // transpose_ce = 1 (default): the transposes are PITCHED PEER COPIES on the copy engines, one stream per destination
    // (cudaMemcpy2DAsync into the peers' IPC-mapped buffers): no SM is taken away from cuFFT, and a pitched copy moves
    // 663 GB/s where SM stores of the same rows reach 555 (tools/p2p_probe.cu, profiles/r1_p2p_probe_2xB200.txt).
    // transpose_ce = 0: the persistent scatter kernels of comm.cuh on one second stream.
    const int NB = c->nloc[2] / c->fft_batch;
    const i64 rb = (i64)c->fft_batch * c->nloc[1];  // rows per chunk
    if (NB > 16) INS_FAIL(c, "more than 16 transform chunks");
    if (!c->comm_stream) {
      INS_CUDA(c, cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking));
      for (int b = 0; b < 16; b++) INS_CUDA(c, cudaEventCreateWithFlags(&c->pipe_ev[b], cudaEventDisableTiming));
      for (int b = 0; b < 16; b++) INS_CUDA(c, cudaEventCreateWithFlags(&c->pack_ev[b], cudaEventDisableTiming));
      INS_CUDA(c, cudaEventCreateWithFlags(&c->pipe_sync, cudaEventDisableTiming));
    }
    // transpose_mode 2 (default): local pack (SM kernel at HBM speed) -> one CONTIGUOUS peer copy per destination on the
    // copy engines -> local unpack on the receiving side after the barrier.  8 GPUs, 1024^3: the direct forms move
    // 512-byte row pieces and reach 290 (pitched copies) / 360 GB/s (scatter kernels) per GPU; contiguous peer copies
    // run at the copy engines' full rate (profiles/r2_sweep_*8gpu*).
    const bool packed = c->transpose_mode == 2 && c->pstage != nullptr;
    for (int q = 0; packed && q < P; q++) if (!c->peerP[q]) INS_FAIL(c, "packed transposes: a peer's staging buffer is not mapped");
    const bool ce = packed || c->transpose_ce != 0;
    C *PS = (C *)c->pstage;
    PeerPtrs pPack;  // local pack = k_fft_scatter2_fwd with every destination aimed at this rank's staging buffer
    for (int q = 0; q < 16; q++) pPack.p[q] = nullptr;
    for (int q = 0; packed && q < P; q++) pPack.p[q] = (void *)(PS + rows * (i64)q * c->kxc - (i64)c->rank * rows * chunk_w(q));
    if (ce && c->ce_streams.empty()) {
      c->ce_streams.resize(P);
      for (int d = 0; d < P; d++) INS_CUDA(c, cudaStreamCreateWithFlags(&c->ce_streams[d], cudaStreamNonBlocking));
      c->ce_events.resize((size_t)16 * P);
      for (auto &e : c->ce_events) INS_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    // transpose of chunk b after event `after`; fwd: S -> peers' B, bwd: B -> peers' S.  Afterwards comm_stream has
    // waited for every piece of the chunk (so an event recorded there means "chunk b has left this rank").
    auto transpose_chunk = [&](int b, bool fwd, cudaEvent_t after) -> int {
      INS_CUDA(c, cudaStreamWaitEvent(c->comm_stream, after, 0));
      ScopedTimer ts(c, TM_SCATTER, c->comm_stream);
      const i64 r0 = b * rb;
      if (packed) {
        cudaEvent_t start = after;
        if (fwd) {  // pack chunk b: S (strided rows) -> staging, one contiguous block per destination
          k_fft_scatter2_fwd<C><<<lgrid(rb), sblk, 0, c->comm_stream>>>(pPack, S, c->nxh, c->nxp, rows, c->kxc, c->rank, P, r0, r0 + rb);
          CHECK_LAUNCH(c);
          start = c->pack_ev[b];
          INS_CUDA(c, cudaEventRecord(start, c->comm_stream));
        }
        for (int d = 0; d < P; d++) {
          const int q = (d + c->rank + 1) % P;  // every rank starts with a different peer
          const int w = fwd ? chunk_w(q) : c->kxl;
          if (w <= 0) continue;
          cudaStream_t st = c->ce_streams[d];
          INS_CUDA(c, cudaStreamWaitEvent(st, start, 0));
          if (fwd)  // my block for q -> q's transposed spectrum W_q[(rank * rows + r0)][.]
            INS_CUDA(c, cudaMemcpyAsync((C *)c->peerB[q] + ((i64)c->rank * rows + r0) * w, PS + rows * (i64)q * c->kxc + r0 * w,
                                        (size_t)rb * w * sizeof(C), cudaMemcpyDeviceToDevice, st));
          else      // my solved block for q's slab -> q's staging R_q[rank][r0 ..][.]
            INS_CUDA(c, cudaMemcpyAsync((C *)c->peerP[q] + rows * (i64)c->kx0 + r0 * w, B + ((i64)q * rows + r0) * w,
                                        (size_t)rb * w * sizeof(C), cudaMemcpyDeviceToDevice, st));
          cudaEvent_t ev = c->ce_events[(size_t)b * P + d];
          INS_CUDA(c, cudaEventRecord(ev, st));
          INS_CUDA(c, cudaStreamWaitEvent(c->comm_stream, ev, 0));
        }
        return 0;
      }
      if (!ce) {
        if (fwd) k_fft_scatter2_fwd<C><<<sgrid(rb), sblk, 0, c->comm_stream>>>(pB, S, c->nxh, c->nxp, rows, c->kxc, c->rank, P, r0, r0 + rb);
        else if (c->kxl > 0) k_fft_scatter2_bwd<C><<<sgrid(rb), sblk, 0, c->comm_stream>>>(pS, B, c->nxp, rows, c->kxl, c->kx0, c->rank, P, r0, r0 + rb);
        CHECK_LAUNCH(c);
        return 0;
      }
      for (int d = 0; d < P; d++) {
        const int q = (d + c->rank + 1) % P;  // every rank starts with a different peer
        const int w = fwd ? chunk_w(q) : c->kxl;
        if (w <= 0) continue;
        cudaStream_t st = c->ce_streams[d];
        INS_CUDA(c, cudaStreamWaitEvent(st, after, 0));
        if (fwd)
          INS_CUDA(c, cudaMemcpy2DAsync((C *)c->peerB[q] + ((i64)c->rank * rows + r0) * w, (size_t)w * sizeof(C), S + r0 * c->nxp + (i64)q * c->kxc,
                                        (size_t)c->nxp * sizeof(C), (size_t)w * sizeof(C), (size_t)rb, cudaMemcpyDeviceToDevice, st));
        else
          INS_CUDA(c, cudaMemcpy2DAsync((C *)c->peerS[q] + r0 * c->nxp + c->kx0, (size_t)c->nxp * sizeof(C), B + ((i64)q * rows + r0) * w,
                                        (size_t)w * sizeof(C), (size_t)w * sizeof(C), (size_t)rb, cudaMemcpyDeviceToDevice, st));
        cudaEvent_t ev = c->ce_events[(size_t)b * P + d];
        INS_CUDA(c, cudaEventRecord(ev, st));
        INS_CUDA(c, cudaStreamWaitEvent(c->comm_stream, ev, 0));
      }
      return 0;
    };
    {
      ScopedTimer tm(c, TM_FFT_FWD);
      for (int b = 0; b < NB; b++) {
        T *r = pr + (i64)b * c->fft_batch * c->csz;
        C *s = S + (i64)b * c->fft_batch * c->fft_cdist;
        if (sizeof(T) == 4) INS_CUFFT(c, cufftExecR2C(c->plan_fwd_in, (cufftReal *)r, (cufftComplex *)s));
        else INS_CUFFT(c, cufftExecD2Z(c->plan_fwd_in, (cufftDoubleReal *)r, (cufftDoubleComplex *)s));
        INS_CUDA(c, cudaEventRecord(c->pipe_ev[b], c->stream));
        INS_TRY(transpose_chunk(b, true, c->pipe_ev[b]));
      }
      c->fft_execs++;
    }
    {
      ScopedTimer tm(c, TM_A2A);
      INS_CUDA(c, cudaEventRecord(c->pipe_sync, c->comm_stream));
      INS_CUDA(c, cudaStreamWaitEvent(c->stream, c->pipe_sync, 0));
      INS_TRY(peer_barrier(c));
    }
    if (c->kxl > 0) INS_TRY((launch_zsolve<T>(c, B, (i64)c->nglob[1] * c->kxl, c->kxl, c->kx0, invN)));
    INS_CUDA(c, cudaEventRecord(c->pipe_sync, c->stream));
    for (int b = 0; b < NB; b++) {
      INS_TRY(transpose_chunk(b, false, c->pipe_sync));
      INS_CUDA(c, cudaEventRecord(c->pipe_ev[b], c->comm_stream));
    }
    {
      ScopedTimer tm(c, TM_FFT_INV);
      for (int b = 0; b < NB; b++) {
        INS_CUDA(c, cudaStreamWaitEvent(c->stream, c->pipe_ev[b], 0));
        INS_TRY(peer_barrier(c));  // chunk b of every rank's spectrum is complete
        if (packed) {              // ... in the staging buffer: unpack it into the in-slab spectrum
          k_fft_unpack<C><<<lgrid(rb), sblk, 0, c->stream>>>(S, PS, c->nxh, c->nxp, rows, c->kxc, P, b * rb, (b + 1) * rb);
          CHECK_LAUNCH(c);
        }
        T *r = pr + (i64)b * c->fft_batch * c->csz;
        C *s = S + (i64)b * c->fft_batch * c->fft_cdist;
        if (sizeof(T) == 4) INS_CUFFT(c, cufftExecC2R(c->plan_inv_in, (cufftComplex *)s, (cufftReal *)r));
        else INS_CUFFT(c, cufftExecZ2D(c->plan_inv_in, (cufftDoubleComplex *)s, (cufftDoubleReal *)r));
      }
      c->fft_execs++;
    }
    return 0;
  }
  INS_TRY((exec_fwd_in<T, C>(c, pr, S)));
  if (c->p2p) {
    {
      ScopedTimer tm(c, TM_SCATTER);
      k_fft_scatter2_fwd<C><<<sgrid(rows), sblk, 0, c->stream>>>(pB, S, c->nxh, c->nxp, rows, c->kxc, c->rank, P, 0, rows);
      CHECK_LAUNCH(c);
    }
    ScopedTimer tm(c, TM_A2A);
    INS_TRY(peer_barrier(c));
  } else {
    {
      ScopedTimer tm(c, TM_PACK);
      k_fft_pack<C><<<pgrd, pblk, 0, c->stream>>>(B, S, c->nxh, c->nxp, rows, c->kxc, P, 0);
      CHECK_LAUNCH(c);
    }
    {
      ScopedTimer tm(c, TM_A2A);
      INS_NCCL(c, ncclGroupStart());
      for (int q = 0; q < P; q++) {
        const i64 ns = rows * chunk_w(q) * 2, nr = rows * c->kxl * 2;
        if (ns) INS_NCCL(c, ncclSend((T *)(B + (i64)q * c->kxc * rows), ns, dt, q, c->comm, c->stream));
        if (nr) INS_NCCL(c, ncclRecv((T *)(S + (i64)q * rows * c->kxl), nr, dt, q, c->comm, c->stream));
      }
      INS_NCCL(c, ncclGroupEnd());
    }
  }
  if (c->kxl > 0 && D == 3 && zsolve_supported(c)) {
    INS_TRY((launch_zsolve<T>(c, W, (i64)c->nglob[1] * c->kxl, c->kxl, c->kx0, invN)));
  } else if (c->kxl > 0) {
    {
      ScopedTimer tm(c, TM_FFT_FWD);
      if (sizeof(T) == 4) INS_CUFFT(c, cufftExecC2C(c->plan_last, (cufftComplex *)W, (cufftComplex *)W, CUFFT_FORWARD));
      else INS_CUFFT(c, cufftExecZ2Z(c->plan_last, (cufftDoubleComplex *)W, (cufftDoubleComplex *)W, CUFFT_FORWARD));
      c->fft_execs++;
    }
    {
      ScopedTimer tm(c, TM_POISSON_SCALE);
      // layout (kxl fastest, ky, kz) in 3-D; (kxl fastest, ky) in 2-D
      const int nky = c->nglob[1], nkz = D == 3 ? c->nglob[2] : 1;
      dim3 blk(64, 4, 1), grd((c->kxl + 63) / 64, (nky + 3) / 4, nkz);
      k_poisson_scale<T, C, D><<<grd, blk, 0, c->stream>>>(W, (const T *)c->ahat[0], (const T *)c->ahat[1], (const T *)c->ahat[2],
                                                         c->kxl, nky, nkz, c->kx0, invN);
      CHECK_LAUNCH(c);
    }
    {
      ScopedTimer tm(c, TM_FFT_INV);
      if (sizeof(T) == 4) INS_CUFFT(c, cufftExecC2C(c->plan_last, (cufftComplex *)W, (cufftComplex *)W, CUFFT_INVERSE));
      else INS_CUFFT(c, cufftExecZ2Z(c->plan_last, (cufftDoubleComplex *)W, (cufftDoubleComplex *)W, CUFFT_INVERSE));
      c->fft_execs++;
    }
  }
  if (c->p2p) {
    if (c->kxl > 0) {
      ScopedTimer tm(c, TM_SCATTER);
      k_fft_scatter2_bwd<C><<<sgrid(rows), sblk, 0, c->stream>>>(pS, W, c->nxp, rows, c->kxl, c->kx0, c->rank, P, 0, rows);
      CHECK_LAUNCH(c);
    }
    ScopedTimer tm(c, TM_A2A);
    INS_TRY(peer_barrier(c));
  } else {
    {
      ScopedTimer tm(c, TM_A2A);
      INS_NCCL(c, ncclGroupStart());
      for (int q = 0; q < P; q++) {
        const i64 ns = rows * c->kxl * 2, nr = rows * chunk_w(q) * 2;
        if (ns) INS_NCCL(c, ncclSend((T *)(S + (i64)q * rows * c->kxl), ns, dt, q, c->comm, c->stream));
        if (nr) INS_NCCL(c, ncclRecv((T *)(B + (i64)q * c->kxc * rows), nr, dt, q, c->comm, c->stream));
      }
      INS_NCCL(c, ncclGroupEnd());
    }
    {
      ScopedTimer tm(c, TM_PACK);
      k_fft_pack<C><<<pgrd, pblk, 0, c->stream>>>(B, S, c->nxh, c->nxp, rows, c->kxc, P, 1);
      CHECK_LAUNCH(c);
    }
  }
  INS_TRY((exec_inv_in<T, C>(c, S, pr)));
  return 0;
}

}  // namespace insb200

// =================================================================================================
// operators on reference-layout device fields
// =================================================================================================
namespace insb200 {

template <typename T, int D>
static int op_apply_bc(insb200_ctx *c, void *f, int nc) { return bc_fill<T, D>(c, f, nc); }

template <typename T, int D>
static int op_set_bodyforce(insb200_ctx *c, const void *f) {
  INS_TRY(ensure_vec(c, &c->force));
  // `F .+= bodyforce` (operators.jl:866) runs over the whole array: keep the ghost values too
  i64 e = D;
  for (int a = 0; a < D; a++) e *= c->nloc[a] + 2;
  INS_TRY(dev_alloc(c, &c->force_g, (size_t)e * sizeof(T)));
  INS_CUDA(c, cudaMemcpyAsync(c->force_g, f, (size_t)e * sizeof(T), cudaMemcpyDeviceToDevice, c->stream));
  return import_field<T, D>(c, c->force, f, D);
}

// unsteady body force: ask the caller for bodyforce.(a, x..., t) (insb200_set_bodyforce_callback) and take it as the
// stored force of the launches that follow
template <typename T, int D>
static int update_unsteady_force(insb200_ctx *c, double t) {
  if (!c->force_cb) return 0;
  const void *f = nullptr;
  const int rc = c->force_cb(c->force_user, t, &f);
  if (rc != 0 || !f) INS_FAIL(c, "the body-force callback failed (returned " + std::to_string(rc) + (f ? ")" : ", no field)"));
  return op_set_bodyforce<T, D>(c, f);
}
// time of RK stage i of the step that starts at c->t_step: `t = tstart + c[i] * dt` is set AFTER momentum! of stage i
// (step_explicit_runge_kutta.jl:21,32), arithmetic in T
template <typename T>
static double stage_time(const insb200_ctx *c, const double *cc, int i, double dt) {
  return i == 0 ? (double)T(c->t_step) : (double)(T(c->t_step) + T(cc[i - 1]) * T(dt));
}

// F = convdiff(u) [+ force], written to any view
template <typename T, int D>
static int launch_momentum(insb200_ctx *c, View<T> F, View<const T> u, Geo g, int terms, bool with_force) {
  ScopedTimer tm(c, TM_MOMENTUM);
  const T visc = T(1) / T(c->Re);
  View<const T> fv{nullptr, 0, 0, 0};
  if (with_force && c->force) fv = view_compact_c<T>(c, c->force);
  dim3 grd = cell_grid(c->nloc), blk = cell_block();
  Rh<T> rh = make_rh<T>(c);
  if (terms == 3) k_momentum<T, D, true, true><<<grd, blk, 0, c->stream>>>(F, u, fv, g, visc, rh);
  else if (terms == 1) k_momentum<T, D, true, false><<<grd, blk, 0, c->stream>>>(F, u, fv, g, visc, rh);
  else k_momentum<T, D, false, true><<<grd, blk, 0, c->stream>>>(F, u, fv, g, visc, rh);
  CHECK_LAUNCH(c);
  return 0;
}

static i64 ghosted_elems(const insb200_ctx *c) {
  i64 e = 1;
  for (int a = 0; a < c->D; a++) e *= c->nloc[a] + 2;
  return e;
}

template <typename T, int D>
static int op_momentum(insb200_ctx *c, void *F, const void *u, int terms, bool with_force) {
  // fill!(F, 0) (operators.jl:971) then, with a force, `F .+= bodyforce` over the whole array
  // (:866): ghosts of F are 0 resp. the ghost values of the stored force field
  const size_t fb = (size_t)ghosted_elems(c) * D * sizeof(T);
  if (with_force && c->force_g) INS_CUDA(c, cudaMemcpyAsync(F, c->force_g, fb, cudaMemcpyDeviceToDevice, c->stream));
  else INS_CUDA(c, cudaMemsetAsync(F, 0, fb, c->stream));
  return launch_momentum<T, D>(c, view_ghosted<T>(c, F), view_ghosted_c<T>(c, u), geo_ghosted(c), terms, with_force);
}

template <typename T>
static T eddy_const(const insb200_ctx *c, double theta) {
  // theta^2 * d^2 with d = sqrt(sum h_a^2)  (operators.jl:1128, 1146)
  T s = T(0);
  for (int a = 0; a < c->D; a++) s = (a == 0) ? T(c->h[a]) * T(c->h[a]) : s + T(c->h[a]) * T(c->h[a]);
  const T d = std::sqrt(s);
  const T th = T(theta);
  return (th * th) * (d * d);
}

// sigma(u) on all local p-cells into c->sigma, then its halo on the decomposed axis
template <typename T, int D>
static int compute_sigma(insb200_ctx *c, View<const T> u, Geo gu, double theta, const T *pin = nullptr) {
  INS_TRY(ensure_sigma(c));
  {
    ScopedTimer tm(c, TM_SMAGTENSOR);
    k_smagtensor<T, D><<<cell_grid(c->nloc), cell_block(), 0, c->stream>>>(view_compact<T>(c, c->sigma), u, pin, gu,
                                                                        eddy_const<T>(c, theta), make_rh<T>(c));
    CHECK_LAUNCH(c);
  }
  return halo_compact<T>(c, c->sigma, D * (D + 1) / 2, 1, 1);
}

template <typename T, int D>
static int op_smagorinsky(insb200_ctx *c, void *s, const void *u, double theta) {
  INS_TRY((compute_sigma<T, D>(c, view_ghosted_c<T>(c, u), geo_ghosted(c), theta)));
  INS_CUDA(c, cudaMemsetAsync(s, 0, (size_t)ghosted_elems(c) * D * sizeof(T), c->stream));
  ScopedTimer tm(c, TM_DIVTENSOR);
  k_divtensor<T, D><<<cell_grid(c->nloc), cell_block(), 0, c->stream>>>(view_ghosted<T>(c, s), view_compact_c<T>(c, c->sigma),
                                                                     geo_compact(c), make_rh<T>(c));
  CHECK_LAUNCH(c);
  return 0;
}

template <typename T, int D>
static int launch_divergence(insb200_ctx *c, View<T> out, View<const T> u, Geo g, bool scale) {
  ScopedTimer tm(c, TM_DIVERGENCE);
  if constexpr (D == 2) {
    if (c->rows2d) {
      constexpr int RPT = 4;
      dim3 grd((c->nloc[0] + 63) / 64, (c->nloc[1] + 4 * RPT - 1) / (4 * RPT), 1);
      k_divergence_y<T, RPT><<<grd, cell_block(), 0, c->stream>>>(out, u, g, make_rh<T>(c), volume<T>(c), scale ? 1 : 0);
      CHECK_LAUNCH(c);
      return 0;
    }
  }
  k_divergence<T, D><<<cell_grid(c->nloc), cell_block(), 0, c->stream>>>(out, u, g, make_rh<T>(c), volume<T>(c), scale ? 1 : 0);
  CHECK_LAUNCH(c);
  return 0;
}

template <typename T, int D>
static int op_divergence(insb200_ctx *c, void *div, const void *u) {
  View<T> o = view_ghosted<T>(c, div);
  return launch_divergence<T, D>(c, o, view_ghosted_c<T>(c, u), geo_ghosted(c), false);
}

template <typename T, int D>
static int op_vorticity(insb200_ctx *c, void *w, const void *u) {
  ScopedTimer tm(c, TM_MISC);
  const int N0 = c->nloc[0] + 2, N1 = c->nloc[1] + 2, N2 = D == 3 ? c->nloc[2] + 2 : 1;
  dim3 blk(64, 4, 1), grd((N0 - 1 + 63) / 64, (N1 - 1 + 3) / 4, D == 3 ? N2 - 1 : 1);
  k_vorticity<T, D><<<grd, blk, 0, c->stream>>>((T *)w, (const T *)u, N0, N1, N2, make_rh<T>(c));
  CHECK_LAUNCH(c);
  return 0;
}
template <typename T, int D>
static int op_interp_p(insb200_ctx *c, void *out, const void *in, int mode) {
  ScopedTimer tm(c, TM_MISC);
  if (mode == 0) k_interp_p<T, D, 0><<<cell_grid(c->nloc), cell_block(), 0, c->stream>>>(view_ghosted<T>(c, out), view_ghosted_c<T>(c, in), geo_ghosted(c));
  else k_interp_p<T, D, 1><<<cell_grid(c->nloc), cell_block(), 0, c->stream>>>(view_ghosted<T>(c, out), view_ghosted_c<T>(c, in), geo_ghosted(c));
  CHECK_LAUNCH(c);
  return 0;
}
template <typename T, int D>
static int op_velocitynorm(insb200_ctx *c, void *out, const void *up) {
  ScopedTimer tm(c, TM_MISC);
  const i64 n = ghosted_elems(c);
  k_velocitynorm<T, D><<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>((T *)out, (const T *)up, n);
  CHECK_LAUNCH(c);
  return 0;
}

template <typename T, int D>
static int op_scalewithvolume(insb200_ctx *c, void *p) {
  ScopedTimer tm(c, TM_MISC);
  const i64 n = ghosted_elems(c);
  k_scale<T><<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>((T *)p, n, volume<T>(c));
  CHECK_LAUNCH(c);
  return 0;
}

template <typename T, int D, int MODE>
static int launch_gradient(insb200_ctx *c, View<T> out, View<const T> p, Geo gp) {
  ScopedTimer tm(c, TM_GRADIENT);
  if constexpr (D == 3) {
    constexpr int ZPT = 4;
    dim3 grd((c->nloc[0] + 63) / 64, (c->nloc[1] + 3) / 4, (c->nloc[2] + ZPT - 1) / ZPT);
    k_gradient_z<T, MODE, ZPT><<<grd, cell_block(), 0, c->stream>>>(out, p, gp, make_rh<T>(c));
  } else if (c->rows2d) {
    constexpr int RPT = 4;
    dim3 grd((c->nloc[0] + 63) / 64, (c->nloc[1] + 4 * RPT - 1) / (4 * RPT), 1);
    k_gradient_y<T, MODE, RPT><<<grd, cell_block(), 0, c->stream>>>(out, p, gp, make_rh<T>(c));
  } else
  k_gradient<T, D, MODE><<<cell_grid(c->nloc), cell_block(), 0, c->stream>>>(out, p, gp, make_rh<T>(c));
  CHECK_LAUNCH(c);
  return 0;
}

template <typename T, int D>
static int op_pressuregradient(insb200_ctx *c, void *G, const void *p) {
  View<const T> pv = view_ghosted_c<T>(c, p);
  return launch_gradient<T, D, 0>(c, view_ghosted<T>(c, G), pv, geo_ghosted(c));
}
template <typename T, int D>
static int op_applypressure(insb200_ctx *c, void *u, const void *p) {
  View<const T> pv = view_ghosted_c<T>(c, p);
  return launch_gradient<T, D, 1>(c, view_ghosted<T>(c, u), pv, geo_ghosted(c));
}

template <typename T, int D>
static int op_psolve(insb200_ctx *c, void *p) {
  INS_TRY(ensure_fft(c));
  INS_TRY((import_field<T, D>(c, c->preal, p, 1)));
  INS_TRY((poisson_solve<T, D>(c)));
  return export_field<T, D>(c, p, c->preal, 1);
}

// project on any view of u (reference layout or compact): divergence*Omega -> solve -> u -= grad p
template <typename T, int D>
static int project_view(insb200_ctx *c, View<T> u, Geo gu) {
  INS_TRY(ensure_fft(c));
  View<const T> uc{u.p, u.sy, u.sz, u.sc};
  INS_TRY((launch_divergence<T, D>(c, view_compact<T>(c, c->preal), uc, gu, true)));
  INS_TRY((poisson_solve<T, D>(c)));
  INS_TRY((halo_compact<T>(c, c->preal, 1, 0, 1)));
  return launch_gradient<T, D, 1>(c, u, view_compact_c<T>(c, c->preal), geo_compact(c));
}

template <typename T, int D>
static int export_p(insb200_ctx *c, void *p) {
  INS_TRY((export_field<T, D>(c, p, c->preal, 1)));
  return bc_fill<T, D>(c, p, 1);
}

template <typename T, int D>
static int op_project(insb200_ctx *c, void *u, void *p) {
  INS_TRY((project_view<T, D>(c, view_ghosted<T>(c, u), geo_ghosted(c))));
  if (p) INS_TRY((export_p<T, D>(c, p)));
  return 0;
}

template <typename T, int D>
static int op_pressure(insb200_ctx *c, void *p, const void *u) {
  INS_TRY(ensure_fft(c));
  INS_TRY(ensure_vec(c, &c->fwork));
  // momentum!(F,u) -> apply_bc_u!(F; dudt) (periodic: wrap / halo) -> divergence -> scale -> solve -> bc
  INS_TRY((launch_momentum<T, D>(c, view_compact<T>(c, c->fwork), view_ghosted_c<T>(c, u), geo_ghosted(c), 3, true)));
  INS_TRY((halo_compact<T>(c, c->fwork, D, 1, 0)));
  INS_TRY((launch_divergence<T, D>(c, view_compact<T>(c, c->preal), view_compact_c<T>(c, c->fwork), geo_compact(c), true)));
  INS_TRY((poisson_solve<T, D>(c)));
  return export_p<T, D>(c, p);
}

// ---- reductions ---------------------------------------------------------------------------------
// out (host, 4 doubles): [0] kinetic energy sum, [1..3] max |u_a|  -- over all ranks
template <typename T, int D>
static int reduce_view(insb200_ctx *c, View<const T> u, Geo g, double *out4) {
  {
    ScopedTimer tm(c, TM_REDUCE);
    dim3 grd = cell_grid(c->nloc);
    const i64 nb = (i64)grd.x * grd.y * grd.z;
    double *part_ke = (double *)c->red, *part_max = part_ke + nb, *out = part_max + 3 * nb;
    k_reduce_partials<T, D><<<grd, cell_block(), 0, c->stream>>>(u, g, volume<T>(c), part_ke, part_max);
    CHECK_LAUNCH(c);
    k_reduce_final<<<1, 1024, 0, c->stream>>>(part_ke, part_max, nb, out);
    CHECK_LAUNCH(c);
    if (c->nranks > 1) {
      INS_NCCL(c, ncclGroupStart());
      INS_NCCL(c, ncclAllReduce(out, out, 1, ncclDouble, ncclSum, c->comm, c->stream));
      INS_NCCL(c, ncclAllReduce(out + 1, out + 1, 3, ncclDouble, ncclMax, c->comm, c->stream));
      INS_NCCL(c, ncclGroupEnd());
    }
    INS_CUDA(c, cudaMemcpyAsync(c->red_host, out, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  }
  INS_CUDA(c, cudaStreamSynchronize(c->stream));
  for (int q = 0; q < 4; q++) out4[q] = c->red_host[q];
  return 0;
}

// get_cfl_timestep! (solver.jl:101-125) from the per-component maxima, arithmetic in T
template <typename T>
static double cfl_from_max(const insb200_ctx *c, const double *out4) {
  T dt = std::numeric_limits<T>::infinity();
  for (int a = 0; a < c->D; a++) {
    const T h = T(c->h[a]);
    const T dt_diff = T(c->Re) * (h * h) / T(2);
    const T dt_conv = h / T(out4[1 + a]);
    dt = std::min(dt, std::min(dt_diff, dt_conv));
  }
  return (double)dt;
}

template <typename T, int D>
static int op_reduce(insb200_ctx *c, const void *u, double *out4) {
  return reduce_view<T, D>(c, view_ghosted_c<T>(c, u), geo_ghosted(c), out4);
}

}  // namespace insb200

// ---- reductions without any allocation (no reference counterpart) ----
// insb200_reduce_noalloc: the fused reduction pass of insb200_kinetic_energy / insb200_maxabs / insb200_cfl_dt
// (k_reduce_partials + k_reduce_final, unchanged) for callers that cannot call cudaMalloc -- code under CUDA graph
// capture, tight serving loops, harnesses that forbid it.  The per-block partials live in the file-scope device array
// below instead of the context's workspace, the result goes to a caller-owned DEVICE buffer, nothing is copied to the
// host and nothing is synchronised.  kNoAllocMaxBlocks blocks of 64 x 4 cells: up to 256^3 in 3-D, 4096^2 in 2-D
// (2 MiB of device memory on every GPU that loads the library, until the process exits).
// In HOST code the array's name is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current (noalloc_scratch below).
namespace insb200 {
constexpr long long kNoAllocMaxBlocks = 65536;
static __device__ __align__(256) double g_noalloc_scratch[4 * kNoAllocMaxBlocks];
static double *g_noalloc_ptr[64] = {nullptr};  // device address of g_noalloc_scratch, per device ordinal

static int noalloc_scratch(insb200_ctx *c, double **out) {
  if (c->device < 0 || c->device >= 64) INS_FAIL(c, "insb200_reduce_noalloc: device ordinal out of range");
  if (!g_noalloc_ptr[c->device]) {  // the context's device is current (CHECK_CTX)
    void *ptr = nullptr;
    INS_CUDA(c, cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
    g_noalloc_ptr[c->device] = (double *)ptr;
  }
  *out = g_noalloc_ptr[c->device];
  return 0;
}

template <typename T, int D>
static int op_reduce_noalloc(insb200_ctx *c, const void *u, double *out4_dev) {
  dim3 grd = cell_grid(c->nloc);
  const i64 nb = (i64)grd.x * grd.y * grd.z;
  if (nb > kNoAllocMaxBlocks)
    INS_FAIL(c, "insb200_reduce_noalloc: grid too large for the static scratch array (" + std::to_string(nb) + " blocks of 64 x 4 cells, maximum " +
                    std::to_string(kNoAllocMaxBlocks) + ": up to 256^3 or 4096^2 cells); use insb200_kinetic_energy / insb200_maxabs");
  double *scr = nullptr;
  INS_TRY(noalloc_scratch(c, &scr));
  // every block writes its own slot and k_reduce_final reads exactly nb slots: nothing has to start at zero, stale
  // contents of an earlier (larger) call are never read
  double *part_ke = scr, *part_max = scr + nb;
  k_reduce_partials<T, D><<<grd, cell_block(), 0, c->stream>>>(view_ghosted_c<T>(c, u), geo_ghosted(c), volume<T>(c), part_ke, part_max);
  CHECK_LAUNCH(c);
  k_reduce_final<<<1, 1024, 0, c->stream>>>(part_ke, part_max, nb, out4_dev);
  CHECK_LAUNCH(c);
  return 0;
}
}  // namespace insb200

extern "C" int insb200_reduce_noalloc(insb200_ctx *c, const void *u, double *out4_dev) {
  CHECK_CTX(c);
  if (!u || !out4_dev) INS_FAIL(c, "insb200_reduce_noalloc: NULL pointer");
  if (c->nranks > 1) INS_FAIL(c, "insb200_reduce_noalloc: one rank only (the all-reduced reductions are insb200_kinetic_energy / insb200_maxabs)");
  return DISPATCH(c, op_reduce_noalloc, c, u, out4_dev);
}

#define API_PROLOG(c) CHECK_CTX(c)

extern "C" int insb200_apply_bc_u(insb200_ctx *c, void *u) {
  API_PROLOG(c);
  if (!u) INS_FAIL(c, "u is NULL");
  return DISPATCH(c, op_apply_bc, c, u, c->D);
}
extern "C" int insb200_apply_bc_p(insb200_ctx *c, void *p) {
  API_PROLOG(c);
  if (!p) INS_FAIL(c, "p is NULL");
  return DISPATCH(c, op_apply_bc, c, p, 1);
}
extern "C" int insb200_set_bodyforce(insb200_ctx *c, const void *f) {
  API_PROLOG(c);
  clear_graphs(c);
  if (!f) {
    if (c->force) {
      INS_CUDA(c, cudaStreamSynchronize(c->stream));
      cudaFree(c->force); cudaFree(c->force_g);
      c->force = nullptr; c->force_g = nullptr;
    }
    return 0;
  }
  return DISPATCH(c, op_set_bodyforce, c, f);
}
extern "C" int insb200_set_bodyforce_callback(insb200_ctx *c, insb200_force_fn fn, void *user) {
  API_PROLOG(c);
  clear_graphs(c);
  c->force_cb = fn;
  c->force_user = user;
  if (!fn) return insb200_set_bodyforce(c, nullptr);
  return 0;
}
extern "C" int insb200_momentum(insb200_ctx *c, void *F, const void *u, double t) {
  API_PROLOG(c);
  if (!F || !u) INS_FAIL(c, "NULL field");
  if (c->force_cb) INS_TRY(DISPATCH(c, update_unsteady_force, c, t));
  return DISPATCH(c, op_momentum, c, F, u, 3, true);
}
extern "C" int insb200_convectiondiffusion(insb200_ctx *c, void *F, const void *u, int terms) {
  API_PROLOG(c);
  if (!F || !u) INS_FAIL(c, "NULL field");
  if (terms < 1 || terms > 3) INS_FAIL(c, "terms must be 1 (convection), 2 (diffusion) or 3 (both)");
  return DISPATCH(c, op_momentum, c, F, u, terms, false);
}
extern "C" int insb200_smagorinsky(insb200_ctx *c, void *s, const void *u, double theta) {
  API_PROLOG(c);
  if (!s || !u) INS_FAIL(c, "NULL field");
  return DISPATCH(c, op_smagorinsky, c, s, u, theta);
}
extern "C" int insb200_divergence(insb200_ctx *c, void *div, const void *u) {
  API_PROLOG(c);
  if (!div || !u) INS_FAIL(c, "NULL field");
  return DISPATCH(c, op_divergence, c, div, u);
}
static int observers_single_rank(insb200_ctx *c) {
  if (c->nranks > 1) INS_FAIL(c, "observer fields are computed per slab by the caller: not available with nranks > 1 yet");
  return 0;
}
extern "C" int insb200_vorticity(insb200_ctx *c, void *w, const void *u) {
  API_PROLOG(c);
  if (!w || !u) INS_FAIL(c, "NULL field");
  INS_TRY(observers_single_rank(c));
  return DISPATCH(c, op_vorticity, c, w, u);
}
extern "C" int insb200_interpolate_u_p(insb200_ctx *c, void *up, const void *u) {
  API_PROLOG(c);
  if (!up || !u) INS_FAIL(c, "NULL field");
  INS_TRY(observers_single_rank(c));
  return DISPATCH(c, op_interp_p, c, up, u, 0);
}
extern "C" int insb200_interpolate_w_p(insb200_ctx *c, void *wp, const void *w) {
  API_PROLOG(c);
  if (!wp || !w) INS_FAIL(c, "NULL field");
  INS_TRY(observers_single_rank(c));
  return DISPATCH(c, op_interp_p, c, wp, w, 1);
}
extern "C" int insb200_velocitynorm(insb200_ctx *c, void *out, const void *up) {
  API_PROLOG(c);
  if (!out || !up) INS_FAIL(c, "NULL field");
  INS_TRY(observers_single_rank(c));
  return DISPATCH(c, op_velocitynorm, c, out, up);
}
extern "C" int insb200_scalewithvolume(insb200_ctx *c, void *p) {
  API_PROLOG(c);
  if (!p) INS_FAIL(c, "NULL field");
  return DISPATCH(c, op_scalewithvolume, c, p);
}
extern "C" int insb200_pressuregradient(insb200_ctx *c, void *G, const void *p) {
  API_PROLOG(c);
  if (!G || !p) INS_FAIL(c, "NULL field");
  return DISPATCH(c, op_pressuregradient, c, G, p);
}
extern "C" int insb200_applypressure(insb200_ctx *c, void *u, const void *p) {
  API_PROLOG(c);
  if (!u || !p) INS_FAIL(c, "NULL field");
  return DISPATCH(c, op_applypressure, c, u, p);
}
extern "C" int insb200_psolve_spectral(insb200_ctx *c, void *p) {
  API_PROLOG(c);
  if (!p) INS_FAIL(c, "NULL field");
  return DISPATCH(c, op_psolve, c, p);
}
extern "C" int insb200_project(insb200_ctx *c, void *u, void *p) {
  API_PROLOG(c);
  if (!u) INS_FAIL(c, "NULL field");
  return DISPATCH(c, op_project, c, u, p);
}
extern "C" int insb200_pressure(insb200_ctx *c, void *p, const void *u, double t) {
  API_PROLOG(c);
  if (!p || !u) INS_FAIL(c, "NULL field");
  if (c->force_cb) INS_TRY(DISPATCH(c, update_unsteady_force, c, t));
  return DISPATCH(c, op_pressure, c, p, u);
}
extern "C" int insb200_cfl_dt(insb200_ctx *c, const void *u, double *dt) {
  API_PROLOG(c);
  if (!u || !dt) INS_FAIL(c, "NULL argument");
  double o[4];
  INS_TRY(DISPATCH(c, op_reduce, c, u, o));
  *dt = c->dtype == INSB200_F32 ? cfl_from_max<float>(c, o) : cfl_from_max<double>(c, o);
  return 0;
}
extern "C" int insb200_kinetic_energy(insb200_ctx *c, const void *u, double *E) {
  API_PROLOG(c);
  if (!u || !E) INS_FAIL(c, "NULL argument");
  double o[4];
  INS_TRY(DISPATCH(c, op_reduce, c, u, o));
  *E = o[0];
  return 0;
}
extern "C" int insb200_maxabs(insb200_ctx *c, const void *u, double *m) {
  API_PROLOG(c);
  if (!u || !m) INS_FAIL(c, "NULL argument");
  double o[4];
  INS_TRY(DISPATCH(c, op_reduce, c, u, o));
  *m = std::max(o[1], std::max(o[2], o[3]));
  return 0;
}

namespace insb200 {
// ---- tiled 3-D stage kernel (tiled.cuh) ----------------------------------------------------------
template <typename T, int D>
static bool tiled_stage_supported(const insb200_ctx *c) {
  if (D != 3 || c->variant < 1) return false;
  const int ty = c->tile_y == 16 ? 16 : 8;
  return c->nloc[0] % 32 == 0 && c->nloc[1] % ty == 0 && c->nloc[2] >= 2;
}

template <typename T, bool SMAG, bool PGRAD, int TX, int TY>
static int launch_stage3d(insb200_ctx *c, const StageArgs<T> &a, Geo g, T visc, Rh<T> rh) {
  typedef Tile3<TX, TY> TL;
  const size_t smem = (size_t)(12 * TL::NP + 2 * TY * (TX + 1) + 2 * (TY + 1) * TX) * sizeof(T);
  auto kern = k_stage3d<T, SMAG, PGRAD, TX, TY>;
  INS_DYN_SMEM(c, kern, smem);
  const int zc = std::max(2, std::min(c->zchunk > 0 ? c->zchunk : 64, c->nloc[2]));
  dim3 grd(c->nloc[0] / TX, c->nloc[1] / TY, (c->nloc[2] + zc - 1) / zc), blk(TX, TY, 1);
  kern<<<grd, blk, smem, c->stream>>>(a, g, visc, rh, volume<T>(c), zc);
  CHECK_LAUNCH(c);
  return 0;
}

template <typename T, int D>
static int launch_stage_tiled(insb200_ctx *c, const StageArgs<T> &a, Geo g, T visc, Rh<T> rh, bool smag) {
  if (D != 3) INS_FAIL(c, "tiled stage kernel is 3-D only");
  const bool pg = a.pin != nullptr;
#define LS(SM, PG)                                                                            \
  (c->tile_y == 16 ? launch_stage3d<T, SM, PG, 32, 16>(c, a, g, visc, rh) : launch_stage3d<T, SM, PG, 32, 8>(c, a, g, visc, rh))
  if (smag) return pg ? LS(true, true) : LS(true, false);
  return pg ? LS(false, true) : LS(false, false);
#undef LS
}

// divergence of u' on the first plane of every z-chunk of the tiled kernel (needs the minus-side halo of u')
template <typename T>
static int launch_divergence_planes(insb200_ctx *c, T *out, const T *u, Geo g) {
  ScopedTimer tm(c, TM_DIVERGENCE);
  const int zc = std::max(2, std::min(c->zchunk > 0 ? c->zchunk : 64, c->nloc[2]));
  dim3 blk(64, 4, 1), grd((c->nloc[0] + 63) / 64, (c->nloc[1] + 3) / 4, (c->nloc[2] + zc - 1) / zc);
  k_divergence_planes<T><<<grd, blk, 0, c->stream>>>(out, u, c->csy, c->csz, c->ccomp, g, make_rh<T>(c), volume<T>(c), zc);
  CHECK_LAUNCH(c);
  return 0;
}

// ---- TMA plane-pipeline kernels (tma_stage.cuh): padded 3-D layout ---------------------------------
static const int kTX = 32, kTY = 16;

template <typename T, int D>
static bool tma_supported(const insb200_ctx *c) {
  // the closure kernels address the three components of one buffer with UNSIGNED 32-bit element offsets
  // (tma_stage.cuh; the stage kernel uses 64-bit byte offsets per component): a padded vector field must stay
  // below 2^32 elements -- 1024^3 on one GPU qualifies (3.35e9); larger slabs take the LDG-tiled kernels
  return D == 3 && c->padded && c->variant >= 2 && c->nloc[0] % kTX == 0 && c->nloc[1] % kTY == 0 && c->nloc[2] >= 2 &&
         3 * c->ccomp < ((i64)1 << 32);
}

static const int kNS = 5;  // velocity plane slots of the stage kernel (prefetch distance 2 planes)

// Planes marched per block of a z-marching kernel.  `user` > 0: the value set through insb200_set_param.  Otherwise
// pick the chunk that minimises  rounds x (zc + extra)  with rounds = ceil(tiles x chunks / (SMs x blocks per SM)):
// a block costs zc planes plus `extra` planes of pipeline prologue, and the grid runs in rounds of resident blocks.
// Measured against a sweep (tools/sweep_zchunk.sh): 128^3 Float64 stage kernel 64 planes -> 32: 0.112 -> 0.063 ms
// (64 blocks left 84 of 148 SMs idle); 512^3 closure kernel 128 -> 256: 1.89 -> 1.86 ms.
int auto_zchunk(const insb200_ctx *c, int user, int tiles, int blocks_per_sm, int extra) {
  const int nz = c->nloc[2];
  if (user > 0) return std::max(2, std::min(user, nz));
  int best = nz;
  double best_cost = 1e300;
  for (int m = 1; m <= std::max(1, nz / 8); m++) {
    const int zc = (nz + m - 1) / m, chunks = (nz + zc - 1) / zc;
    const double rounds = std::ceil((double)tiles * chunks / ((double)std::max(1, c->nsm) * blocks_per_sm));
    const double cost = rounds * (zc + extra);
    if (cost < best_cost - 1e-9) { best_cost = cost; best = zc; }
  }
  return std::max(2, best);
}

template <typename T, bool PG, bool SA, bool GEN>
static int launch_stage_tma_i(insb200_ctx *c, const TmaStageArgs<T> &a, Geo g, T visc, Rh<T> rh, const void *ubase,
                              const void *pbase) {
  typedef StageCfg<T, PG, kTX, kTY, kNS> Cfg;
  CUtensorMap tu, tp;
  INS_TRY(get_tmap(c, ubase, 3, Cfg::Pipe::BX, kTY + 3, &tu));
  INS_TRY(get_tmap(c, pbase, 1, Cfg::Pipe::BX, kTY + 4, &tp));
  const int zc = auto_zchunk(c, c->zchunk, (c->nloc[0] / kTX) * (c->nloc[1] / kTY), sizeof(T) == 4 ? 2 : 1, 4);
  dim3 grd(c->nloc[0] / kTX, c->nloc[1] / kTY, (c->nloc[2] + zc - 1) / zc), blk(kTX, kTY, 1);
#ifdef INSB200_EXPERIMENTAL
  if constexpr (std::is_same<T, float>::value) if (c->stage_x2) {  // two cells per thread on packed FP32 (tma_stage_x2.cuh)
    auto kex = k_stage3d_tma_x2<f2t<true>, PG, SA, GEN, kTX, kTY, kNS>;    // 1: exact (bit-identical to the scalar kernel)
    auto kfu = k_stage3d_tma_x2<f2t<false>, PG, SA, GEN, kTX, kTY, kNS>;   // 2: products contracted into FFMA2
    INS_DYN_SMEM(c, kex, Cfg::SMEM);
    INS_DYN_SMEM(c, kfu, Cfg::SMEM);
    (c->stage_x2 == 1 ? kex : kfu)<<<grd, dim3(kTX, kTY / 2, 1), Cfg::SMEM, c->stream>>>(tu, tp, a, g, visc, rh, volume<T>(c), zc, c->org[0],
                                                                                       c->org[1], c->org[2]);
    CHECK_LAUNCH(c);
    return 0;
  }
#endif
  auto kern = k_stage3d_tma<T, PG, SA, GEN, kTX, kTY, kNS>;
  INS_DYN_SMEM(c, kern, Cfg::SMEM);
  kern<<<grd, blk, Cfg::SMEM, c->stream>>>(tu, tp, a, g, visc, rh, volume<T>(c), zc, c->org[0], c->org[1], c->org[2]);
  CHECK_LAUNCH(c);
  return 0;
}

// ubase / pbase: START of the padded buffers (descriptors address the whole box); pbase == nullptr: first
// stage (no projection on load, ustart is the stage input itself)
template <typename T>
static int launch_stage_tma(insb200_ctx *c, const TmaStageArgs<T> &a, Geo g, T visc, Rh<T> rh, const void *ubase,
                            const void *pbase) {
  const bool sa = a.sadd[0] != nullptr, gen = a.nk > 0 || a.kout[0] == nullptr;
  const void *pb = pbase ? pbase : c->preal;
#define LST(PG, SA, GEN) launch_stage_tma_i<T, PG, SA, GEN>(c, a, g, visc, rh, ubase, pb)
  if (pbase) return sa ? (gen ? LST(true, true, true) : LST(true, true, false)) : (gen ? LST(true, false, true) : LST(true, false, false));
  return sa ? (gen ? LST(false, true, true) : LST(false, true, false)) : (gen ? LST(false, false, true) : LST(false, false, false));
#undef LST
}

template <typename T, bool PG>
static int launch_smag_tma_i(insb200_ctx *c, T *sout, Geo g, T c_eddy, Rh<T> rh, const void *ubase, const void *pbase) {
  typedef SmagCfg<T, PG, kTX, kTY> Cfg;
  CUtensorMap tu, tp;
  INS_TRY(get_tmap(c, ubase, 3, Cfg::Pipe::BX, kTY + 4, &tu));
  INS_TRY(get_tmap(c, pbase, 1, Cfg::Pipe::BX, kTY + 5, &tp));
  if constexpr (sizeof(T) == 4) if (c->smag_v >= 2) {  // register-carried z-march (F32)
    const int ty2 = (c->smag_ty == 8 && c->nloc[1] % 8 == 0) ? 8 : kTY;
    const int zc2 = auto_zchunk(c, c->smag_zchunk, (c->nloc[0] / kTX) * (c->nloc[1] / ty2), ty2 == 8 ? 2 : 1, 5);
    if (c->smag_ty == 8 && c->nloc[1] % 8 == 0) {  // two blocks of 32 x 12 threads per SM
      typedef Smag2Cfg<T, PG, kTX, 8> Cfg2;
      CUtensorMap tu8, tp8;
      INS_TRY(get_tmap(c, ubase, 3, Cfg2::Pipe::BX, 8 + 4, &tu8));
      INS_TRY(get_tmap(c, pbase, 1, Cfg2::Pipe::BX, 8 + 5, &tp8));
      auto kern2 = k_smag3d_v2<T, PG, kTX, 8>;
      INS_DYN_SMEM(c, kern2, Cfg2::SMEM);
      dim3 grd2(c->nloc[0] / kTX, c->nloc[1] / 8, (c->nloc[2] + zc2 - 1) / zc2), blk2(kTX, 8 + 4, 1);
      kern2<<<grd2, blk2, Cfg2::SMEM, c->stream>>>(tu8, tp8, sout, c->csy, c->csz, c->ccomp, g, c_eddy, rh, zc2, c->org[0],
                                                c->org[1], c->org[2]);
      CHECK_LAUNCH(c);
      return 0;
    }
    // one block of 32 x 20 threads per SM
    typedef Smag2Cfg<T, PG, kTX, kTY> Cfg2;
    auto kern2 = k_smag3d_v2<T, PG, kTX, kTY>;
    INS_DYN_SMEM(c, kern2, Cfg2::SMEM);
    dim3 grd2(c->nloc[0] / kTX, c->nloc[1] / kTY, (c->nloc[2] + zc2 - 1) / zc2), blk2(kTX, kTY + 4, 1);
    kern2<<<grd2, blk2, Cfg2::SMEM, c->stream>>>(tu, tp, sout, c->csy, c->csz, c->ccomp, g, c_eddy, rh, zc2, c->org[0], c->org[1],
                                              c->org[2]);
    CHECK_LAUNCH(c);
    return 0;
  }
  auto kern = k_smag3d_tma<T, PG, kTX, kTY>;
  INS_DYN_SMEM(c, kern, Cfg::SMEM);
  const int zc = auto_zchunk(c, c->zchunk, (c->nloc[0] / kTX) * (c->nloc[1] / kTY), sizeof(T) == 4 ? 2 : 1, 5);
  dim3 grd(c->nloc[0] / kTX, c->nloc[1] / kTY, (c->nloc[2] + zc - 1) / zc), blk(kTX, kTY, 1);
  kern<<<grd, blk, Cfg::SMEM, c->stream>>>(tu, tp, sout, c->csy, c->csz, c->ccomp, g, c_eddy, rh, zc, c->org[0], c->org[1],
                                          c->org[2]);
  CHECK_LAUNCH(c);
  return 0;
}

// s = smagorinsky_closure(u) into c->sbuf (interior) from the padded velocity buffer (ghosts valid, 2 layers)
template <typename T>
static int launch_smag_tma(insb200_ctx *c, double theta, const void *ubase, const void *pbase) {
  ScopedTimer tm(c, TM_SMAG_FUSED);
  Geo g = geo_compact(c);
  T *sout = (T *)c->sbuf + c->corg;
  if (pbase) return launch_smag_tma_i<T, true>(c, sout, g, eddy_const<T>(c, theta), make_rh<T>(c), ubase, pbase);
  return launch_smag_tma_i<T, false>(c, sout, g, eddy_const<T>(c, theta), make_rh<T>(c), ubase, c->preal);
}

// fused momentum + stage combination, one thread per cell (any D, any n)
template <typename T, int D>
static int launch_stage(insb200_ctx *c, const StageArgs<T> &a, Geo g, T visc, Rh<T> rh, bool smag) {
  dim3 grd = cell_grid(c->nloc), blk = cell_block();
  if (smag) k_momentum_stage<T, D, true><<<grd, blk, 0, c->stream>>>(a, g, visc, rh);
  else k_momentum_stage<T, D, false><<<grd, blk, 0, c->stream>>>(a, g, visc, rh);
  CHECK_LAUNCH(c);
  return 0;
}
}  // namespace insb200

// =================================================================================================
// time stepping (compact layout)
// =================================================================================================
namespace insb200 {

struct Tableau {
  int s;
  const double *A;  // shifted, row-major s*s
  const double *c;
  bool need[INSB200_MAX_STAGES];  // k_j is read by a later stage
};

static int make_tableau(insb200_ctx *c, int s, const double *A, const double *cc, Tableau &tb) {
  if (s < 1 || s > INSB200_MAX_STAGES) INS_FAIL(c, "number of stages out of range");
  if (!A || !cc) INS_FAIL(c, "tableau is NULL");
  tb.s = s; tb.A = A; tb.c = cc;
  for (int j = 0; j < s; j++) {
    tb.need[j] = false;
    for (int l = j + 1; l < s; l++) if (A[l * s + j] != 0.0) tb.need[j] = true;
    for (int l = 0; l < j; l++) if (A[l * s + j] != 0.0) INS_FAIL(c, "tableau is not explicit (shifted A must be lower triangular)");
  }
  return 0;
}

static int ensure_rk(insb200_ctx *c, const Tableau &tb, bool smag) {
  INS_TRY(ensure_fft(c));
  for (int b = 0; b < 3; b++) INS_TRY(ensure_vec(c, &c->ubuf[b]));
  for (int j = 0; j < tb.s; j++) if (tb.need[j]) INS_TRY(ensure_vec(c, &c->kbuf[j]));
  if (c->nranks > 1) {  // halo planes of the stage velocities and of the multiplier by peer copies
    INS_TRY(dev_alloc(c, &c->preal2, (size_t)c->ccomp * esize(c)));
    for (int b = 0; b < 3; b++) INS_TRY(register_halo_buffer(c, c->ubuf[b]));
    INS_TRY(register_halo_buffer(c, c->preal));
    INS_TRY(register_halo_buffer(c, c->preal2));
  }
  if (smag && !(c->D == 3 && c->padded && c->variant >= 2)) INS_TRY(ensure_sigma(c));
  return 0;
}

// divergence(u') * Omega on plane 0 only (multi-rank TMA path: the chunk that starts at the slab's first
// plane cannot recompute u'_z of the neighbour rank's last plane)
template <typename T>
static int launch_divergence_plane0(insb200_ctx *c, T *out, const T *u, Geo g) {
  ScopedTimer tm(c, TM_DIVERGENCE);
  dim3 blk(64, 4, 1), grd((c->nloc[0] + 63) / 64, (c->nloc[1] + 3) / 4, 1);
  k_divergence_planes<T><<<grd, blk, 0, c->stream>>>(out, u, c->csy, c->csz, c->ccomp, g, make_rh<T>(c), volume<T>(c), c->nloc[2]);
  CHECK_LAUNCH(c);
  return 0;
}

// one timestep! on the padded 3-D layout with the TMA kernels: per stage [closure kernel] + stage kernel
// (k_i, u', Poisson right-hand side) + ghost refresh + cuFFT solve; the projection u = u' - grad p of
// stages 1..s-1 is applied by the NEXT stage's plane pipeline, the last one by the gradient kernel.
template <typename T>
static int rk_step_tma(insb200_ctx *c, const Tableau &tb, double dt, double theta) {
  const bool smag = !std::isnan(theta);
  const Geo g = geo_compact(c);
  const T visc = T(1) / T(c->Re);
  const Rh<T> rh = make_rh<T>(c);
  const i64 goff = c->corg;
  const bool multi = c->nranks > 1;
  INS_TRY(dev_alloc(c, &c->preal2, (size_t)c->ccomp * sizeof(T)));
  INS_TRY((fill_ghosts<T>(c, c->ubuf[0], 3, 2, 2)));  // apply_bc_u! before momentum (:19)
  int cur = 0;
  bool pending = false;  // ubuf[cur] holds u' of the previous stage; its multiplier p is in preal2
  for (int i = 0; i < tb.s; i++) {
    const int nxt = (cur == 1) ? 2 : 1;
    const bool last = i == tb.s - 1;
    const void *pbase = nullptr;
    if (pending) {
      std::swap(c->preal, c->preal2);  // p of the previous stage -> preal2; preal receives the new right-hand side
      pbase = c->preal2;
    }
    if (c->force_cb) INS_TRY((update_unsteady_force<T, 3>(c, stage_time<T>(c, tb.c, i, dt))));
    int nkrow = 0;  // earlier stage derivatives this tableau row reads
    for (int j = 0; j < i; j++) nkrow += tb.A[i * tb.s + j] != 0.0;
    // fused closure + stage kernel (fast.cu) for rows with at most 3 earlier k_j; `exact` and longer rows keep the two kernels below
    const bool gen_row = nkrow > 0 || !tb.need[i] || c->force;
    const bool fast = stage_fast_supported(c, smag) && nkrow <= 3 && (smag || !gen_row || c->fast_gen);
    if (smag && !fast) {
      INS_TRY(ensure_vec(c, &c->sbuf));
      INS_TRY((launch_smag_tma<T>(c, theta, c->ubuf[cur], pbase)));
    }
    TmaStageArgs<T> a;
    memset(&a, 0, sizeof a);
    for (int q = 0; q < 3; q++) {
      const i64 co = goff + q * c->ccomp;
      a.u0[q] = (const T *)c->ubuf[0] + co;
      a.sadd[q] = (smag && !fast) ? (const T *)c->sbuf + co : nullptr;
      a.force[q] = c->force ? (const T *)c->force + co : nullptr;
      a.kout[q] = tb.need[i] ? (T *)c->kbuf[i] + co : nullptr;
      a.unew[q] = (T *)c->ubuf[nxt] + co;
    }
    a.div = (T *)c->preal + goff;
    a.nk = 0;
    for (int j = 0; j < i; j++) {
      if (tb.A[i * tb.s + j] == 0.0) continue;  // exact: adds +-0
      for (int q = 0; q < 3; q++) a.kin[a.nk][q] = (const T *)c->kbuf[j] + goff + q * c->ccomp;
      a.coef[a.nk] = T(dt) * T(tb.A[i * tb.s + j]);  // `dt * A[i,j]` formed first (:37)
      a.nk++;
    }
    a.coef_self = T(dt) * T(tb.A[i * tb.s + i]);
    a.sy = c->csy; a.sz = c->csz; a.sc = c->ccomp;
    {
      ScopedTimer tm(c, (a.nk > 0 || !a.kout[0]) ? TM_STAGE_GEN : TM_STAGE);
      if (fast) INS_TRY((launch_stage_fast<T>(c, a, g, theta, c->ubuf[cur], pbase)));
      else INS_TRY((launch_stage_tma<T>(c, a, g, visc, rh, c->ubuf[cur], pbase)));
    }
    // apply_bc_u! (:48): ghosts of u' for the next stage's boxes (and, multi-rank, for the first plane's divergence)
    if (!last) INS_TRY((fill_ghosts<T>(c, c->ubuf[nxt], 3, 2, 2)));
    else if (multi) INS_TRY((halo_compact<T>(c, c->ubuf[nxt], 3, 1, 0)));
    if (multi) INS_TRY((launch_divergence_plane0<T>(c, (T *)c->preal + goff, (const T *)c->ubuf[nxt] + goff, g)));
    INS_TRY((poisson_solve<T, 3>(c)));
    if (!last) {
      INS_TRY((fill_ghosts<T>(c, c->preal, 1, 2, 3)));  // apply_bc_p! (pressure.jl:78)
      pending = true;
    } else {
      INS_TRY((halo_compact<T>(c, c->preal, 1, 0, 1)));
      INS_TRY((launch_gradient<T, 3, 1>(c, view_compact<T>(c, c->ubuf[nxt]), view_compact_c<T>(c, c->preal), g)));
      pending = false;
    }
    cur = nxt;
  }
  std::swap(c->ubuf[0], c->ubuf[cur]);
  return 0;
}

// fused 2-D stage kernel (stage2d.cuh): rows wrap by index on one rank, ghost rows on slabs; rows of at least 3 cells
template <typename T, int D>
static bool fused2d_supported(const insb200_ctx *c) {
  return D == 2 && c->fuse2d && c->nloc[0] >= 4 && c->nloc[1] >= 4;
}
template <typename T>
static int launch_stage2d(insb200_ctx *c, const StageArgs<T> &a, Geo g, T visc, Rh<T> rh) {
  const int bx = ((c->nloc[0] + kS2Cols - 1) / kS2Cols + kS2Warps - 1) / kS2Warps;
  // rows per block: the march is a serial chain per warp, so many short chunks beat few long ones (4096^2 F32:
  // 0.434 / 0.266 / 0.251 / 0.231 / 0.225 ms at 512 / 256 / 128 / 64 / 32 rows); the pre-iteration costs one row per chunk
  int rc = c->rows_chunk;
  if (rc <= 0) {
    const int chunks = std::max((c->nsm * 16) / bx, (c->nloc[1] + 31) / 32);
    rc = std::min(32, std::max(8, (c->nloc[1] + chunks - 1) / chunks));
  }
  rc = std::min(rc, c->nloc[1]);
  dim3 grd(bx, (c->nloc[1] + rc - 1) / rc, 1);
  if (a.pin) k_stage2d_warp<T, true><<<grd, 32 * kS2Warps, 0, c->stream>>>(a, g, visc, rh, volume<T>(c), rc);
  else k_stage2d_warp<T, false><<<grd, 32 * kS2Warps, 0, c->stream>>>(a, g, visc, rh, volume<T>(c), rc);
  CHECK_LAUNCH(c);
  return 0;
}

// one timestep!: state in ubuf[0] (compact), result in ubuf[0] (buffers rotated).
//
// Tiled path (3-D, tile-aligned n): per stage ONE kernel produces k_i, u' and the Poisson right-hand
// side; the projection u = u' - grad p of stages 1..s-1 is applied by the NEXT stage's loader, so only
// the last stage runs the gradient kernel.  Otherwise: stage kernel, divergence, solve, gradient.
template <typename T, int D>
static int rk_step_compact(insb200_ctx *c, const Tableau &tb, double dt, double theta) {
  if (tma_supported<T, D>(c)) return rk_step_tma<T>(c, tb, dt, theta);
  const bool smag = !std::isnan(theta);
  const Geo g = geo_compact(c);
  const T visc = T(1) / T(c->Re);
  const Rh<T> rh = make_rh<T>(c);
  const bool f2d = !smag && fused2d_supported<T, D>(c);
  const bool tiled = f2d || tiled_stage_supported<T, D>(c);
  const bool fuse_grad = f2d || (tiled && c->fuse_gradient);
  const i64 goff = c->corg;
  T *pr = (T *)c->preal + goff;
  int cur = 0;
  bool pending = false;  // ubuf[cur] holds u' of the previous stage, its gradient correction is still in preal
  for (int i = 0; i < tb.s; i++) {
    const int nxt = (cur == 1) ? 2 : 1;
    if (c->force_cb) INS_TRY((update_unsteady_force<T, D>(c, stage_time<T>(c, tb.c, i, dt))));
    // apply_bc_u! before momentum (step_explicit_runge_kutta.jl:19): wrap by index, halos over NCCL
    if (!pending) INS_TRY((halo_compact<T>(c, c->ubuf[cur], D, f2d ? 2 : 1, 1)));  // (pending: exchanged after the stage kernel)
    if (smag) INS_TRY((compute_sigma<T, D>(c, view_compact_c<T>(c, c->ubuf[cur]), g, theta, pending ? pr : nullptr)));
    StageArgs<T> a;
    memset(&a, 0, sizeof a);
    a.u = (const T *)c->ubuf[cur] + goff;
    a.u0 = (const T *)c->ubuf[0] + goff;
    a.nk = 0;
    for (int j = 0; j < i; j++) {
      if (tb.A[i * tb.s + j] == 0.0) continue;  // exact: adds +-0
      a.kin[a.nk] = (const T *)c->kbuf[j] + goff;
      a.coef[a.nk] = T(dt) * T(tb.A[i * tb.s + j]);  // `dt * A[i,j]` formed first (:37)
      a.nk++;
    }
    a.coef_self = T(dt) * T(tb.A[i * tb.s + i]);
    a.kout = tb.need[i] ? (T *)c->kbuf[i] + goff : nullptr;
    a.unew = (T *)c->ubuf[nxt] + goff;
    a.force = c->force ? (const T *)c->force + goff : nullptr;
    a.sigma = smag ? (const T *)c->sigma + goff : nullptr;
    a.pin = pending ? pr : nullptr;
    a.div = tiled ? pr : nullptr;
    a.sy = c->csy; a.sz = c->csz; a.sc = c->ccomp; a.ssc = c->ccomp;
    if (tiled) {
      // with a pending gradient the kernel reads p while it writes the new right-hand side into the
      // same scalar buffer: keep the multiplier in a second buffer
      if (pending) {
        INS_TRY(dev_alloc(c, &c->preal2, (size_t)c->ccomp * sizeof(T)));
        std::swap(c->preal, c->preal2);
        pr = (T *)c->preal + goff;
        a.pin = (const T *)c->preal2 + goff;
        a.div = pr;
      }
      {
        ScopedTimer tm(c, TM_STAGE);
        if constexpr (D == 2) INS_TRY((launch_stage2d<T>(c, a, g, visc, rh)));  // f2d (tiled implies it in 2-D)
        else INS_TRY((launch_stage_tiled<T, D>(c, a, g, visc, rh, smag)));
      }
      // u' halos: the minus side for the chunk-first planes of the divergence, both for the next stage (the fused 2-D
      // kernel reads two rows below its first one)
      INS_TRY((halo_compact<T>(c, c->ubuf[nxt], D, f2d ? 2 : 1, 1)));
      if constexpr (D == 3) INS_TRY((launch_divergence_planes<T>(c, pr, (const T *)c->ubuf[nxt] + goff, g)));
      if constexpr (D == 2) if (c->nranks > 1) {
        ScopedTimer tm(c, TM_DIVERGENCE);
        k_divergence_row0<T><<<(c->nloc[0] + 255) / 256, 256, 0, c->stream>>>(pr, (const T *)c->ubuf[nxt] + goff, c->csy, c->ccomp, c->nloc[0], rh, volume<T>(c));
        CHECK_LAUNCH(c);
      }
      INS_TRY((poisson_solve<T, D>(c)));
      const bool last = i == tb.s - 1;
      if (fuse_grad && !last) {
        INS_TRY((halo_compact<T>(c, c->preal, 1, f2d ? 2 : 1, 2)));
        pending = true;
      } else {
        INS_TRY((halo_compact<T>(c, c->preal, 1, 0, 1)));
        INS_TRY((launch_gradient<T, D, 1>(c, view_compact<T>(c, c->ubuf[nxt]), view_compact_c<T>(c, c->preal), g)));
        pending = false;
      }
    } else {
      {
        ScopedTimer tm(c, TM_STAGE);
        INS_TRY((launch_stage<T, D>(c, a, g, visc, rh, smag)));
      }
      // apply_bc_u! + project! (:48-49): only the minus-side halo of u' is consumed by the divergence
      INS_TRY((halo_compact<T>(c, c->ubuf[nxt], D, 1, 0)));
      INS_TRY((project_view<T, D>(c, view_compact<T>(c, c->ubuf[nxt]), g)));
    }
    cur = nxt;
  }
  std::swap(c->ubuf[0], c->ubuf[cur]);
  return 0;
}

// ---- CUDA graphs --------------------------------------------------------------------------------------------------
// A step is a fixed sequence of 20-40 launches (own kernels + cuFFT) on one stream; on small grids (128^3, 256^2) the
// launch overhead is a large part of it.  With a fixed dt the sequence repeats with the buffers rotated (period 2-4).
// The first step after any (re)configuration runs eagerly -- lazy allocations, function attributes and tensor-map
// descriptors happen there --; from then on every new (buffers, dt, theta, tableau) state is captured once and
// replayed afterwards.  One rank only (NCCL and the second stream of the pipelined transposes stay outside), not on
// the legacy default stream (capture is not allowed there), not while the per-kernel timers record, and by default
// only for slabs of up to 2^22 cells (above that a step takes milliseconds and the capture would not pay).
static void clear_graphs(insb200_ctx *c) {
  for (auto &e : c->graphs) if (e.exec) cudaGraphExecDestroy(e.exec);
  c->graphs.clear();
  c->graph_warm = false;
  c->graph_misses = 0;
}

template <typename T, int D>
static int rk_step_graphed(insb200_ctx *c, const Tableau &tb, double dt, double theta) {
  const bool wanted = c->use_graphs > 0 || (c->use_graphs < 0 && c->cint <= ((i64)1 << 22));
  const bool usable = wanted && c->nranks == 1 && !c->profiling && !c->force_cb && c->stream != nullptr && c->stream != cudaStreamLegacy &&
                      c->stream != cudaStreamPerThread;
  if (!usable) return rk_step_compact<T, D>(c, tb, dt, theta);
  if (!c->graph_warm) {
    c->graph_warm = true;
    return rk_step_compact<T, D>(c, tb, dt, theta);
  }
  std::vector<unsigned long long> key;
  auto push_ptr = [&](const void *p) { key.push_back((unsigned long long)(uintptr_t)p); };
  auto push_f64 = [&](double v) { unsigned long long b; memcpy(&b, &v, sizeof b); key.push_back(b); };
  for (void *p : c->ubuf) push_ptr(p);
  push_ptr(c->preal); push_ptr(c->preal2); push_ptr((const void *)c->stream);
  push_f64(dt); push_f64(theta);
  key.push_back((unsigned long long)tb.s);
  for (int i = 0; i < tb.s * tb.s; i++) push_f64(tb.A[i]);
  for (auto &e : c->graphs)
    if (e.key == key) {
      INS_CUDA(c, cudaGraphLaunch(e.exec, c->stream));
      for (int b = 0; b < 3; b++) c->ubuf[b] = e.post[b];
      c->preal = e.post[3]; c->preal2 = e.post[4];
      c->launches += e.launches; c->fft_execs += e.fft_execs;
      c->graph_replays++;
      c->graph_misses = 0;
      return 0;
    }
  // adaptive dt: the keys never repeat -- after 8 captures in a row without a replay stop capturing
  if (c->graph_misses >= 8) return rk_step_compact<T, D>(c, tb, dt, theta);
  c->graph_misses++;
  if (c->graphs.size() >= 32) { clear_graphs(c); c->graph_warm = true; }
  // new state: capture.  The step mutates the buffer rotation on the host; keep the state to fall back
  void *pre[5] = {c->ubuf[0], c->ubuf[1], c->ubuf[2], c->preal, c->preal2};
  const long long l0 = c->launches, f0 = c->fft_execs;
  cudaGraph_t graph = nullptr;
  if (cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    c->use_graphs = 0;
    return rk_step_compact<T, D>(c, tb, dt, theta);
  }
  const int r = rk_step_compact<T, D>(c, tb, dt, theta);
  const cudaError_t ce = cudaStreamEndCapture(c->stream, &graph);
  insb200_ctx::StepGraph e;
  if (r == 0 && ce == cudaSuccess && graph && cudaGraphInstantiate(&e.exec, graph, 0) == cudaSuccess) {
    cudaGraphDestroy(graph);
    e.key = key;
    e.post[0] = c->ubuf[0]; e.post[1] = c->ubuf[1]; e.post[2] = c->ubuf[2]; e.post[3] = c->preal; e.post[4] = c->preal2;
    e.launches = c->launches - l0; e.fft_execs = c->fft_execs - f0;
    INS_CUDA(c, cudaGraphLaunch(e.exec, c->stream));
    c->graphs.push_back(e);
    return 0;
  }
  // capture failed: nothing ran; restore the rotation, stop using graphs on this ctx and run the step eagerly
  if (graph) cudaGraphDestroy(graph);
  cudaGetLastError();
  for (int b = 0; b < 3; b++) c->ubuf[b] = pre[b];
  c->preal = pre[3]; c->preal2 = pre[4];
  c->launches = l0; c->fft_execs = f0;
  c->use_graphs = 0;
  return rk_step_compact<T, D>(c, tb, dt, theta);
}

template <typename T, int D>
static int reduce_compact(insb200_ctx *c, double *out4) {
  INS_TRY((halo_compact<T>(c, c->ubuf[0], D, 1, 0)));
  return reduce_view<T, D>(c, view_compact_c<T>(c, c->ubuf[0]), geo_compact(c), out4);
}

template <typename T, int D>
static int export_state(insb200_ctx *c, void *u) {
  INS_TRY((export_field<T, D>(c, u, c->ubuf[0], D)));
  return bc_fill<T, D>(c, u, D);  // final apply_bc_u! (:55): all ghosts valid on return
}

template <typename T, int D>
static int op_rk_steps(insb200_ctx *c, void *u, double *t, double dt, int nsteps, const Tableau &tb, double theta,
                       double *ke_hist, void *p_out) {
  const bool smag = !std::isnan(theta);
  INS_TRY(ensure_rk(c, tb, smag));
  INS_TRY((import_field<T, D>(c, c->ubuf[0], u, D)));
  double o[4];
  if (ke_hist) { INS_TRY((reduce_compact<T, D>(c, o))); ke_hist[0] = o[0]; }
  T tt = T(*t);
  for (int n = 0; n < nsteps; n++) {
    c->t_step = (double)tt;
    INS_TRY((rk_step_graphed<T, D>(c, tb, dt, theta)));
    tt = tt + T(tb.c[tb.s - 1]) * T(dt);  // t = tstart + c[i]*dt (:32)
    if (ke_hist) { INS_TRY((reduce_compact<T, D>(c, o))); ke_hist[n + 1] = o[0]; }
  }
  *t = (double)tt;
  INS_TRY((export_state<T, D>(c, u)));
  if (p_out) INS_TRY((export_p<T, D>(c, p_out)));
  return 0;
}

template <typename T, int D>
static int op_solve_host(insb200_ctx *c, void *u_host, double t0, double tend, double dt, double cfl, int n_adapt,
                         int max_steps, const Tableau &tb, double theta, double *ke_hist, int ke_cap, int *nsteps_out,
                         double *t_out) {
  const bool smag = !std::isnan(theta);
  INS_TRY(ensure_rk(c, tb, smag));
  const size_t bytes = (size_t)ghosted_elems(c) * D * sizeof(T);
  INS_TRY(dev_alloc(c, &c->hostio, bytes));
  INS_CUDA(c, cudaMemcpyAsync(c->hostio, u_host, bytes, cudaMemcpyHostToDevice, c->stream));
  INS_TRY((import_field<T, D>(c, c->ubuf[0], c->hostio, D)));
  double o[4];
  int nk = 0;
  if (ke_hist && nk < ke_cap) { INS_TRY((reduce_compact<T, D>(c, o))); ke_hist[nk++] = o[0]; }
  T t = T(t0);
  const T te = T(tend);
  int n = 0;
  if (dt > 0) {
    // fixed step (solver.jl:71-79): nstep = round((tend-tstart)/dt); dt = (tend-tstart)/nstep
    long long nstep = llround((double)((te - t) / T(dt)));
    if (nstep < 1) nstep = 1;
    const T d = (te - t) / T((double)nstep);
    if (max_steps > 0 && nstep > max_steps) nstep = max_steps;
    for (; n < nstep; n++) {
      c->t_step = (double)t;
      INS_TRY((rk_step_graphed<T, D>(c, tb, (double)d, theta)));
      t = t + T(tb.c[tb.s - 1]) * d;
      if (ke_hist && nk < ke_cap) { INS_TRY((reduce_compact<T, D>(c, o))); ke_hist[nk++] = o[0]; }
    }
  } else {
    // adaptive (solver.jl:53-69)
    if (n_adapt < 1) n_adapt = 1;
    T d = T(0);
    while (t < te) {
      if (n % n_adapt == 0) {
        INS_TRY((reduce_compact<T, D>(c, o)));
        d = T(cfl) * T(cfl_from_max<T>(c, o));
      }
      d = std::min(d, te - t);
      c->t_step = (double)t;
      INS_TRY((rk_step_compact<T, D>(c, tb, (double)d, theta)));
      t = t + T(tb.c[tb.s - 1]) * d;
      n++;
      if (ke_hist && nk < ke_cap) { INS_TRY((reduce_compact<T, D>(c, o))); ke_hist[nk++] = o[0]; }
      if (max_steps > 0 && n >= max_steps) break;
    }
  }
  INS_TRY((export_state<T, D>(c, c->hostio)));
  INS_CUDA(c, cudaMemcpyAsync(u_host, c->hostio, bytes, cudaMemcpyDeviceToHost, c->stream));
  INS_CUDA(c, cudaStreamSynchronize(c->stream));
  if (nsteps_out) *nsteps_out = n;
  if (t_out) *t_out = (double)t;
  return 0;
}

}  // namespace insb200

extern "C" int insb200_rk_step(insb200_ctx *c, void *u, double t, double dt, int s, const double *A, const double *cc,
                               double theta, void *p_out) {
  API_PROLOG(c);
  if (!u) INS_FAIL(c, "u is NULL");
  Tableau tb;
  INS_TRY(make_tableau(c, s, A, cc, tb));
  double tt = t;
  return DISPATCH(c, op_rk_steps, c, u, &tt, dt, 1, tb, theta, nullptr, p_out);
}

extern "C" int insb200_rk_steps(insb200_ctx *c, void *u, double *t, double dt, int nsteps, int s, const double *A,
                                const double *cc, double theta, double *ke_hist) {
  API_PROLOG(c);
  if (!u || !t) INS_FAIL(c, "NULL argument");
  if (nsteps < 0) INS_FAIL(c, "nsteps < 0");
  Tableau tb;
  INS_TRY(make_tableau(c, s, A, cc, tb));
  return DISPATCH(c, op_rk_steps, c, u, t, dt, nsteps, tb, theta, ke_hist, nullptr);
}

extern "C" int insb200_solve_unsteady_host(insb200_ctx *c, void *u_host, double t0, double tend, double dt, double cfl,
                                           int n_adapt_dt, int max_steps, int s, const double *A, const double *cc,
                                           double theta, double *ke_hist, int ke_cap, int *nsteps_out, double *t_out) {
  API_PROLOG(c);
  if (!u_host) INS_FAIL(c, "u_host is NULL");
  if (!(tend >= t0)) INS_FAIL(c, "tend < t0");
  Tableau tb;
  INS_TRY(make_tableau(c, s, A, cc, tb));
  return DISPATCH(c, op_solve_host, c, u_host, t0, tend, dt, cfl, n_adapt_dt, max_steps, tb, theta, ke_hist, ke_cap,
                  nsteps_out, t_out);
}

// =================================================================================================
// introspection
// =================================================================================================
extern "C" double insb200_model_bytes_per_cell_step(const insb200_ctx *c, int nstages) {
  if (!c) return 0;
  // RK44 = 3*4D + 6D + 4*[(D+1) + 4D + (2D+1)] = 46D + 8 words; generalised per stage:
  // stages with one tableau entry 4D, final stage (3 + (s-1)) D; + D per stage with a stored force.
  const int D = c->D, s = nstages;
  double words = (s - 1) * 4.0 * D + (3.0 + (s - 1)) * D + s * ((D + 1) + 4.0 * D + (2.0 * D + 1));
  if (c->force) words += (double)s * D;
  return words * (double)esize(c);
}
extern "C" long long insb200_launch_count(const insb200_ctx *c) { return c ? c->launches : 0; }
extern "C" long long insb200_fft_exec_count(const insb200_ctx *c) { return c ? c->fft_execs : 0; }
extern "C" long long insb200_workspace_bytes(const insb200_ctx *c) { return c ? c->ws_bytes : 0; }
extern "C" long long insb200_graph_replays(const insb200_ctx *c) { return c ? c->graph_replays : 0; }

static int drain_timers(insb200_ctx *c) {
  INS_CUDA(c, cudaStreamSynchronize(c->stream));
  for (auto &r : c->recs) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) { c->tm_ms[r.id] += ms; c->tm_cnt[r.id]++; }
    c->pool.push_back(r.a);
    c->pool.push_back(r.b);
  }
  c->recs.clear();
  return 0;
}
extern "C" int insb200_set_profiling(insb200_ctx *c, int enable) {
  API_PROLOG(c);
  INS_TRY(drain_timers(c));
  c->profiling = enable != 0;
  return 0;
}
extern "C" int insb200_reset_timers(insb200_ctx *c) {
  API_PROLOG(c);
  INS_TRY(drain_timers(c));
  for (int i = 0; i < TM_COUNT; i++) { c->tm_ms[i] = 0; c->tm_cnt[i] = 0; }
  return 0;
}
extern "C" int insb200_timers(insb200_ctx *c, const char **names, double *ms, long long *counts, int cap, int *n) {
  API_PROLOG(c);
  INS_TRY(drain_timers(c));
  int k = 0;
  for (int i = 0; i < TM_COUNT && k < cap; i++) {
    if (!c->tm_cnt[i]) continue;
    names[k] = kTimerNames[i]; ms[k] = c->tm_ms[i]; counts[k] = c->tm_cnt[i];
    k++;
  }
  if (n) *n = k;
  return 0;
}
extern "C" int insb200_set_variant(insb200_ctx *c, int variant) {
  if (!c) return 1;
  clear_graphs(c);
  c->variant = variant;
  return 0;
}
extern "C" int insb200_set_param(insb200_ctx *c, const char *name, double value) {
  if (!c) return 1;
  if (!name) INS_FAIL(c, "parameter name is NULL");
  const std::string k(name);
  const int v = (int)value;
  clear_graphs(c);  // a captured step holds the launch configuration of the parameters it was captured with
  if (k == "variant") c->variant = v;
  else if (k == "exact") c->exact = v != 0;   // 1: bit-exact kernels (reference operation order, no contraction); 0: fast kernels (fast.cu)
  else if (k == "fast_gen") c->fast_gen = v != 0;
  else if (k == "graphs") c->use_graphs = v;  // 1 on, 0 off, -1 automatic (small grids)
  else if (k == "tile_y") { if (v != 8 && v != 16) INS_FAIL(c, "tile_y must be 8 or 16"); c->tile_y = v; }
  else if (k == "zchunk") { if (v != 0 && v < 2) INS_FAIL(c, "zchunk must be 0 (automatic) or >= 2"); c->zchunk = v; }
  else if (k == "fuse_gradient") c->fuse_gradient = v != 0;
  else if (k == "zsolve") c->zsolve = v != 0;
  else if (k == "rows2d") c->rows2d = v != 0;
  else if (k == "fuse2d") c->fuse2d = v != 0;
  else if (k == "rows_chunk") { if (v != 0 && v < 2) INS_FAIL(c, "rows_chunk must be 0 (automatic) or >= 2"); c->rows_chunk = v; }
  else if (k == "stage_x2") {
#ifdef INSB200_EXPERIMENTAL
    if (v < 0 || v > 2) INS_FAIL(c, "stage_x2 must be 0, 1 or 2");
    c->stage_x2 = v;
#else
    if (v != 0) INS_FAIL(c, "stage_x2: the packed-FP32 stage kernel is a negative result kept behind -DINSB200_EXPERIMENTAL (DESIGN.md section 4)");
#endif
  }
  else if (k == "smag_v") { if (v != 1 && v != 2) INS_FAIL(c, "smag_v must be 1 or 2"); c->smag_v = v; }
  else if (k == "fft_pipe") {
    if (c->plans) INS_FAIL(c, "fft_pipe must be set before the first pressure solve");
    if (v < 1 || v > 16) INS_FAIL(c, "fft_pipe must be in [1, 16]");
    c->fft_pipe = v;
  }
  else if (k == "fft_pipe_min_ranks") c->fft_pipe_min_ranks = v;
  else if (k == "fft_chunk") {
    if (c->plans) INS_FAIL(c, "fft_chunk must be set before the first pressure solve");
    c->fft_chunk = v;
  }
  else if (k == "smag_ty") { if (v != 8 && v != 16) INS_FAIL(c, "smag_ty must be 8 or 16"); c->smag_ty = v; }
  else if (k == "smag_zchunk") { if (v != 0 && v < 2) INS_FAIL(c, "smag_zchunk must be 0 (automatic) or >= 2"); c->smag_zchunk = v; }
  else if (k == "p2p_halo") c->p2p_halo = v != 0;
  else if (k == "fast_barrier") c->fast_barrier = v != 0;
  else if (k == "transpose_ce") { c->transpose_ce = v != 0; c->transpose_mode_req = c->transpose_mode = c->transpose_ce; }  // alias of transpose_mode 0 / 1
  else if (k == "transpose_mode") {
    if (v < -1 || v > 2) INS_FAIL(c, "transpose_mode must be -1 (automatic), 0 (scatter kernels), 1 (pitched peer copies) or 2 (pack + contiguous peer copies)");
    if (v < 0) { c->transpose_mode_req = -1; return 0; }
    if (c->plans && c->nranks > 1 && v == 2 && !c->pstage) INS_FAIL(c, "transpose_mode 2 must be set before the first pressure solve");
    c->transpose_mode_req = c->transpose_mode = v;
    if (v < 2) c->transpose_ce = v;
  }
  else if (k == "scatter_blocks") { if (v < 1 || v > 65535) INS_FAIL(c, "scatter_blocks must be in [1, 65535]"); c->scatter_blocks = v; }
  else if (k == "p2p_fused") { if (v != 0) INS_FAIL(c, "p2p_fused: the fused peer transposes (k_zsolve_peer) are a negative result kept behind -DINSB200_EXPERIMENTAL (DESIGN.md)"); }
  else if (k == "p2p") {
    if (c->plans && c->nranks > 1) INS_FAIL(c, "p2p must be set before the first pressure solve");
    c->p2p_enable = v != 0;
  }
  else if (k == "zline") { if (v < -1 || v > 1) INS_FAIL(c, "zline must be -1 (automatic), 0 or 1"); c->zline_req = v; }
  else if (k == "zline_chunks") { if (v < 0) INS_FAIL(c, "zline_chunks must be >= 0"); c->zline_chunks = v; }
  else if (k == "zcols") { if (v != 0 && v != 4 && v != 8 && v != 16) INS_FAIL(c, "zcols must be 0, 4, 8 or 16"); c->zcols = v; }
  else if (k == "zradix16") c->zradix16 = v != 0;
  else if (k == "zthreads") { if (v < 64 || v > 512 || v % 32) INS_FAIL(c, "zthreads must be a multiple of 32 in [64, 512]"); c->zthreads = v; }
  else INS_FAIL(c, "unknown parameter: " + k);
  return 0;
}
