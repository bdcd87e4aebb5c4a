// ctx.cuh -- context, views and error plumbing of libinsb200 (sm_100a).
//
// Data layout in HBM
//   reference layout ("ghosted"): what the Julia caller owns, column-major (N1,..,ND[,D]),
//     N_a = nloc_a + 2 (src/grid.jl:121).  Row pitch (n+2)*sizeof(T) is NOT a multiple of 16 B in
//     Float32, so rows are neither float4- nor TMA-addressable.
//   compact layout (internal): interior cells only, x/y periodic by index wrap, the last axis
//     either wrapped (nranks == 1) or padded with G ghost planes filled over NCCL (nranks > 1).
//     Row pitch n*sizeof(T) is a multiple of 16 B for every admissible n (n even and, for the
//     vector paths, n % 4 == 0), so rows can be moved with 128-bit loads / bulk copies.
//   padded layout (internal, 3-D): (PX, PY, PZ) with REAL ghost cells -- gm = 2 layers on the minus
//     side and gp = 3 on the plus side in y and z, 64 B (left) + 64 B (right) in x -- so PX*sizeof(T)
//     is a multiple of 128 B and the interior origin is 64-B aligned.  Any tile + halo is one
//     in-bounds TMA box (tma_stage.cuh); the ghosts are refreshed by k_fill_ghosts (x, y, and z on one
//     rank) or by NCCL plane exchange (z, nranks > 1).  The one-thread-per-cell kernels address the
//     same buffers through (origin, strides) and wrap indices instead of reading ghosts.
//   The RK state (ustart, stage buffers, k_i, s, FFT buffers) lives in the compact (2-D) or padded
//   (3-D) layout; the reference layout is touched once on entry and once on exit of a stepping call.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cufft.h>
#include <nccl.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/insb200.h"

namespace insb200 {

typedef long long i64;

// A field view: pointer to cell (0,0,0) of component 0 of the INTERIOR, strides in elements.
template <typename T>
struct View {
  T *p;
  i64 sy, sz, sc;  // x stride is 1
};

// Local geometry handed to kernels.
struct Geo {
  int n[3];     // local interior cells per axis (n[2] = 1 in 2-D)
  int wrap[3];  // 1: periodic wrap by index (compact layout); 0: ghost cells are addressable
};

enum TimerId {
  TM_IMPORT = 0, TM_EXPORT, TM_BC, TM_MOMENTUM, TM_STAGE, TM_SMAGTENSOR, TM_DIVTENSOR, TM_DIVERGENCE,
  TM_FFT_FWD, TM_FFT_INV, TM_POISSON_SCALE, TM_GRADIENT, TM_REDUCE, TM_HALO, TM_PACK, TM_A2A,
  TM_MISC, TM_STAGE_GEN, TM_SMAG_FUSED, TM_GHOSTS, TM_SCATTER, TM_ZSUMS, TM_COUNT
};
static const char *const kTimerNames[TM_COUNT] = {
    "import_ghosted", "export_ghosted", "apply_bc", "momentum", "momentum_stage", "smagtensor",
    "divoftensor", "divergence", "cufft_forward", "cufft_inverse", "poisson_scale", "pressure_gradient",
    "reduce", "halo_exchange", "fft_pack_unpack", "fft_alltoall", "misc", "momentum_stage_last", "smagorinsky_fused",
    "fill_ghosts", "fft_transpose", "poisson_forward"};

struct TimerRec {
  cudaEvent_t a, b;
  int id;
};

}  // namespace insb200

struct insb200_ctx {
  int D = 0;
  int dtype = 0;           // 0 f32, 1 f64
  int nglob[3] = {1, 1, 1};
  int nloc[3] = {1, 1, 1};
  int last = 0;            // index of the decomposed (last) axis
  int off_last = 0;        // first global plane of this rank on the last axis
  double h[3] = {1, 1, 1};
  double Re = 1;
  int device = 0;
  int rank = 0, nranks = 1;
  int exact = 0;           // param "exact": 1 = the bit-exact kernels of insb200.cu everywhere; 0 = fast.cu's contracted kernels where they exist
  int fast_gen = 0;        // param "fast_gen": without closure, tableau rows that read earlier k_j (RK44's last row): 1 = k_stage3d_fast<GEN> (one block per SM, 128 registers), 0 = the exact TMA kernel (measured faster)
  int variant = 2;         // 0: one thread per cell, 1: LDG-tiled stage kernel, 2: TMA plane pipeline (3-D)
  int tile_y = 16;         // tiled stage kernel: 32 x tile_y cells per block
  int fft_pipe = 4;        // param: nranks > 1, peer transposes: the slab is transformed in this many z-chunks and the scatter of
                           // one chunk overlaps the cuFFT of the next on a second stream (1 = no pipelining)
  int fft_pipe_min_ranks = 2;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t pipe_ev[16] = {nullptr}, pack_ev[16] = {nullptr}, pipe_sync = nullptr;
  int fft_chunk = 0;       // param: planes per in-plane cuFFT exec (0 = the whole slab); small chunks keep the x->y intermediate in L2
  int fft_batch = 0;       // planes per exec of plan_fwd_in / plan_inv_in as planned
  long long fft_cdist = 0; // complex words between consecutive planes of the in-slab spectrum
  int stage_x2 = 0;        // F32 stage kernel: 0 = k_stage3d_tma (default: fastest overall); two cells per thread on packed FP32 (k_stage3d_tma_x2): 1 = exact, 2 = FFMA2-contracted
  int smag_v = 2;          // closure kernel: 1 = shared-memory stencil (k_smag3d_tma), 2 = register-carried (k_smag3d_v2)
  int smag_ty = 16;        // tile height of k_smag3d_v2: 16 (one 640-thread block per SM) or 8 (two 384-thread blocks)
  int smag_zchunk = 0;     // planes marched per block of k_smag3d_v2 (0 = automatic, auto_zchunk)
  int zchunk = 0;          // stage kernels: planes marched per block (0 = automatic for the TMA kernels, 64 for the LDG-tiled ones)
  int nsm = 148;           // SMs of the device
  int fuse2d = 1;          // 2-D, one rank: fused warp-marching stage kernel (stage2d.cuh); 0 = stage + divergence + gradient kernels
  int rows_chunk = 0;      // rows marched per block of k_stage2d_warp (0 = automatic)
  int rows2d = 1;          // 2-D: divergence / gradient kernels march several rows per thread (0 = one cell per thread)
  int padded = 0;          // 1: padded layout with real ghost cells (3-D)
  int org[3] = {0, 0, 0};  // interior origin inside the padded box (elements)
  int ext[3] = {1, 1, 1};  // padded extents PX, PY, PZ
  int gm = 0, gp = 0;      // ghost layers kept valid on the minus / plus side (y, z; x has more room)
  int fuse_gradient = 1;   // apply the projection of stage i in the loader of stage i+1
  int zradix16 = 1;        // k_zsolve, Float32: radix-16 passes where they save a trip through shared memory
  int zcols = 0, zthreads = 256;  // k_zsolve tuning: columns per tile (0 = largest that fits 64 KB), threads per block
  int zsolve = 1;          // 3-D: own fused last-axis transform + symbol kernel instead of cuFFT's z passes
  cudaStream_t stream = nullptr;
  cudaStream_t own_stream = nullptr;
  ncclComm_t comm = nullptr;
  std::string err;

  // compact-layout geometry
  int G = 0;               // ghost planes on the last axis of compact vector buffers (0 if nranks==1)
  insb200::i64 csy = 0, csz = 0;      // compact strides
  insb200::i64 cplane = 0;            // elements of one last-axis plane
  insb200::i64 ccomp = 0;             // elements per component incl. ghost planes
  insb200::i64 cint = 0;              // interior elements per component
  insb200::i64 corg = 0;              // offset of interior cell (0,0,0) from the buffer start

  // workspaces (device), element type = dtype
  void *ubuf[3] = {nullptr, nullptr, nullptr};          // rotating state buffers (compact vector)
  void *kbuf[INSB200_MAX_STAGES] = {nullptr};           // stage derivatives
  void *force = nullptr;                                // compact steady body force (nullable)
  insb200_force_fn force_cb = nullptr;                  // unsteady body force (operators.jl:870-877): the caller evaluates it per stage time
  void *force_user = nullptr;
  double t_step = 0;                                    // start time of the RK step being taken (the stage times of the callback)
  void *force_g = nullptr;                              // the same force in the reference layout (ghost values of F)
  void *sigma = nullptr;                                // compact symmetric tensor, D(D+1)/2 comps
  void *sbuf = nullptr;                                 // closure term s = div sigma (vector, TMA path)
  std::map<std::tuple<const void *, int, int>, CUtensorMap> tmaps;  // (buffer, ncomp, box rows) -> descriptor
  void *preal = nullptr;                                // compact scalar: FFT real buffer (+ ghost planes)
  void *preal2 = nullptr;                               // second scalar buffer (multiplier of the previous stage)
  void *pcplx = nullptr;                                // complex spectrum
  void *pcplx2 = nullptr;                               // transposed spectrum / a2a buffers (nranks>1)
  void *ahat[3] = {nullptr, nullptr, nullptr};          // Laplacian symbol per axis (T)
  void *peerS[16] = {nullptr}, *peerB[16] = {nullptr};  // pcplx / pcplx2 of every rank mapped through CUDA IPC
  struct NeighbourBuf { void *prev = nullptr, *next = nullptr; };  // ring neighbours' mappings of a local buffer
  std::map<void *, NeighbourBuf> halo_peers;             // registered by register_halo_buffer (halo planes by peer copies)
  int p2p_halo = 1;                                     // param "p2p_halo": 0 = halo planes over ncclSend/ncclRecv
  std::map<int, std::vector<CUtensorMap>> peer_tmaps;   // k_zsolve_peer descriptors of the peers' spectra, key = COLS*64 + CX
  int p2p_fused = 0;                                    // param "p2p_fused": k_zsolve_peer gathers / scatters the spectrum itself (3-D, needs p2p)
  int nxp = 0;                                          // row pitch (complex words) of the in-slab spectrum, >= nxh
  int p2p = 0;                                          // 1: peer mappings established, transposes go over peer stores
  int p2p_enable = 1;                                   // param "p2p": 0 forces the NCCL send/recv transpose
  void *barbuf = nullptr;                               // one int: all-reduced as the barrier after a scatter (NCCL fallback)
  void *flagbuf = nullptr, *peerF[16] = {nullptr};      // barrier flag words of this rank / of every rank (CUDA IPC): k_peer_barrier
  unsigned bar_epoch = 0;                               // barriers issued so far (same sequence on every rank)
  int transpose_mode_req = -1;                          // as set by the caller; -1 = automatic: by the size of the row pieces (ensure_fft)
  int transpose_mode = 2;                               // param "transpose_mode" (pipelined path): 2 = local pack -> CONTIGUOUS peer copies on the copy engines -> local unpack; 1 = pitched peer copies (transpose_ce); 0 = scatter kernels
  void *pstage = nullptr, *peerP[16] = {nullptr};       // staging buffer of the packed transposes (own / every rank's, CUDA IPC)
  int transpose_ce = 1;                                 // param "transpose_ce": 1 = transposes as pitched peer copies on the copy engines, 0 = scatter kernels
  std::vector<cudaStream_t> ce_streams;                 // one stream per destination slot of the copy-engine transposes
  std::vector<cudaEvent_t> ce_events;                   // [chunk][destination slot]
  int scatter_blocks = 128;                             // param "scatter_blocks": blocks of the persistent transpose kernels (comm.cuh)
  int fast_barrier = 1;                                 // param "fast_barrier": 0 = ncclAllReduce barriers
  std::map<const void *, size_t> smem_attr;             // dynamic shared memory opted into, per kernel: a context belongs to ONE device, a process-wide flag would not do
  int zline_req = -1;                                   // param "zline": last-axis solve of a decomposed grid without transposes (zline.cuh); -1 = automatic (on when the peers are mapped), 0 off, 1 on (3-D: also on one rank)
  int zline_chunks = 0;                                 // param "zline_chunks": sub-chunks per slab (0 = automatic)
  insb200::i64 cplx_slots = 0;                          // complex words of pcplx / pcplx2 on a decomposed grid
  void *zl_x = nullptr;                                 // exchange areas of the line solve on ONE rank (no peer buffer there)
  unsigned zl_count = 0;                                // line solves so far: the two exchange areas alternate
  void *ztw = nullptr;                                  // twiddles exp(-2 pi i t / n_last) of k_zsolve (complex T)
  void *fwork = nullptr;                                // compact vector scratch (insb200_pressure, API ops)
  void *hostio = nullptr;                               // reference-layout device image for *_host calls
  void *red = nullptr;                                  // reduction partials (double)
  double *red_host = nullptr;                           // pinned
  void *cufft_work = nullptr;
  size_t cufft_work_bytes = 0;
  insb200::i64 ws_bytes = 0;
  bool rk_alloc = false;
  int rk_nk = 0;

  // temperature equation (temperature.cu): coefficients of temperature_equation (setup.jl:56-86) and the
  // reference-layout RK cache (ustart, ku, tempstart, ktemp, diff, p) of the temperature stepper
  struct TempEq { bool set = false; double a1 = 0, a2 = 0, a3 = 0, a4 = 0, gamma = 1; int dodissipation = 1; int gdir = 1; } tq;
  void *tws = nullptr;
  size_t tws_bytes = 0;

  // CUDA graphs of whole RK steps (insb200.cu: rk_step_graphed): one instantiated graph per buffer permutation
  struct StepGraph {
    std::vector<unsigned long long> key;   // buffers, stream, dt, theta, tableau
    cudaGraphExec_t exec = nullptr;
    void *post[5] = {nullptr};             // ubuf[0..2], preal, preal2 after the step (the step rotates its buffers)
    long long launches = 0, fft_execs = 0; // per replay
  };
  std::vector<StepGraph> graphs;
  bool graph_warm = false;                              // one eager step has run since the last (re)configuration
  int use_graphs = -1;                                  // param "graphs": 1 on, 0 off, -1 automatic (slabs of <= 2^22 cells)
  long long graph_replays = 0;
  int graph_misses = 0;                                 // captures since the last replay

  // FFT plans
  cufftHandle plan_fwd = 0, plan_inv = 0;               // nranks==1: D-dim R2C / C2R
  cufftHandle plan_fwd_in = 0, plan_inv_in = 0;         // nranks>1: in-slab transforms
  cufftHandle plan_last = 0;                            // nranks>1: C2C along the last axis
  bool plans = false;
  int nxh = 0;                                          // n[0]/2+1
  int kxc = 0, kx0 = 0, kxl = 0;                        // kx chunk size, my first kx, my kx count

  // accounting
  long long launches = 0, fft_execs = 0;
  bool profiling = false;
  std::vector<insb200::TimerRec> recs;
  std::vector<cudaEvent_t> pool;
  double tm_ms[insb200::TM_COUNT] = {0};
  long long tm_cnt[insb200::TM_COUNT] = {0};
};

namespace insb200 {

extern thread_local std::string g_create_error;

inline int fail(insb200_ctx *c, const char *file, int line, const std::string &msg) {
  char buf[96];
  const char *base = strrchr(file, '/');
  snprintf(buf, sizeof buf, " [%s:%d]", base ? base + 1 : file, line);
  std::string m = msg + buf;
  if (c) c->err = m; else g_create_error = m;
  return 1;
}

#define INS_FAIL(ctx, msg) return ::insb200::fail((ctx), __FILE__, __LINE__, (msg))
#define INS_CUDA(ctx, call)                                                                     \
  do {                                                                                          \
    cudaError_t e_ = (call);                                                                    \
    if (e_ != cudaSuccess)                                                                      \
      return ::insb200::fail((ctx), __FILE__, __LINE__, std::string("CUDA: ") + cudaGetErrorString(e_) + " in " #call); \
  } while (0)
#define INS_CUFFT(ctx, call)                                                                    \
  do {                                                                                          \
    cufftResult r_ = (call);                                                                    \
    if (r_ != CUFFT_SUCCESS)                                                                    \
      return ::insb200::fail((ctx), __FILE__, __LINE__, std::string("cuFFT error ") + std::to_string((int)r_) + " in " #call); \
  } while (0)
#define INS_NCCL(ctx, call)                                                                     \
  do {                                                                                          \
    ncclResult_t r_ = (call);                                                                   \
    if (r_ != ncclSuccess)                                                                      \
      return ::insb200::fail((ctx), __FILE__, __LINE__, std::string("NCCL: ") + ncclGetErrorString(r_) + " in " #call); \
  } while (0)
// opt a kernel into `bytes` of dynamic shared memory once per context (and again if a launch needs more)
#define INS_DYN_SMEM(c, kern, bytes)                                                                                  \
  do {                                                                                                                \
    size_t &cur_ = (c)->smem_attr[reinterpret_cast<const void *>(kern)];                                              \
    if ((size_t)(bytes) > cur_) {                                                                                     \
      INS_CUDA((c), cudaFuncSetAttribute((kern), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));         \
      cur_ = (size_t)(bytes);                                                                                         \
    }                                                                                                                 \
  } while (0)

#ifndef CHECK_LAUNCH
#define CHECK_LAUNCH(c)                      \
  do {                                       \
    (c)->launches++;                         \
    INS_CUDA((c), cudaGetLastError());       \
  } while (0)
#endif
#define INS_TRY(call)            \
  do {                           \
    int r__ = (call);            \
    if (r__) return r__;         \
  } while (0)

// scoped device timer: records an event pair around the launches of one logical kernel
struct ScopedTimer {
  insb200_ctx *c;
  int id;
  cudaEvent_t a = nullptr, b = nullptr;
  cudaStream_t st;
  ScopedTimer(insb200_ctx *c_, int id_, cudaStream_t st_ = nullptr) : c(c_), id(id_), st(st_ ? st_ : c_->stream) {
    if (!c->profiling) return;
    a = take();
    b = take();
    cudaEventRecord(a, st);
  }
  ~ScopedTimer() {
    if (!c->profiling) return;
    cudaEventRecord(b, st);
    c->recs.push_back({a, b, id});
  }
  cudaEvent_t take() {
    if (!c->pool.empty()) {
      cudaEvent_t e = c->pool.back();
      c->pool.pop_back();
      return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
  }
};

inline size_t esize(const insb200_ctx *c) { return c->dtype == INSB200_F64 ? 8 : 4; }

}  // namespace insb200
