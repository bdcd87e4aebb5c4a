// fast.cuh -- host interface between insb200.cu (compiled with -fmad=false: the bit-exact kernels) and fast.cu
// (compiled with FMA contraction: the default kernels of the 3-D step).
#pragma once
#include "tiled.cuh"
#include "tma_stage.cuh"
#include "zline.cuh"

namespace insb200 {

// insb200.cu
int get_tmap(insb200_ctx *c, const void *base, int nc, int box_x, int box_y, CUtensorMap *out);
int auto_zchunk(const insb200_ctx *c, int user, int tiles, int blocks_per_sm, int extra);

// fast.cu: one RK stage up to the Poisson right-hand side with k_stage3d_fast (fast_stage.cuh).
// theta = NaN: no closure.  ubase / pbase: START of the padded velocity / multiplier buffers (pbase == nullptr:
// first stage, no projection on load).  The caller checks stage_fast_supported() and that the tableau row reads at
// most 3 earlier k_j (a.nk <= 3); everything else takes the exact kernels of insb200.cu.
template <typename T>
int launch_stage_fast(insb200_ctx *c, const TmaStageArgs<T> &a, Geo g, double theta, const void *ubase, const void *pbase);
bool stage_fast_supported(const insb200_ctx *c, bool smag);

// fast.cu: k_zsolve (zsolve.cuh) built with FMA contraction
struct ZRadix;
template <typename T>
int launch_zsolve_fast(insb200_ctx *c, void *S, i64 ncols, int nxw, int kx0, T invN, const ZRadix &rad);

// fast.cu: the transpose-free last-axis solve of a slab-decomposed spectrum (zline.cuh).  launch_zline_forward = pass A (the
// chunk sums go to this rank's exchange area X.p[g.rank]); the caller then runs the peer barrier; launch_zline_backward =
// the constant mode's column and pass B (they read the other ranks' exchange areas).
template <typename T>
int launch_zline_forward(insb200_ctx *c, const ZlGeo &g, const PeerPtrs &X, void *S);
template <typename T>
int launch_zline_backward(insb200_ctx *c, const ZlGeo &g, const PeerPtrs &X, void *S);

}  // namespace insb200
