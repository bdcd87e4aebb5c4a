// tiled.cuh -- shared-memory tiled variants of the hot kernels (compact layout).
#pragma once
#include "stencil.cuh"

namespace insb200 {

template <typename T, int D>
static bool tiled_stage_supported(const insb200_ctx *, const Geo &) { return false; }

template <typename T, int D>
static int launch_stage_tiled(insb200_ctx *c, const StageArgs<T> &, Geo, T, Rh<T>, bool) {
  INS_FAIL(c, "tiled stage kernel not built");
}

}  // namespace insb200
