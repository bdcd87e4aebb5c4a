// zsolve_launch.cuh -- host launch of k_zsolve (zsolve.cuh), shared by insb200.cu (TAG 0: exact arithmetic) and fast.cu
// (TAG 1: FMA contraction).
#pragma once
#include <algorithm>

#include "zsolve.cuh"

namespace insb200 {

typedef CUresult (*PFN_tmap_encode)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <typename T, int COLS, int TAG>
static int launch_zsolve_i(insb200_ctx *c, void *S, i64 ncols, int nxw, int kx0, T invN, const ZRadix &rad) {
  const int nz = c->nglob[2];
  const size_t smem = 128 + (size_t)(nz * COLS + nz) * sizeof(Cx<T>) + (size_t)nz * sizeof(T) + 16;
  auto kern = k_zsolve<T, COLS, TAG>;
  // 2-D descriptor of the spectrum seen as [nz][2*ncols] words of T, box = (2*COLS) x min(nz, 256)
  CUtensorMap tm;
  {
    auto key = std::make_tuple((const void *)S, -COLS, (int)(ncols & 0x7fffffff));
    auto it = c->tmaps.find(key);
    if (it != c->tmaps.end()) tm = it->second;
    else {
      void *fn = nullptr;
      cudaDriverEntryPointQueryResult qr;
      INS_CUDA(c, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
      if (!fn || qr != cudaDriverEntryPointSuccess) INS_FAIL(c, "cuTensorMapEncodeTiled is not available from the driver");
      cuuint64_t dims[2] = {(cuuint64_t)(2 * ncols), (cuuint64_t)nz};
      cuuint64_t strides[1] = {(cuuint64_t)(2 * ncols) * sizeof(T)};
      cuuint32_t box[2] = {(cuuint32_t)(2 * COLS), (cuuint32_t)std::min(nz, 256)};
      cuuint32_t estr[2] = {1, 1};
      CUresult r = ((PFN_tmap_encode)fn)(&tm, sizeof(T) == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, S,
                                         dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) INS_FAIL(c, "cuTensorMapEncodeTiled (spectrum) failed with CUresult " + std::to_string((int)r));
      c->tmaps[key] = tm;
    }
  }
  INS_DYN_SMEM(c, kern, smem);
  const unsigned nb = (unsigned)((ncols + COLS - 1) / COLS);
  kern<<<nb, c->zthreads, smem, c->stream>>>(tm, (const Cx<T> *)c->ztw, (const T *)c->ahat[0], (const T *)c->ahat[1],
                                     (const T *)c->ahat[2], nz, ncols, nxw, kx0, T(-1) / invN, rad);
  CHECK_LAUNCH(c);
  return 0;
}

// tile width: the largest number of columns whose [nz][COLS] tile fits 64 KB (two blocks per SM), unless set_param("zcols")
template <typename T, int TAG>
static int launch_zsolve_cols(insb200_ctx *c, void *S, i64 ncols, int nxw, int kx0, T invN, const ZRadix &rad) {
  const size_t col_bytes = (size_t)c->nglob[2] * sizeof(Cx<T>);
  if (c->zcols == 4) return launch_zsolve_i<T, 4, TAG>(c, S, ncols, nxw, kx0, invN, rad);
  if (c->zcols == 16 && 16 * col_bytes <= 131072) return launch_zsolve_i<T, 16, TAG>(c, S, ncols, nxw, kx0, invN, rad);  // explicit request only
  if (c->zcols == 8 && 8 * col_bytes <= 131072) return launch_zsolve_i<T, 8, TAG>(c, S, ncols, nxw, kx0, invN, rad);
  if (16 * col_bytes <= 65536) return launch_zsolve_i<T, 16, TAG>(c, S, ncols, nxw, kx0, invN, rad);
  if (8 * col_bytes <= 65536) return launch_zsolve_i<T, 8, TAG>(c, S, ncols, nxw, kx0, invN, rad);
  return launch_zsolve_i<T, 4, TAG>(c, S, ncols, nxw, kx0, invN, rad);
}

}  // namespace insb200
