// p2p_probe.cu -- how fast can one B200 move a strided half-spectrum into its NVLink peer?
// Single process, two devices, cudaDeviceEnablePeerAccess.  Build (here) and run (2-GPU box):
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/p2p_probe tools/p2p_probe.cu
//   gpurun --gpus 2 -- ./tools/p2p_probe
// Variants: SM stores of 8 / 16 bytes per thread (1 or 4 in flight), SM loads (pull), bulk-async
// shared->remote-global stores, copy-engine contiguous and pitched copies; one direction and both at once.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

template <typename V, int ILP>
__global__ void k_push(V *__restrict__ dst, const V *__restrict__ src, size_t n) {
  size_t i = ((size_t)blockIdx.x * blockDim.x) * ILP + threadIdx.x;
  V v[ILP];
#pragma unroll
  for (int j = 0; j < ILP; j++) if (i + (size_t)j * blockDim.x < n) v[j] = src[i + (size_t)j * blockDim.x];
#pragma unroll
  for (int j = 0; j < ILP; j++) if (i + (size_t)j * blockDim.x < n) dst[i + (size_t)j * blockDim.x] = v[j];
}

// strided rows like the spectrum transpose: src row pitch ps, dst row pitch w (elements of 8 bytes)
__global__ void k_push_rows(float2 *__restrict__ dst, const float2 *__restrict__ src, int w, int ps, size_t rows) {
  const int i = blockIdx.y * blockDim.x + threadIdx.x;
  const size_t r = (size_t)blockIdx.x * blockDim.y + threadIdx.y;
  if (i < w && r < rows) dst[r * w + i] = src[r * ps + i];
}

// persistent grid-stride version (148*k CTAs), 16 B per access
__global__ void k_push_persist(uint4 *__restrict__ dst, const uint4 *__restrict__ src, size_t n) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    uint4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
    dst[i] = a; dst[i + stride] = b; dst[i + 2 * stride] = c; dst[i + 3 * stride] = d;
  }
  for (; i < n; i += stride) dst[i] = src[i];
}

// bulk async store: stage CH bytes in shared memory (plain loads), then one cp.async.bulk shared -> global
template <int CH>
__global__ void k_push_bulk(char *__restrict__ dst, const char *__restrict__ src, size_t nbytes) {
  extern __shared__ __align__(128) char sm[];
  const size_t nchunks = nbytes / CH;
  int buf = 0;
  for (size_t ch = blockIdx.x; ch < nchunks; ch += gridDim.x, buf ^= 1) {
    char *s = sm + buf * CH;
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");  // the buffer used 2 chunks ago is free
    __syncthreads();
    const uint4 *g = reinterpret_cast<const uint4 *>(src + ch * CH);
    for (int k = threadIdx.x; k < CH / 16; k += blockDim.x) reinterpret_cast<uint4 *>(s)[k] = g[k];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + ch * CH),
                   "r"((uint32_t)__cvta_generic_to_shared(s)), "r"(CH)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

struct Dev {
  int id;
  cudaStream_t st;
  cudaEvent_t a, b;
  char *src, *dst;  // src: local data; dst: landing buffer for the peer's pushes
};

int main() {
  int nd = 0;
  CK(cudaGetDeviceCount(&nd));
  if (nd < 2) { printf("need 2 devices\n"); return 0; }
  const size_t bytes = (size_t)129 * 8 * 262144;  // 270 MB: half of a 257 x 512 x 512 complex64 spectrum
  Dev d[2];
  for (int i = 0; i < 2; i++) {
    d[i].id = i;
    CK(cudaSetDevice(i));
    int can = 0;
    CK(cudaDeviceCanAccessPeer(&can, i, 1 - i));
    printf("device %d can access peer %d: %d\n", i, 1 - i, can);
    if (!can) return 0;
    CK(cudaDeviceEnablePeerAccess(1 - i, 0));
    CK(cudaStreamCreate(&d[i].st));
    CK(cudaEventCreate(&d[i].a));
    CK(cudaEventCreate(&d[i].b));
    CK(cudaMalloc(&d[i].src, 2 * bytes + 4096));
    CK(cudaMalloc(&d[i].dst, bytes + 4096));
    CK(cudaMemset(d[i].src, 1, 2 * bytes));
    CK(cudaMemset(d[i].dst, 0, bytes));
  }
  const size_t rows = 262144;
  const int w = 129, ps = 257;
  const char *names[] = {"sm store  8B x1", "sm store 16B x1", "sm store 16B x4", "sm store 16B persistent 148x4x512",
                         "sm store rows 8B (w=129, pitch 257)", "bulk-async store 16KB chunks, 148x2 CTAs",
                         "bulk-async store 32KB chunks, 148x2 CTAs", "copy engine contiguous", "copy engine pitched (1032 B rows)",
                         "sm load (pull) 16B x4", "local copy 16B x4 (HBM reference)"};
  const int nvar = 11;
  for (int var = 0; var < nvar; var++) {
    for (int bidir = 0; bidir < 2; bidir++) {
      float best[2] = {1e30f, 1e30f};
      for (int rep = 0; rep < 4; rep++) {
        for (int i = 0; i <= bidir; i++) {
          CK(cudaSetDevice(i));
          Dev &me = d[i], &pe = d[1 - i];
          CK(cudaEventRecord(me.a, me.st));
          const size_t n8 = bytes / 8, n16 = bytes / 16;
          switch (var) {
            case 0: k_push<float2, 1><<<(unsigned)((n8 + 255) / 256), 256, 0, me.st>>>((float2 *)pe.dst, (const float2 *)me.src, n8); break;
            case 1: k_push<uint4, 1><<<(unsigned)((n16 + 255) / 256), 256, 0, me.st>>>((uint4 *)pe.dst, (const uint4 *)me.src, n16); break;
            case 2: k_push<uint4, 4><<<(unsigned)((n16 + 1023) / 1024), 256, 0, me.st>>>((uint4 *)pe.dst, (const uint4 *)me.src, n16); break;
            case 3: k_push_persist<<<148 * 4, 512, 0, me.st>>>((uint4 *)pe.dst, (const uint4 *)me.src, n16); break;
            case 4: k_push_rows<<<dim3((unsigned)(rows / 4), (w + 63) / 64), dim3(64, 4), 0, me.st>>>((float2 *)pe.dst, (const float2 *)me.src, w, ps, rows); break;
            case 5: {
              CK(cudaFuncSetAttribute(k_push_bulk<16384>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768));
              k_push_bulk<16384><<<148 * 2, 256, 32768, me.st>>>(pe.dst, me.src, bytes);
              break;
            }
            case 6: {
              CK(cudaFuncSetAttribute(k_push_bulk<32768>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
              k_push_bulk<32768><<<148 * 2, 256, 65536, me.st>>>(pe.dst, me.src, bytes);
              break;
            }
            case 7: CK(cudaMemcpyPeerAsync(pe.dst, pe.id, me.src, me.id, bytes, me.st)); break;
            case 8: CK(cudaMemcpy2DAsync(pe.dst, (size_t)w * 8, me.src, (size_t)ps * 8, (size_t)w * 8, rows, cudaMemcpyDeviceToDevice, me.st)); break;
            case 9: k_push<uint4, 4><<<(unsigned)((n16 + 1023) / 1024), 256, 0, me.st>>>((uint4 *)me.dst, (const uint4 *)pe.src, n16); break;
            case 10: k_push<uint4, 4><<<(unsigned)((n16 + 1023) / 1024), 256, 0, me.st>>>((uint4 *)me.dst, (const uint4 *)me.src, n16); break;
          }
          CK(cudaGetLastError());
          CK(cudaEventRecord(me.b, me.st));
        }
        for (int i = 0; i <= bidir; i++) {
          CK(cudaSetDevice(i));
          CK(cudaStreamSynchronize(d[i].st));
          float ms = 0;
          CK(cudaEventElapsedTime(&ms, d[i].a, d[i].b));
          if (rep > 0 && ms < best[i]) best[i] = ms;
        }
      }
      if (bidir) printf("%-46s both dirs : %7.3f / %7.3f ms  -> %7.1f / %7.1f GB/s per direction\n", names[var], best[0], best[1], bytes / best[0] * 1e-6, bytes / best[1] * 1e-6);
      else printf("%-46s one dir   : %7.3f ms            -> %7.1f GB/s\n", names[var], best[0], bytes / best[0] * 1e-6);
    }
  }
  return 0;
}
