// fp32x2_probe.cu -- issue rate of scalar vs packed FP32 instructions on sm_100a (FADD/FMUL/FFMA vs FADD2/FMUL2/FFMA2).
// One block of 1024 threads per SM (8 warps per scheduler), 8 independent dependency chains per thread, so the
// pipe rate, not the latency, bounds the loop.  Prints warp-instructions per cycle per SM sub-partition.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o tools/fp32x2_probe tools/fp32x2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float lo(u64 v) { float a, b; asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); return a + b; }

template <int MODE>
__global__ void __launch_bounds__(1024, 1) probe(float *out, long long *cyc, int iters, float s) {
  float a[8];
  u64 p[8];
#pragma unroll
  for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 0.001f + i; p[i] = pk(a[i], a[i] + 0.5f); }
  const u64 ps = pk(s, s);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(s));
      if (MODE == 1) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(s));
      if (MODE == 2) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(a[i]) : "f"(s));
      if (MODE == 3) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(ps));
      if (MODE == 4) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(ps));
      if (MODE == 5) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(p[i]) : "l"(ps));
      if (MODE == 6) {  // alternating FADD / FMUL (the stencil mix)
        if (i & 1) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(s));
        else asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(s));
      }
      if (MODE == 7) {  // alternating FADD2 / FMUL2
        if (i & 1) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(ps));
        else asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(ps));
      }
      if (MODE == 8) {  // FADD2 with operands assembled from two scalar registers each time (pack cost)
        u64 q = pk(a[i], a[(i + 1) & 7]);
        asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(q));
      }
    }
  }
  const long long t1 = clock64();
  float r = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) r += a[i] + lo(p[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char *name, float *out, long long *cyc, int nsm) {
  const int iters = 4096;
  probe<MODE><<<nsm, 1024>>>(out, cyc, iters, 1.0000001f);
  probe<MODE><<<nsm, 1024>>>(out, cyc, iters, 1.0000001f);
  cudaDeviceSynchronize();
  long long h[256];
  cudaMemcpy(h, cyc, nsm * sizeof(long long), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < nsm; i++) avg += (double)h[i];
  avg /= nsm;
  const double winstr = (double)iters * 8 * 32;  // warp-instructions per SM (32 warps)
  printf("%-28s %10.0f cycles  %.3f warp-instr/clk/SM  %.3f per scheduler  (%s)\n", name, avg, winstr / avg, winstr / avg / 4,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  cudaDeviceProp pr;
  cudaGetDeviceProperties(&pr, 0);
  const int nsm = pr.multiProcessorCount;
  float *out; long long *cyc;
  cudaMalloc(&out, (size_t)nsm * 1024 * sizeof(float));
  cudaMalloc(&cyc, 256 * sizeof(long long));
  printf("%s, %d SMs\n", pr.name, nsm);
  run<0>("FADD", out, cyc, nsm);
  run<1>("FMUL", out, cyc, nsm);
  run<2>("FFMA", out, cyc, nsm);
  run<3>("FADD2 (add.f32x2)", out, cyc, nsm);
  run<4>("FMUL2 (mul.f32x2)", out, cyc, nsm);
  run<5>("FFMA2 (fma.f32x2)", out, cyc, nsm);
  run<6>("FADD+FMUL alternating", out, cyc, nsm);
  run<7>("FADD2+FMUL2 alternating", out, cyc, nsm);
  run<8>("FADD2 with per-use packing", out, cyc, nsm);
  return 0;
}
