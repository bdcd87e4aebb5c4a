// p2p_ipc_probe.cu -- two PROCESSES, one GPU each, buffers shared through CUDA IPC: times the spectrum
// scatter kernel of the distributed FFT exactly as the library runs it (local + remote halves).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/p2p_ipc_probe tools/p2p_ipc_probe.cu
#include <cuda_runtime.h>
#include <unistd.h>
#include <sys/wait.h>
#include <cstdint>
#include <cstdio>
#include <cstring>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("[%d] CUDA error %s at line %d\n", me, cudaGetErrorString(e), __LINE__); _exit(1); } } while (0)
typedef long long i64;
struct PeerPtrs { void *p[16]; };

template <typename C>
__global__ void k_scatter(PeerPtrs W, const C *__restrict__ S, int nxh, i64 rows, int kxc, int rank, int only) {
  const int i = blockIdx.y * blockDim.x + threadIdx.x;
  const i64 r = (i64)blockIdx.x * blockDim.y + threadIdx.y;
  if (i >= nxh || r >= rows) return;
  const int q = i / kxc;
  if (only >= 0 && q != only) return;
  const int kq = i - q * kxc;
  const int w = min(kxc, nxh - q * kxc);
  reinterpret_cast<C *>(W.p[q])[((i64)rank * rows + r) * w + kq] = S[r * nxh + i];
}

// one warp per row segment, x fastest over the whole row, 4 rows in flight per thread
template <typename C>
__global__ void k_scatter2(PeerPtrs W, const C *__restrict__ S, int nxh, i64 rows, int kxc, int rank) {
  const i64 r0 = ((i64)blockIdx.x * blockDim.y + threadIdx.y) * 4;
  for (int i = threadIdx.x; i < nxh; i += blockDim.x) {
    const int q = i / kxc, kq = i - q * kxc, w = min(kxc, nxh - q * kxc);
    C v[4];
#pragma unroll
    for (int j = 0; j < 4; j++) if (r0 + j < rows) v[j] = S[(r0 + j) * nxh + i];
#pragma unroll
    for (int j = 0; j < 4; j++) if (r0 + j < rows) reinterpret_cast<C *>(W.p[q])[((i64)rank * rows + r0 + j) * w + kq] = v[j];
  }
}

static void xwrite(int fd, const void *b, size_t n) { if (write(fd, b, n) != (ssize_t)n) _exit(2); }
static void xread(int fd, void *b, size_t n) { size_t g = 0; while (g < n) { ssize_t k = read(fd, (char *)b + g, n - g); if (k <= 0) _exit(3); g += k; } }

int main() {
  int ab[2], ba[2];
  if (pipe(ab) || pipe(ba)) return 1;
  pid_t pid = fork();
  const int me = pid == 0 ? 1 : 0;
  const int rd = me == 0 ? ba[0] : ab[0], wr = me == 0 ? ab[1] : ba[1];
  auto barrier = [&]() { char c = 1; xwrite(wr, &c, 1); xread(rd, &c, 1); };
  CK(cudaSetDevice(me));
  const int nxh = 257, kxc = 129;
  const i64 rows = 512 * 512;
  const size_t bytes = (size_t)rows * 2 * kxc * 8 + 256;
  float2 *S, *B;
  CK(cudaMalloc(&S, bytes));
  CK(cudaMalloc(&B, bytes));
  CK(cudaMemset(S, 1, bytes));
  CK(cudaMemset(B, 0, bytes));
  cudaIpcMemHandle_t hm, hp;
  CK(cudaIpcGetMemHandle(&hm, B));
  xwrite(wr, &hm, sizeof(hm));
  xread(rd, &hp, sizeof(hp));
  void *pB = nullptr;
  CK(cudaIpcOpenMemHandle(&pB, hp, cudaIpcMemLazyEnablePeerAccess));
  PeerPtrs W;
  memset(&W, 0, sizeof(W));
  W.p[me] = B;
  W.p[1 - me] = pB;
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  const char *names[] = {"app kernel, both halves", "app kernel, local half only", "app kernel, remote half only",
                         "row-loop kernel (32x8 threads, 4 rows/thread)", "row-loop kernel (64x4 threads)", "remote half then local half (2 launches)"};
  for (int var = 0; var < 6; var++) {
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
      CK(cudaDeviceSynchronize());
      barrier();
      CK(cudaEventRecord(a, st));
      dim3 blk(64, 4), grd((unsigned)((rows + 3) / 4), (nxh + 63) / 64);
      switch (var) {
        case 0: k_scatter<float2><<<grd, blk, 0, st>>>(W, S, nxh, rows, kxc, me, -1); break;
        case 1: k_scatter<float2><<<grd, blk, 0, st>>>(W, S, nxh, rows, kxc, me, me); break;
        case 2: k_scatter<float2><<<grd, blk, 0, st>>>(W, S, nxh, rows, kxc, me, 1 - me); break;
        case 3: k_scatter2<float2><<<(unsigned)((rows + 31) / 32), dim3(32, 8), 0, st>>>(W, S, nxh, rows, kxc, me); break;
        case 4: k_scatter2<float2><<<(unsigned)((rows + 15) / 16), dim3(64, 4), 0, st>>>(W, S, nxh, rows, kxc, me); break;
        case 5:
          k_scatter<float2><<<grd, blk, 0, st>>>(W, S, nxh, rows, kxc, me, 1 - me);
          k_scatter<float2><<<grd, blk, 0, st>>>(W, S, nxh, rows, kxc, me, me);
          break;
      }
      CK(cudaGetLastError());
      CK(cudaEventRecord(b, st));
      CK(cudaStreamSynchronize(st));
      float ms;
      CK(cudaEventElapsedTime(&ms, a, b));
      if (rep && ms < best) best = ms;
    }
    barrier();
    printf("[rank %d] %-48s %7.3f ms\n", me, names[var], best);
    fflush(stdout);
  }
  barrier();
  CK(cudaIpcCloseMemHandle(pB));
  barrier();
  if (me == 0) { int stt; waitpid(pid, &stt, 0); }
  return 0;
}
