// factor_diag in isolation: cycles of one 128x128 tile factorisation, one CTA of 256 threads (compile with
// -DDGIBBS_FACTOR_LOOKAHEAD=0/1)
#include <cstdio>
#include <cuda_runtime.h>
__device__ long long g_fd[64];
#define FD_T(k)                                                                                        \
  do {                                                                                                 \
    if ((threadIdx.x & 31) == 0 && blockIdx.x == 0 && threadIdx.x < 64) {                              \
      long long t_ = clock64();                                                                        \
      if ((k) > 0) atomicAdd((unsigned long long*)&g_fd[(threadIdx.x >> 5) * 8 + (k)], (unsigned long long)(t_ - fd_last)); \
      fd_last = t_;                                                                                    \
    }                                                                                                  \
  } while (0)
#define FD_DECL long long fd_last = 0
#include "../draco_b200/csrc/dense_core.cuh"
using namespace dg;

__global__ void __launch_bounds__(THREADS, 1) probe(double* out, long long* cyc, int reps) {
  extern __shared__ __align__(16) double smem[];
  double* sWp = smem;
  double* sT = sWp + WP_DOUBLES;
  double* misc = sT + PIPE_DOUBLES;
  const Lane ln;
  long long tot = 0, tpiv = 0;
  bool ok = true;
  for (int rep = 0; rep < reps; rep++) {
    for (int e = threadIdx.x; e < TILE * TILE; e += THREADS) {
      int r = e >> 7, c = e & 127;
      sT[r * LDW + c] = 1.0 / (1.0 + abs(r - c)) + (r == c ? 2.0 + 1e-3 * rep : 0.0);
    }
    __syncthreads();
    long long t0 = clock64();
    ok = factor_diag(sT, sWp, misc, ln) && ok;
    long long t1 = clock64();
    tot += t1 - t0;
    __syncthreads();
    // pivot block alone, the other warps waiting at the barrier
    for (int e = threadIdx.x; e < 64; e += THREADS) {
      int r = e >> 3, c = e & 7;
      sT[r * LDW + c] = 1.0 / (1.0 + abs(r - c)) + (r == c ? 2.0 : 0.0);
    }
    __syncthreads();
    t0 = clock64();
    if (ln.warp == 0) ok = pivot_block(sT, sWp, 0) && ok;
    __syncthreads();
    t1 = clock64();
    tpiv += t1 - t0;
  }
  if (threadIdx.x == 0) {
    cyc[blockIdx.x * 2] = tot / reps;
    cyc[blockIdx.x * 2 + 1] = tpiv / reps;
  }
  out[blockIdx.x * THREADS + threadIdx.x] = sWp[threadIdx.x] + (ok ? 1.0 : 0.0);
}

int main() {
  const int grid = 148;
  double* out; long long* cyc;
  cudaMalloc(&out, grid * THREADS * 8); cudaMalloc(&cyc, grid * 16);
  const size_t smem = sizeof(double) * (WP_DOUBLES + PIPE_DOUBLES + MISC_DOUBLES);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int rep = 0; rep < 2; rep++) probe<<<grid, THREADS, smem>>>(out, cyc, 8);
  long long h[2 * grid];
  cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
  printf("err=%s lookahead=%d\n", cudaGetErrorString(cudaGetLastError()), DGIBBS_FACTOR_LOOKAHEAD);
  printf("factor_diag : %lld cyc (cta 0), %lld (cta 147)\n", h[0], h[2 * 147]);
  printf("pivot_block + barrier, 256-thread CTA : %lld cyc\n", h[1]);
  long long fd[64];
  cudaMemcpyFromSymbol(fd, g_fd, sizeof fd);
  for (int w = 0; w < 2; w++) {
    printf("warp %d:", w);
    for (int k = 1; k < 7; k++) printf("  T%d %lld", k, fd[w * 8 + k] / 16);
    printf("   (cycles per factorisation)\n");
  }
  return 0;
}
