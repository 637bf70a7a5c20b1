// cost of the trailing-update building block: 4 blocks of one row strip (LDS, 8 DMMA, STS) per call
#include <cstdio>
#include <cuda_runtime.h>
#include "../draco_b200/csrc/dense_core.cuh"
using namespace dg;

template <int U>
__device__ __forceinline__ void update_rowU(double* sT, int rb, int cb0, int p, int g, int q) {
  const double* ap = sT + (rb * 8 + g) * LDW + 8 * p + q;
  const double a0 = -ap[0], a1 = -ap[4];
  double2 c[U];
  double bb0[U], bb1[U];
#pragma unroll
  for (int u = 0; u < U; u++) {
    c[u] = *reinterpret_cast<const double2*>(sT + (rb * 8 + g) * LDW + 8 * (cb0 + u) + 2 * q);
    const double* bp = sT + ((cb0 + u) * 8 + g) * LDW + 8 * p + q;
    bb0[u] = bp[0];
    bb1[u] = bp[4];
  }
#pragma unroll
  for (int u = 0; u < U; u++) dmma(c[u].x, c[u].y, a0, bb0[u]);
#pragma unroll
  for (int u = 0; u < U; u++) dmma(c[u].x, c[u].y, a1, bb1[u]);
#pragma unroll
  for (int u = 0; u < U; u++) *reinterpret_cast<double2*>(sT + (rb * 8 + g) * LDW + 8 * (cb0 + u) + 2 * q) = c[u];
}

__global__ void __launch_bounds__(THREADS, 1) probe(double* out, long long* cyc, int nwarps) {
  extern __shared__ __align__(16) double smem[];
  double* sT = smem;
  const Lane ln;
  for (int e = threadIdx.x; e < TILE * LDW; e += THREADS) sT[e] = 1e-3 * (e % 97);
  __syncthreads();
  long long t0 = 0, t1 = 0, t2 = 0, t3 = 0;
  if (ln.warp < nwarps) {
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 64; i++) update_rowU<4>(sT, 15 - (ln.warp & 7), (i & 1) * 4 + 1, 0, ln.g, ln.q);
    t1 = clock64();
#pragma unroll 1
    for (int i = 0; i < 64; i++) update_rowU<8>(sT, 15 - (ln.warp & 7), 1, 0, ln.g, ln.q);
    t2 = clock64();
#pragma unroll 1
    for (int i = 0; i < 64; i++) update_rowU<1>(sT, 15 - (ln.warp & 7), 1 + (i & 7), 0, ln.g, ln.q);
    t3 = clock64();
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) {
    cyc[ln.warp * 3] = (t1 - t0) / 64;
    cyc[ln.warp * 3 + 1] = (t2 - t1) / 64;
    cyc[ln.warp * 3 + 2] = (t3 - t2) / 64;
  }
  out[threadIdx.x] = sT[threadIdx.x * 7];
}

int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, THREADS * 8); cudaMalloc(&cyc, 8 * 3 * 8);
  const size_t smem = sizeof(double) * TILE * LDW;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int nw : {1, 2, 4, 8}) {
    probe<<<1, THREADS, smem>>>(out, cyc, nw);
    probe<<<1, THREADS, smem>>>(out, cyc, nw);
    long long h[24];
    cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    printf("%d warps: cycles per call  U=4: %lld  U=8: %lld  U=1: %lld   (warp 0)   err=%s\n", nw, h[0], h[1], h[2], cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
