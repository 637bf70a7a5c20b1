// latency probe: dependent DFMA / rsqrt / rcp / DMMA chains and the 8x8 pivot block, one warp, clock64
#include <cstdio>
#include <cuda_runtime.h>
#include "../draco_b200/csrc/dense_core.cuh"
using namespace dg;

__global__ void probe(double* out, long long* cyc, double seed) {
  __shared__ double sT[16 * LDW];
  __shared__ double sWp[256];
  const int lane = threadIdx.x;
  double x = seed + lane * 1e-3;
  long long t0, t1;
  // 1. dependent DFMA chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) x = fma(x, 0.999999, 1e-9);
  t1 = clock64();
  if (lane == 0) cyc[0] = (t1 - t0);
  // 2. dependent rsqrt chain
  double y = x + 2.0;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++) y = rsqrt(y) + 1.5;
  t1 = clock64();
  if (lane == 0) cyc[1] = (t1 - t0);
  // 3. dependent reciprocal chain
  double z = y + 2.0;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++) z = 1.0 / z + 1.5;
  t1 = clock64();
  if (lane == 0) cyc[2] = (t1 - t0);
  // 4. dependent DMMA chain
  double c0 = z, c1 = y;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++) dmma(c0, c1, 1e-3, 1e-3);
  t1 = clock64();
  if (lane == 0) cyc[3] = (t1 - t0);
  // 5. pivot block on an SPD 8x8 block
  for (int e = lane; e < 64; e += 32) {
    int r = e >> 3, c = e & 7;
    sT[r * LDW + c] = (r == c ? 10.0 + x * 1e-3 : 0.0) + 1.0 / (1.0 + r + c);
  }
  __syncwarp();
  t0 = clock64();
  bool ok = pivot_block(sT, sWp, 0);
  __syncwarp();
  t1 = clock64();
  if (lane == 0) cyc[4] = (t1 - t0);
  // 6. float rsqrt + 2 Newton steps in double
  double u = c0 + 3.0;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++) {
    double r = (double)rsqrtf((float)u);
    r = r * (1.5 - 0.5 * u * r * r);
    r = r * (1.5 - 0.5 * u * r * r);
    u = r + 1.5;
  }
  t1 = clock64();
  if (lane == 0) cyc[5] = (t1 - t0);
  // 7. smem store->load round trip chain (double)
  double v = u;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++) {
    sT[lane] = v;
    __syncwarp();
    v = sT[(lane + 1) & 31] + 1.0;
    __syncwarp();
  }
  t1 = clock64();
  if (lane == 0) cyc[6] = (t1 - t0);
  // 8. shuffle chain (double = 2 shfl)
  double w = v;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++) w = __shfl_xor_sync(0xffffffffu, w, 1) + 1.0;
  t1 = clock64();
  if (lane == 0) cyc[7] = (t1 - t0);
  out[lane] = x + y + z + c0 + c1 + (ok ? 1.0 : 0.0) + sWp[lane] + u + v + w;
}

int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 32 * 8); cudaMalloc(&cyc, 8 * 8);
  for (int rep = 0; rep < 2; rep++) probe<<<1, 32>>>(out, cyc, 1.0 + rep);
  long long h[8];
  cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
  printf("DFMA dependent      : %.1f cyc\n", h[0] / 256.0);
  printf("rsqrt(double)+add   : %.1f cyc\n", h[1] / 64.0);
  printf("1/x (double)+add    : %.1f cyc\n", h[2] / 64.0);
  printf("DMMA dependent      : %.1f cyc\n", h[3] / 64.0);
  printf("pivot_block 8x8     : %lld cyc\n", h[4]);
  printf("rsqrtf+2 Newton+add : %.1f cyc\n", h[5] / 64.0);
  printf("smem st->ld+add     : %.1f cyc\n", h[6] / 64.0);
  printf("shfl(double)+add    : %.1f cyc\n", h[7] / 64.0);
  return 0;
}
