#include <cstdio>
#include <cuda_runtime.h>
// Microbenchmark: fp64 pipes on sm_100a. DMMA shapes vs DFMA.
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1684(double *c, double a0, double a1, double b) {
  asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a0), "d"(a1), "d"(b));
}
__device__ __forceinline__ void dmma1688(double *c, const double *a, const double *b) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void dmma16816(double *c, const double *a, const double *b) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}
template <int ILP>
__global__ void k884(double *out, int iters, double a, double b) {
  double c[ILP][2];
#pragma unroll
  for (int i = 0; i < ILP; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) dmma884(c[i][0], c[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
__global__ void k1684(double *out, int iters, double a, double b) {
  double c[ILP][4];
#pragma unroll
  for (int i = 0; i < ILP; i++) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) dmma1684(c[i], a, b, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
__global__ void k1688(double *out, int iters, double a, double b) {
  double c[ILP][4]; double av[4] = {a, b, a, b}; double bv[2] = {b, a};
#pragma unroll
  for (int i = 0; i < ILP; i++) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) dmma1688(c[i], av, bv);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
__global__ void k16816(double *out, int iters, double a, double b) {
  double c[ILP][4]; double av[8] = {a, b, a, b, a, b, a, b}; double bv[4] = {b, a, b, a};
#pragma unroll
  for (int i = 0; i < ILP; i++) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) dmma16816(c[i], av, bv);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
__global__ void kdfma(double *out, int iters, double a, double b) {
  double c[ILP];
#pragma unroll
  for (int i = 0; i < ILP; i++) c[i] = threadIdx.x + i;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) c[i] = fma(c[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <typename F>
float timeit(F f) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
  double *out; cudaMalloc(&out, sizeof(double) * 148 * 8 * 1024);
  int iters = 20000;
  int nsm = 148;
  for (int warps : {4, 8, 16, 32}) {
    for (int cps : {1, 2}) {
      int grid = nsm * cps, thr = warps * 32;
      if (warps * cps > 64) continue;
      double nw = (double)grid * warps;
#define RUN(name, kern, ilp, flop_per) { float ms = timeit([&] { kern<ilp><<<grid, thr>>>(out, iters, 1.0000001, 0.9999999); }); \
      printf("%-8s ilp=%d warps/cta=%d cta/sm=%d : %.3f ms  %.2f TFLOP/s\n", name, ilp, warps, cps, ms, nw * iters * ilp * (double)(flop_per) / ms / 1e9); }
      RUN("m8n8k4", k884, 4, 512);
      RUN("m8n8k4", k884, 8, 512);
      RUN("m16n8k4", k1684, 4, 1024);
      RUN("m16n8k4", k1684, 8, 1024);
      RUN("m16n8k8", k1688, 4, 2048);
      RUN("m16n8k8", k1688, 8, 2048);
      RUN("m16n8k16", k16816, 4, 4096);
      RUN("m16n8k16", k16816, 8, 4096);
      RUN("dfma", kdfma, 8, 64);
      RUN("dfma", kdfma, 16, 64);
    }
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return 0;
}
