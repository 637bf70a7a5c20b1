// host-free compile check + device test harness (run on the GPU box): compares the in-smem FFT with a direct DFT
#include <cstdio>
#include <cmath>
#include <vector>
#include "dense_core.cuh"
using namespace dg;
__global__ void fft_kernel(const double2* in, double2* out, const double2* tw, int N) {
  extern __shared__ double2 sm[];
  for (int i = threadIdx.x; i < N; i += THREADS) sm[i] = in[i];
  __syncthreads();
  const double2* r = fft_inverse_smem(sm, sm + N, tw, N);
  for (int i = threadIdx.x; i < N; i += THREADS) out[i] = r[i];
}
int main() {
  for (int N = 8; N <= 4096; N *= 2) {
    std::vector<double2> h(N), tw(N / 2), o(N);
    for (int i = 0; i < N; i++) h[i] = make_double2(sin(1.3 * i) + 0.1 * i / N, cos(0.7 * i * i));
    for (int m = 0; m < N / 2; m++) tw[m] = make_double2(cos(2 * M_PI * m / N), sin(2 * M_PI * m / N));
    double2 *din, *dout, *dtw;
    cudaMalloc(&din, N * 16); cudaMalloc(&dout, N * 16); cudaMalloc(&dtw, N * 8);
    cudaMemcpy(din, h.data(), N * 16, cudaMemcpyHostToDevice); cudaMemcpy(dtw, tw.data(), N * 8, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(fft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * N * 16);
    fft_kernel<<<1, THREADS, 2 * N * 16>>>(din, dout, dtw, N);
    cudaMemcpy(o.data(), dout, N * 16, cudaMemcpyDeviceToHost);
    double err = 0, nrm = 0;
    for (int n = 0; n < N; n += (N > 256 ? 37 : 1)) {
      long double sr = 0, si = 0;
      for (int i = 0; i < N; i++) { long double a = 2 * M_PIl * ((long long)i * n % N) / N; sr += h[i].x * cosl(a) - h[i].y * sinl(a); si += h[i].x * sinl(a) + h[i].y * cosl(a); }
      err = fmax(err, fmax(fabs(o[n].x - (double)sr), fabs(o[n].y - (double)si))); nrm = fmax(nrm, fabs((double)sr));
    }
    printf("N=%d max err %.3e (scale %.3e) %s\n", N, err, nrm, cudaGetErrorString(cudaGetLastError()));
  }
}
