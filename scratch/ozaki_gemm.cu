// Stand-alone timing / accuracy harness of the pipelined Ozaki-split fp64 GEMM on the int8 tensor cores
// (draco_b200/csrc/ozaki_gemm.cuh: the kernels are the product's; this file only feeds and checks them).
//
//   C[b] = A[b] B[b]^T,  A [M][K], B [N][K] fp64 row-major (K contiguous), M % 128 == N % 64 == K % 128 == 0.
//
// scratch/ozaki_probe.cu showed the arithmetic (bit-exact against the host emulation) and the two rates that bracket a real
// kernel: 81.8 TFLOP/s-equivalent with the tensor pipe alone, 34.2 with load -> MMA -> epilogue one after the other.
// History of the pipelined kernel (profiles/): r04f one thread as producer + issuer, exponent blocks of 128 columns: 53.2;
// r04j separate producer warp, exponent blocks of 256 columns (half the epilogues).
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o scratch/_build/ozaki_gemm scratch/ozaki_gemm.cu -lcublas
//   ozaki_gemm [M N K batch]        (default 512 1024 1024 296: the A = T1 P^H contraction of the Wiener operator at F = 512)
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                       \
  do {                                                                                              \
    cudaError_t e_ = (x);                                                                           \
    if (e_ != cudaSuccess) {                                                                        \
      fprintf(stderr, "CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_), __FILE__, __LINE__, #x); \
      exit(3);                                                                                      \
    }                                                                                               \
  } while (0)

#include "../draco_b200/csrc/ozaki_gemm.cuh"  // the kernels under test are the product's (solver 4 of dgwiener_construct)
using namespace dgw::oz;

// ------------------------------------------------------------------------------------------------ test data
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// rows with a spread of scales (x 2^[-20, 20] per row) and entries spanning ~6 decades inside a row: what a Cholesky
// panel of a windowed system looks like to the split
__global__ void fill_kernel(double* X, size_t n, int ld, uint64_t seed) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint64_t h = mix64(seed + 0x9E3779B97F4A7C15ull * (i + 1)), hr = mix64(seed ^ (0xD1B54A32D192ED03ull * (i / ld + 1)));
    const double u = ((double)(h >> 11) + 0.5) * (1.0 / 9007199254740992.0) * 2.0 - 1.0;
    const int er = (int)(hr % 41) - 20, ee = -(int)((h >> 3) % 20);
    X[i] = ldexp(u, er + ee);
  }
}

static double time_ms(cudaEvent_t a, cudaEvent_t b) {
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, a, b));
  return ms;
}

template <int NS>
static void run(int M, int N, int K, int batch, const double* dA, const double* dB, double* dC, int nsm, const std::vector<double>& hA,
                const std::vector<double>& hB, cudaEvent_t e0, cudaEvent_t e1, double ms_cublas) {
  int8_t *qa, *qb;
  int *ea, *eb;
  const int nkb = (K / KS + BST - 1) / BST;
  CK(cudaMalloc(&qa, (size_t)batch * M * K * NS));
  CK(cudaMalloc(&qb, (size_t)batch * N * K * NS));
  CK(cudaMalloc(&ea, (size_t)batch * nkb * M * 4));
  CK(cudaMalloc(&eb, (size_t)batch * nkb * N * 4));
  auto split = [&]() {
    const long long wa = (long long)batch * (M / 8) * nkb, wb = (long long)batch * (N / 8) * nkb;
    oz_split_kernel<NS><<<(unsigned)((wa * 32 + 255) / 256), 256>>>(dA, (long long)M * K, K, M, K, TM, qa, (long long)M * K * NS, ea,
                                                                   (long long)nkb * M, batch);
    oz_split_kernel<NS><<<(unsigned)((wb * 32 + 255) / 256), 256>>>(dB, (long long)N * K, K, N, K, TN, qb, (long long)N * K * NS, eb,
                                                                   (long long)nkb * N, batch);
  };
  split();
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  split();
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  const double ms_split = time_ms(e0, e1);

  constexpr int smem = 2 * NS * (TM + TN) * KS;
  CK(cudaFuncSetAttribute(oz_gemm_kernel<NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  GemmParams p{qa, (long long)M * K * NS, qb, (long long)N * K * NS, ea, (long long)nkb * M, eb, (long long)nkb * N, dC, (long long)M * N, N, M, N, K,
               batch, 0};
  const long long ntiles = (long long)batch * (M / TM) * (N / TN);
  const unsigned grid = (unsigned)(ntiles < nsm ? ntiles : nsm);
  CK(cudaMemset(dC, 0, (size_t)batch * M * N * 8));
  oz_gemm_kernel<NS><<<grid, NTHREADS, smem>>>(p);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  // accuracy on batch 0, rows 0..63 and 3 further rows per row tile, against long double
  std::vector<double> hC((size_t)M * N);
  CK(cudaMemcpy(hC.data(), dC, hC.size() * 8, cudaMemcpyDeviceToHost));
  double worst = 0, worst_dgemm = 0;
  long long checked = 0;
  for (int i = 0; i < M; i++) {
    if (!(i < 64 || i % TM == 5 || i % TM == 77 || i % TM == 127)) continue;
    for (int j = 0; j < N; j++) {
      long double ref = 0, bound = 0;
      double plain = 0;
      for (int k = 0; k < K; k++) {
        const double a = hA[(size_t)i * K + k], bb = hB[(size_t)j * K + k];
        ref += (long double)a * bb;
        bound += fabsl((long double)a * bb);
        plain = fma(a, bb, plain);
      }
      if (bound > 0) {
        worst = fmax(worst, (double)(fabsl((long double)hC[(size_t)i * N + j] - ref) / bound));
        worst_dgemm = fmax(worst_dgemm, (double)(fabsl((long double)plain - ref) / bound));
      }
      checked++;
    }
  }
  CK(cudaEventRecord(e0));
  const int reps = 3;
  for (int r = 0; r < reps; r++) oz_gemm_kernel<NS><<<grid, NTHREADS, smem>>>(p);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  const double ms = time_ms(e0, e1) / reps, flop = 2.0 * M * N * K * batch;
  printf("{\"kernel\": \"ozaki_gemm\", \"slices\": %d, \"int8_gemms\": %d, \"M\": %d, \"N\": %d, \"K\": %d, \"batch\": %d, \"ctas\": %u, "
         "\"max_err_over_absA_absB\": %.3e, \"fp64_fma_chain_err_over_absA_absB\": %.3e, \"entries_checked\": %lld, "
         "\"ms_gemm\": %.4f, \"tflops_equiv\": %.2f, \"ms_split_A_and_B\": %.4f, \"tflops_equiv_with_split\": %.2f, "
         "\"ms_cublas_dgemm_strided_batched\": %.4f, \"tflops_cublas\": %.2f}\n",
         NS, NS * (NS + 1) / 2, M, N, K, batch, grid, worst, worst_dgemm, checked, ms, flop / ms * 1e-9, ms_split, flop / (ms + ms_split) * 1e-9,
         ms_cublas, flop / ms_cublas * 1e-9);
  fflush(stdout);
  CK(cudaFree(qa));
  CK(cudaFree(qb));
  CK(cudaFree(ea));
  CK(cudaFree(eb));
}

int main(int argc, char** argv) {
  const int M = argc > 4 ? atoi(argv[1]) : 512, N = argc > 4 ? atoi(argv[2]) : 1024, K = argc > 4 ? atoi(argv[3]) : 1024;
  const int batch = argc > 4 ? atoi(argv[4]) : 296;
  if (M % TM || N % TN || K % 128) {
    fprintf(stderr, "M %% 128, N %% 64, K %% 128 must be 0\n");
    return 2;
  }
  int nsm = 0;
  CK(cudaSetDevice(0));
  CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0));
  double *dA, *dB, *dC;
  CK(cudaMalloc(&dA, (size_t)batch * M * K * 8));
  CK(cudaMalloc(&dB, (size_t)batch * N * K * 8));
  CK(cudaMalloc(&dC, (size_t)batch * M * N * 8));
  fill_kernel<<<nsm * 8, 256>>>(dA, (size_t)batch * M * K, K, 1);
  fill_kernel<<<nsm * 8, 256>>>(dB, (size_t)batch * N * K, K, 2);
  CK(cudaDeviceSynchronize());
  std::vector<double> hA((size_t)M * K), hB((size_t)N * K);
  CK(cudaMemcpy(hA.data(), dA, hA.size() * 8, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hB.data(), dB, hB.size() * 8, cudaMemcpyDeviceToHost));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  // cuBLAS on the same shape (column-major view: C^T [N x M] = B^T-as-stored^T ... = op(B) op(A) with op = T on the first)
  cublasHandle_t hb;
  cublasCreate(&hb);
  const double one = 1.0, zero = 0.0;
  auto dgemm = [&]() {
    cublasDgemmStridedBatched(hb, CUBLAS_OP_T, CUBLAS_OP_N, N, M, K, &one, dB, K, (long long)N * K, dA, K, (long long)M * K, &zero, dC, N,
                              (long long)M * N, batch);
  };
  dgemm();
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int r = 0; r < 3; r++) dgemm();
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  const double ms_cublas = time_ms(e0, e1) / 3;
  run<8>(M, N, K, batch, dA, dB, dC, nsm, hA, hB, e0, e1, ms_cublas);
  run<7>(M, N, K, batch, dA, dB, dC, nsm, hA, hB, e0, e1, ms_cublas);
  run<6>(M, N, K, batch, dA, dB, dC, nsm, hA, hB, e0, e1, ms_cublas);
  return 0;
}
