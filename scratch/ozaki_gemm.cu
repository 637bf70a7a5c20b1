// Pipelined Ozaki-split fp64 GEMM on the int8 tensor cores (second stage of the probe of DESIGN.md 4.2; NOT product code).
//
//   C[b] = A[b] B[b]^T,  A [M][K], B [N][K] fp64 row-major (K contiguous), M % 128 == N % 64 == K % 128 == 0.
//
// scratch/ozaki_probe.cu showed the arithmetic (signed 7-bit slices, exponent per row x 128 columns, S (S + 1) / 2 int8
// GEMMs, bit-exact against the host emulation) and the two rates that bracket a real kernel: 81.8 TFLOP/s-equivalent with
// the tensor pipe alone, 34.2 with load -> MMA -> epilogue one after the other.  This kernel is the pipelined form:
//   * split_kernel writes the slices of a 128- (A) or 64-row (B) tile and a 64-column K stage as ONE contiguous block in the
//     canonical no-swizzle K-major core-matrix order, so a stage arrives as flat cp.async.bulk copies;
//   * warp 8: one elected thread is TMA producer and MMA issuer (two stages of 64 columns in flight; a stage is refilled as
//     soon as tcgen05.commit reports its MMAs done), tcgen05.mma.cta_group::1.kind::i8 M = 128, N = 64, K = 32, the S
//     accumulators D_d side by side in TMEM;
//   * warps 0-7: epilogue per 128-column K block (tcgen05.ld 32x32b, int64 combination, x 2^(e_i + f_j), fp64 accumulate in
//     registers), two warps per TMEM lane quarter; the next block's operands stream in underneath it.  TMEM is full (8 x 64
//     columns), so the epilogue of a block and the MMAs of the next cannot overlap: the tensor pipe idles ~30 %.
// Output tile 128 x 64 per CTA at a time, persistent grid.
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

constexpr int TM = 128, TN = 64, KB = 128, KS = 64;  // output tile, exponent block, K stage
constexpr int NTHREADS = 288;                       // 8 epilogue warps + 1 producer / issuer warp

// ------------------------------------------------------------------------------------------------ split
// One warp per (row, 128-column block).  Slices of tile (r / TR) and stage (k / 64) form one block of S * TR * 64 bytes:
//   [tile][stage][s][k16 (4)][row group (TR / 8)][row % 8][k % 16]
template <int S>
__global__ void split_kernel(const double* __restrict__ X, long long sX, int ld, int rows, int K, int TR, int8_t* __restrict__ q,
                             int* __restrict__ ex, int batch) {
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31, nkb = K / KB;
  if (warp >= (long long)batch * rows * nkb) return;
  const int r = (int)(warp % rows), kb = (int)((warp / rows) % nkb), b = (int)(warp / ((long long)rows * nkb));
  const double4 v = *reinterpret_cast<const double4*>(X + (size_t)b * sX + (size_t)r * ld + kb * KB + 4 * lane);
  double amax = fmax(fmax(fabs(v.x), fabs(v.y)), fmax(fabs(v.z), fabs(v.w)));
#pragma unroll
  for (int o = 16; o; o >>= 1) amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  int e = 0;
  if (amax > 0.0) {
    const double f = frexp(amax, &e);
    if (f == 0.5) e -= 1;
  }
  if (lane == 0) ex[((size_t)b * nkb + kb) * rows + r] = e;
  double m[4] = {ldexp(v.x, 6 - e), ldexp(v.y, 6 - e), ldexp(v.z, 6 - e), ldexp(v.w, 6 - e)};
  const int tile = r / TR, rr = r % TR, ks = kb * 2 + (lane >> 4), kk = (4 * lane) & 63, nks = K / KS, ntile = rows / TR;
  const size_t blk = (((size_t)b * ntile + tile) * nks + ks) * S * TR * KS;
  const size_t base = blk + (size_t)(kk >> 4) * TR * 16 + (size_t)(rr >> 3) * 128 + (rr & 7) * 16 + (kk & 15);
#pragma unroll
  for (int s = 0; s < S; s++) {
    uint32_t w = 0;
#pragma unroll
    for (int c = 0; c < 4; c++) {
      const double qv = rint(m[c]);
      m[c] = (m[c] - qv) * 128.0;
      w |= ((uint32_t)(uint8_t)(int8_t)(int)qv) << (8 * c);
    }
    *reinterpret_cast<uint32_t*>(q + base + (size_t)s * TR * KS) = w;
  }
}

// ------------------------------------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  unsigned ok = 0;
  long long spins = 0;
  while (!ok) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (!ok && ++spins > (1ll << 26)) {  // a probe must not hang the box
      printf("mbarrier wait timed out (block %d thread %d bar %u parity %u)\n", (int)blockIdx.x, (int)threadIdx.x, smem_u32(bar), parity);
      __trap();
    }
  }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version of sm_100
  return d;
}
__host__ __device__ constexpr uint32_t umma_idesc_i8(int M, int N) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u), "r"(0u), "r"(0u), "r"(0u)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred P;\n"
      "elect.sync _|P, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, P;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ double pow2i(int e) { return __longlong_as_double((long long)(e + 1023) << 52); }

// sum_{d < S} 2^-(12 + 7 d) D_d  =  hi 2^-33 + lo 2^-(12 + 7 (S - 1)),  hi from D_0..D_3, lo from D_4..D_{S-1}  (exact)
template <int S>
__device__ __forceinline__ double combine(const int* D) {
  long long hi = D[0], lo = 0;
#pragma unroll
  for (int d = 1; d < 4; d++) hi = hi * 128 + D[d];
#pragma unroll
  for (int d = 4; d < S; d++) lo = lo * 128 + D[d];
  return (double)hi * 0x1p-33 + (double)lo * pow2i(-(12 + 7 * (S - 1)));
}

struct GemmParams {
  const int8_t* qa;  // [batch][M / 128][K / 64][S][128 x 64]
  const int8_t* qb;  // [batch][N / 64][K / 64][S][64 x 64]
  const int* ea;     // [batch][K / 128][M]
  const int* eb;     // [batch][K / 128][N]
  double* C;         // [batch][M][ldc]
  long long sC;
  int ldc, M, N, K, batch;
};

template <int S>
__global__ void __launch_bounds__(NTHREADS, 1) ozaki_gemm_kernel(GemmParams p) {
  static_assert(S >= 5 && S <= 8, "slices");
  constexpr int A_ST = S * TM * KS, B_ST = S * TN * KS, STAGE = A_ST + B_ST;  // bytes
  constexpr int NCOL = 512;                                                   // TMEM columns (S * 64 <= 512)
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar_full[2], bar_free[2], bar_acc, bar_empty;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int mt = p.M / TM, nt = p.N / TN, nkb = p.K / KB, nks = p.K / KS;
  const long long ntiles = (long long)p.batch * mt * nt;
  if (tid == 0) {
    mbar_init(&bar_full[0], 1);
    mbar_init(&bar_full[1], 1);
    mbar_init(&bar_free[0], 1);
    mbar_init(&bar_free[1], 1);
    mbar_init(&bar_acc, 1);
    mbar_init(&bar_empty, 8);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"((uint32_t)NCOL) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;

  if (warp == 8) {
    // ------------------------------------------------------------ producer + MMA issuer
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_i8(TM, TN);
      uint64_t da0[2], db0[2];
      for (int st = 0; st < 2; st++) {
        da0[st] = umma_desc(smem_u32(smem + st * STAGE), TM * 16, 128);
        db0[st] = umma_desc(smem_u32(smem + st * STAGE + A_ST), TN * 16, 128);
      }
      auto load_stage = [&](long long tile, int kb, int st) {
        const int b = (int)(tile / ((long long)mt * nt)), rem = (int)(tile % ((long long)mt * nt)), im = rem / nt, in = rem % nt;
        const int ks = kb * 2 + st;
        const int8_t* ga = p.qa + (((size_t)b * mt + im) * nks + ks) * A_ST;
        const int8_t* gb = p.qb + (((size_t)b * nt + in) * nks + ks) * B_ST;
        uint8_t* sa = smem + st * STAGE;
        mbar_expect_tx(&bar_full[st], STAGE);
        for (int c = 0; c < A_ST; c += 8192) bulk_g2s(sa + c, ga + c, 8192, &bar_full[st]);
        for (int c = 0; c < B_ST; c += 4096) bulk_g2s(sa + A_ST + c, gb + c, 4096, &bar_full[st]);
      };
      long long tile = blockIdx.x;
      if (tile < ntiles) {
        load_stage(tile, 0, 0);
        load_stage(tile, 0, 1);
      }
      unsigned n = 0;  // K blocks issued so far by this CTA
      for (; tile < ntiles; tile += gridDim.x) {
        for (int kb = 0; kb < nkb; kb++, n++) {
          if (n > 0) mbar_wait(&bar_empty, (n - 1) & 1);  // the epilogue has drained the accumulators of the previous block
          tc_fence_after();
#pragma unroll
          for (int st = 0; st < 2; st++) {
            mbar_wait(&bar_full[st], n & 1);
            tc_fence_after();
#pragma unroll
            for (int j = 0; j < KS / 32; j++)
#pragma unroll
              for (int s = 0; s < S; s++) {
                const uint64_t da = da0[st] + (uint64_t)((s * TM * KS + j * 2 * TM * 16) >> 4);
#pragma unroll
                for (int t = 0; t < S - s; t++) {
                  const uint64_t db = db0[st] + (uint64_t)((t * TN * KS + j * 2 * TN * 16) >> 4);
                  umma_i8(tmem + (uint32_t)(s + t) * TN, da, db, idesc, (st > 0 || j > 0 || s > 0) ? 1u : 0u);
                }
              }
            umma_commit(&bar_free[st]);  // arrives when the MMAs issued so far have read their operands
          }
          umma_commit(&bar_acc);
          // refill both stages with the next K block of this CTA's work (possibly the next tile's first block)
          long long ntile = tile;
          int nkb2 = kb + 1;
          if (nkb2 == nkb) {
            nkb2 = 0;
            ntile += gridDim.x;
          }
          if (ntile < ntiles)
            for (int st = 0; st < 2; st++) {
              mbar_wait(&bar_free[st], n & 1);
              load_stage(ntile, nkb2, st);
            }
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue warps
    const int q = warp & 3, h = warp >> 2;  // TMEM lane quarter, column half
    const int row = q * 32 + lane;
    unsigned n = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int b = (int)(tile / ((long long)mt * nt)), rem = (int)(tile % ((long long)mt * nt)), im = rem / nt, in = rem % nt;
      double acc[32];
#pragma unroll
      for (int c = 0; c < 32; c++) acc[c] = 0.0;
      for (int kb = 0; kb < nkb; kb++, n++) {
        const int ea_row = __ldg(p.ea + ((size_t)b * nkb + kb) * p.M + im * TM + row);
        const int* ebp = p.eb + ((size_t)b * nkb + kb) * p.N + in * TN + h * 32;
        mbar_wait(&bar_acc, n & 1);
        tc_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < 32; c0 += 8) {
          uint32_t D[S][8];
#pragma unroll
          for (int d = 0; d < S; d++) tmem_ld8(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(d * TN + h * 32 + c0), D[d]);
          tmem_wait_ld();
#pragma unroll
          for (int c = 0; c < 8; c++) {
            int Dd[S];
#pragma unroll
            for (int d = 0; d < S; d++) Dd[d] = (int)D[d][c];
            acc[c0 + c] += combine<S>(Dd) * pow2i(ea_row + __ldg(ebp + c0 + c));
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_empty);
      }
      double* o = p.C + (size_t)b * p.sC + (size_t)(im * TM + row) * p.ldc + in * TN + h * 32;
#pragma unroll
      for (int c = 0; c < 32; c += 2) *reinterpret_cast<double2*>(o + c) = make_double2(acc[c], acc[c + 1]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)NCOL) : "memory");
}

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

template <int S>
static void run(int M, int N, int K, int batch, const double* dA, const double* dB, double* dC, int nsm, const std::vector<double>& hA,
                const std::vector<double>& hB, cudaEvent_t e0, cudaEvent_t e1, double ms_cublas) {
  int8_t *qa, *qb;
  int *ea, *eb;
  const int nkb = K / KB;
  CK(cudaMalloc(&qa, (size_t)batch * M * K * S));
  CK(cudaMalloc(&qb, (size_t)batch * N * K * S));
  CK(cudaMalloc(&ea, (size_t)batch * nkb * M * 4));
  CK(cudaMalloc(&eb, (size_t)batch * nkb * N * 4));
  auto split = [&]() {
    const long long wa = (long long)batch * M * nkb, wb = (long long)batch * N * nkb;
    split_kernel<S><<<(unsigned)((wa * 32 + 255) / 256), 256>>>(dA, (long long)M * K, K, M, K, TM, qa, ea, batch);
    split_kernel<S><<<(unsigned)((wb * 32 + 255) / 256), 256>>>(dB, (long long)N * K, K, N, K, TN, qb, eb, batch);
  };
  split();
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  split();
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  const double ms_split = time_ms(e0, e1);

  constexpr int smem = 2 * S * (TM + TN) * KS;
  CK(cudaFuncSetAttribute(ozaki_gemm_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  GemmParams p{qa, qb, ea, eb, dC, (long long)M * N, N, M, N, K, batch};
  const long long ntiles = (long long)batch * (M / TM) * (N / TN);
  const unsigned grid = (unsigned)(ntiles < nsm ? ntiles : nsm);
  CK(cudaMemset(dC, 0, (size_t)batch * M * N * 8));
  ozaki_gemm_kernel<S><<<grid, NTHREADS, smem>>>(p);
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
  for (int r = 0; r < reps; r++) ozaki_gemm_kernel<S><<<grid, NTHREADS, smem>>>(p);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  const double ms = time_ms(e0, e1) / reps, flop = 2.0 * M * N * K * batch;
  printf("{\"kernel\": \"ozaki_gemm\", \"slices\": %d, \"int8_gemms\": %d, \"M\": %d, \"N\": %d, \"K\": %d, \"batch\": %d, \"ctas\": %u, "
         "\"max_err_over_absA_absB\": %.3e, \"fp64_fma_chain_err_over_absA_absB\": %.3e, \"entries_checked\": %lld, "
         "\"ms_gemm\": %.4f, \"tflops_equiv\": %.2f, \"ms_split_A_and_B\": %.4f, \"tflops_equiv_with_split\": %.2f, "
         "\"ms_cublas_dgemm_strided_batched\": %.4f, \"tflops_cublas\": %.2f}\n",
         S, S * (S + 1) / 2, M, N, K, batch, grid, worst, worst_dgemm, checked, ms, flop / ms * 1e-9, ms_split, flop / (ms + ms_split) * 1e-9,
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
  if (M % TM || N % TN || K % KB) {
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
