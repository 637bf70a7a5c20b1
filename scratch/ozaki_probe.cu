// Ozaki-split fp64 GEMM probe (stretch): one 128 x 64 x K trailing update  OUT = A B^T  of the dense
// sampler's Cholesky, computed on the INT8 tensor cores from signed 7-bit slices of the fp64 panels, against the
// exact host emulation of scratch/ozaki_study.py.  NOT product code: nothing under draco_b200/ uses it.
//
//   A [128][K], B [64][K] fp64 row-major (row panels of L).  Per K block of 128 columns and per row:
//   x = 2^e (q_0 2^-6 + q_1 2^-13 + ... + q_7 2^-55), q_s in [-64, 64]  (split_kernel; the exponents per row AND K block
//   keep the full solve at the fp64 rounding floor, profiles/r04a_ozaki_study.log), and
//   OUT_ij = sum_kb 2^(e_i,kb + f_j,kb) sum_{d < 8} 2^-(12 + 7 d) D_d,   D_d = sum_{s + t = d} Q_s R_t^T   (int32, exact).
//   36 int8 GEMMs per fp64 GEMM.
//
// Two engines:
//   tcgen05  tcgen05.mma.cta_group::1.kind::i8, M = 128, N = 64, K = 32 per instruction, the 8 accumulators D_d in the 512
//            TMEM columns; operands in shared memory in the canonical no-swizzle K-major layout (8 x 16 B core
//            matrices; the split kernel writes them in that order, so a K block arrives as ONE flat TMA bulk copy);
//            one thread issues 144 MMAs per K block + tcgen05.commit; epilogue tcgen05.ld 32x32b -> int64 -> fp64.
//   imma     mma.sync.m16n8k32.s8 from the same slices (the path a recompiled pre-Blackwell kernel would take).
// Modes: check (dump D of every K block + OUT for the host to compare bit for bit), time (full update, un-overlapped
// load -> MMA -> epilogue), mma (tensor-rate ceiling: the 144 MMAs of a K block re-issued on resident operands).
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o scratch/_build/ozaki_probe scratch/ozaki_probe.cu
//   ozaki_probe <panel.bin> <engine: tcgen05|imma> <out prefix> [lbo_sbo_swapped 0|1] [desc_version 0|1] [reps]
#include <cuda_runtime.h>
#include <stdint.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <type_traits>
#include <vector>

#define CK(x)                                                                                       \
  do {                                                                                              \
    cudaError_t e_ = (x);                                                                           \
    if (e_ != cudaSuccess) {                                                                        \
      fprintf(stderr, "CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_), __FILE__, __LINE__, #x); \
      exit(3);                                                                                      \
    }                                                                                               \
  } while (0)

constexpr int TM = 128, TN = 64, KB = 128, NS = 8;
constexpr int A_SLICE = TM * KB, B_SLICE = TN * KB;          // bytes of one slice of one K block
constexpr int A_BLOCK = NS * A_SLICE, B_BLOCK = NS * B_SLICE;  // bytes of one K block (all slices)
constexpr int NPAIR = NS * (NS + 1) / 2;

// ------------------------------------------------------------------------------------------------ split
// One warp per (row, K block): row maximum -> exponent, eight signed slices per element, written in the UMMA K-major
// core-matrix order  [kb][s][k / 16][row / 8][row % 8][k % 16]  (rows = tile height).
__global__ void split_kernel(const double* __restrict__ X, int rows, int K, int8_t* __restrict__ q, int* __restrict__ ex) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nkb = K / KB;
  if (warp >= rows * nkb) return;
  const int r = warp % rows, kb = warp / rows;
  const double4 v = *reinterpret_cast<const double4*>(X + (size_t)r * K + kb * KB + 4 * lane);
  double amax = fmax(fmax(fabs(v.x), fabs(v.y)), fmax(fabs(v.z), fabs(v.w)));
#pragma unroll
  for (int o = 16; o; o >>= 1) amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  int e = 0;
  if (amax > 0.0) {
    const double f = frexp(amax, &e);  // amax = f 2^e, f in [0.5, 1)
    if (f == 0.5) e -= 1;              // exact power of two: |m| <= 1 still holds
  }
  if (lane == 0) ex[kb * rows + r] = e;
  double m[4] = {ldexp(v.x, 6 - e), ldexp(v.y, 6 - e), ldexp(v.z, 6 - e), ldexp(v.w, 6 - e)};
  const size_t base = (size_t)kb * NS * rows * KB + (size_t)(lane >> 2) * rows * 16 + (size_t)(r >> 3) * 128 + (r & 7) * 16 + (lane & 3) * 4;
#pragma unroll
  for (int s = 0; s < NS; s++) {
    uint32_t w = 0;
#pragma unroll
    for (int c = 0; c < 4; c++) {
      const double qv = rint(m[c]);
      m[c] = (m[c] - qv) * 128.0;
      w |= ((uint32_t)(uint8_t)(int8_t)(int)qv) << (8 * c);
    }
    *reinterpret_cast<uint32_t*>(q + base + (size_t)s * rows * KB) = w;
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
    if (!ok && ++spins > (1ll << 26)) {  // a probe must not hang the box: give up after ~seconds
      printf("mbarrier wait timed out (block %d thread %d)\n", (int)blockIdx.x, (int)threadIdx.x);
      __trap();
    }
  }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared-memory matrix descriptor, no swizzle, K-major: in 16-byte units ((8, n), 2) : ((1, SBO), LBO)
// (cute/atom/mma_traits_sm100.hpp make_umma_desc<Major::K>, LayoutType::INTERLEAVE; version 1 = Blackwell)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t version) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)(version & 3) << 46;
  return d;
}
// instruction descriptor (cute/arch/mma_sm100_desc.hpp InstrDescriptor): D = S32, A = B = signed int8, both K-major
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
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ double pow2i(int e) { return __longlong_as_double((long long)(e + 1023) << 52); }  // 2^e, normal range
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ bool elect_one() {  // one lane of a converged warp
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

struct Params {
  const int8_t* qa;   // [nkb][NS][A_SLICE]
  const int8_t* qb;   // [nkb][NS][B_SLICE]
  const int* ea;      // [nkb][TM]
  const int* eb;      // [nkb][TN]
  double* out;        // [grid][TM][TN]
  int* dump;          // [nkb][NS][TM][TN] int32 accumulators (CTA 0; may be null)
  int nkb, reps, mode;  // mode 0: full update; 1: MMA only (operands of K block 0 resident); 2: MMA rate of shape 128 x sweepN x 32
  uint32_t swapped, version;
  int sweepN;
};

// hi = ((D0 128 + D1) 128 + D2) 128 + D3,  lo likewise from D4..D7  (|.| < 2^45: exact in int64 and in fp64);
// sum_d 2^-(12 + 7 d) D_d = hi 2^-33 + lo 2^-61
__device__ __forceinline__ double combine8(const int* D) {
  long long hi = D[0], lo = D[4];
#pragma unroll
  for (int d = 1; d < 4; d++) {
    hi = hi * 128 + D[d];
    lo = lo * 128 + D[4 + d];
  }
  return (double)hi * 0x1p-33 + (double)lo * 0x1p-61;
}

// ------------------------------------------------------------------------------------------------ tcgen05 engine
__global__ void __launch_bounds__(128, 1) ozaki_tcgen05_kernel(Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                 // [NS][KB / 16][TM / 8][8][16]
  uint8_t* sB = smem + A_BLOCK;       // [NS][KB / 16][TN / 8][8][16]
  __shared__ __align__(8) uint64_t bar_load, bar_mma;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    mbar_init(&bar_load, 1);
    mbar_init(&bar_mma, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  const uint32_t idesc = umma_idesc_i8(TM, TN);
  // core matrices of one slice: [k / 16][row / 8] x 128 B  ->  LBO (K-direction stride) = rows * 16, SBO (row-group stride) = 128;
  // r04a: with the two swapped the MMA faults (illegal address), descriptor version 0 and 1 give the same bits
  constexpr uint32_t a_k = TM * 16, b_k = TN * 16, rg = 128;
  unsigned ph_load = 0, ph_mma = 0;
  // descriptors = base + constant: the start-address field is the low 14 bits (16-byte units) and cannot carry inside 228 KB.
  // One uniform add per operand per MMA -- with the descriptor rebuilt from a pointer each time (r04a/b) the issuing
  // thread needed ~150 cycles per MMA whatever its shape.
  const uint64_t da0 = umma_desc(smem_u32(sA), a_k, rg, p.version), db0 = umma_desc(smem_u32(sB), b_k, rg, p.version);

  auto issue_block = [&]() {  // 4 k-steps x 36 slice pairs into the 8 accumulators (fresh at the first pair of each d)
#pragma unroll
    for (int j = 0; j < KB / 32; j++)
#pragma unroll
      for (int s = 0; s < NS; s++) {
        const uint64_t da = da0 + (uint64_t)((s * A_SLICE + j * 2 * a_k) >> 4);
#pragma unroll
        for (int t = 0; t < NS - s; t++) {
          const uint64_t db = db0 + (uint64_t)((t * B_SLICE + j * 2 * b_k) >> 4);
          umma_i8(tmem + (uint32_t)(s + t) * TN, da, db, idesc, (j > 0 || s > 0) ? 1u : 0u);
        }
      }
    umma_commit(&bar_mma);
  };
  auto load_block = [&](int kb) {
    mbar_expect_tx(&bar_load, A_BLOCK + B_BLOCK);
    for (int c = 0; c < A_BLOCK; c += 16384) bulk_g2s(sA + c, p.qa + (size_t)kb * A_BLOCK + c, 16384, &bar_load);
    for (int c = 0; c < B_BLOCK; c += 16384) bulk_g2s(sB + c, p.qb + (size_t)kb * B_BLOCK + c, 16384, &bar_load);
  };
  auto sweep = [&](auto nconst) {  // issue rate of one MMA shape: 144 MMAs per commit, round-robin over the 512 / N accumulators
    constexpr int n = decltype(nconst)::value, nacc = 512 / n;
    const uint32_t id = umma_idesc_i8(TM, n);
    const uint64_t dbn = umma_desc(smem_u32(sB), (uint32_t)n * 16, rg, p.version);
    for (int r = 0; r < p.reps; r++) {
#pragma unroll
      for (int i = 0; i < 144; i++) {
        const uint64_t da = da0 + (uint64_t)(((i & 7) * A_SLICE + ((i >> 3) & 3) * 2 * a_k) >> 4);
        const uint64_t db = dbn + (uint64_t)((((i >> 3) & 3) * 2 * n * 16) >> 4);
        umma_i8(tmem + (uint32_t)((i % nacc) * n), da, db, id, 1u);
      }
      umma_commit(&bar_mma);
      mbar_wait(&bar_mma, ph_mma);
      ph_mma ^= 1;
    }
  };

  if (p.mode == 2) {
    if (warp == 0) {
      if (elect_one()) {
        load_block(0);
        mbar_wait(&bar_load, ph_load);
        tc_fence_after();
        if (p.sweepN == 64) sweep(std::integral_constant<int, 64>{});
        else if (p.sweepN == 128) sweep(std::integral_constant<int, 128>{});
        else sweep(std::integral_constant<int, 256>{});
      }
      __syncwarp();
    }
  } else if (p.mode == 1) {  // tensor-rate ceiling
    if (warp == 0) {
      if (elect_one()) {
        load_block(0);
        mbar_wait(&bar_load, ph_load);
        tc_fence_after();
        for (int r = 0; r < p.reps; r++) {
          issue_block();
          mbar_wait(&bar_mma, ph_mma);
          ph_mma ^= 1;
        }
      }
      __syncwarp();
    }
  } else {
    const int row = tid;  // TMEM lane = row of the tile; warp w owns lanes 32 w .. 32 w + 31
    for (int rep = 0; rep < p.reps; rep++) {
      double acc[TN];
#pragma unroll
      for (int c = 0; c < TN; c++) acc[c] = 0.0;
      for (int kb = 0; kb < p.nkb; kb++) {
        if (warp == 0) {
          if (elect_one()) {
            load_block(kb);
            mbar_wait(&bar_load, ph_load);
            tc_fence_after();
            issue_block();
          }
          __syncwarp();  // the other lanes of warp 0 park here instead of spinning beside the issuing lane
        }
        ph_load ^= 1;
        mbar_wait(&bar_mma, ph_mma);
        ph_mma ^= 1;
        tc_fence_after();
        const int ea_row = p.ea[kb * TM + row];
#pragma unroll
        for (int c0 = 0; c0 < TN; c0 += 8) {
          uint32_t D[NS][8];
#pragma unroll
          for (int d = 0; d < NS; d++) tmem_ld8(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(d * TN + c0), D[d]);
          tmem_wait_ld();
#pragma unroll
          for (int c = 0; c < 8; c++) {
            int Dd[NS];
#pragma unroll
            for (int d = 0; d < NS; d++) Dd[d] = (int)D[d][c];
            if (p.dump && blockIdx.x == 0 && rep == 0) {
#pragma unroll
              for (int d = 0; d < NS; d++) p.dump[(((size_t)kb * NS + d) * TM + row) * TN + c0 + c] = Dd[d];
            }
            acc[c0 + c] += combine8(Dd) * pow2i(ea_row + p.eb[kb * TN + c0 + c]);
          }
        }
        tc_fence_before();
        __syncthreads();  // every warp has drained TMEM (and the MMAs have released the operands) before the next block
      }
      double* o = p.out + ((size_t)blockIdx.x * TM + row) * TN;
#pragma unroll
      for (int c = 0; c < TN; c += 2) *reinterpret_cast<double2*>(o + c) = make_double2(acc[c], acc[c + 1]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

// ------------------------------------------------------------------------------------------------ mma.sync engine
__device__ __forceinline__ void imma16832(int* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// byte offset of element (row, k) inside one slice of a K block (core-matrix order, see split_kernel)
__device__ __forceinline__ int core_off(int rows, int r, int k) { return (k >> 4) * rows * 16 + (r >> 3) * 128 + (r & 7) * 16 + (k & 15); }

// 4 warps, warp w owns rows 32 w .. 32 w + 31 (two m16 fragments) x all 64 columns (eight n8 fragments).  Per K block and
// diagonal d: iacc = sum_{s + t = d} Q_s R_t^T (int32), folded into fp64 with 2^-(12 + 7 d) and the block's exponents.
__global__ void __launch_bounds__(128, 1) ozaki_imma_kernel(Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + A_BLOCK;
  __shared__ __align__(8) uint64_t bar_load;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
  if (tid == 0) {
    mbar_init(&bar_load, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  unsigned ph = 0;
  for (int rep = 0; rep < p.reps; rep++) {
    double acc[2][8][4];
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
      for (int j = 0; j < 8; j++)
#pragma unroll
        for (int c = 0; c < 4; c++) acc[i][j][c] = 0.0;
    const int nkb = p.mode == 1 ? 1 : p.nkb;
    for (int kb = 0; kb < nkb; kb++) {
      __syncthreads();  // the previous block's fragments have been read
      if (tid == 0) {
        mbar_expect_tx(&bar_load, A_BLOCK + B_BLOCK);
        for (int c = 0; c < A_BLOCK; c += 16384) bulk_g2s(sA + c, p.qa + (size_t)kb * A_BLOCK + c, 16384, &bar_load);
        for (int c = 0; c < B_BLOCK; c += 16384) bulk_g2s(sB + c, p.qb + (size_t)kb * B_BLOCK + c, 16384, &bar_load);
      }
      mbar_wait(&bar_load, ph);
      ph ^= 1;
      double sa[2][2], sb[8][2];
#pragma unroll
      for (int i = 0; i < 2; i++) {
        sa[i][0] = ldexp(1.0, p.ea[kb * TM + warp * 32 + i * 16 + g]);
        sa[i][1] = ldexp(1.0, p.ea[kb * TM + warp * 32 + i * 16 + g + 8]);
      }
#pragma unroll
      for (int j = 0; j < 8; j++) {
        sb[j][0] = ldexp(1.0, p.eb[kb * TN + j * 8 + 2 * q]);
        sb[j][1] = ldexp(1.0, p.eb[kb * TN + j * 8 + 2 * q + 1]);
      }
      for (int d = 0; d < NS; d++) {
        int iacc[2][8][4];
#pragma unroll
        for (int i = 0; i < 2; i++)
#pragma unroll
          for (int j = 0; j < 8; j++)
#pragma unroll
            for (int c = 0; c < 4; c++) iacc[i][j][c] = 0;
        for (int s = 0; s <= d; s++) {
          const uint8_t* a_s = sA + s * A_SLICE;
          const uint8_t* b_t = sB + (d - s) * B_SLICE;
#pragma unroll
          for (int ks = 0; ks < KB / 32; ks++) {
            uint32_t a[2][4];
#pragma unroll
            for (int i = 0; i < 2; i++) {
              const int r0 = warp * 32 + i * 16 + g;
              a[i][0] = *reinterpret_cast<const uint32_t*>(a_s + core_off(TM, r0, ks * 32 + 4 * q));
              a[i][1] = *reinterpret_cast<const uint32_t*>(a_s + core_off(TM, r0 + 8, ks * 32 + 4 * q));
              a[i][2] = *reinterpret_cast<const uint32_t*>(a_s + core_off(TM, r0, ks * 32 + 16 + 4 * q));
              a[i][3] = *reinterpret_cast<const uint32_t*>(a_s + core_off(TM, r0 + 8, ks * 32 + 16 + 4 * q));
            }
#pragma unroll
            for (int j = 0; j < 8; j++) {
              const uint32_t b0 = *reinterpret_cast<const uint32_t*>(b_t + core_off(TN, j * 8 + g, ks * 32 + 4 * q));
              const uint32_t b1 = *reinterpret_cast<const uint32_t*>(b_t + core_off(TN, j * 8 + g, ks * 32 + 16 + 4 * q));
              imma16832(iacc[0][j], a[0], b0, b1);
              imma16832(iacc[1][j], a[1], b0, b1);
            }
          }
        }
        const double w = ldexp(1.0, -(12 + 7 * d));
#pragma unroll
        for (int i = 0; i < 2; i++)
#pragma unroll
          for (int j = 0; j < 8; j++)
#pragma unroll
            for (int c = 0; c < 4; c++) {
              if (p.dump && blockIdx.x == 0 && rep == 0 && p.mode == 0) {
                const int row = warp * 32 + i * 16 + g + (c >> 1) * 8, col = j * 8 + 2 * q + (c & 1);
                p.dump[(((size_t)kb * NS + d) * TM + row) * TN + col] = iacc[i][j][c];
              }
              acc[i][j][c] += (double)iacc[i][j][c] * (w * sa[i][c >> 1] * sb[j][c & 1]);
            }
      }
    }
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
      for (int j = 0; j < 8; j++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int row = warp * 32 + i * 16 + g + h * 8, col = j * 8 + 2 * q;
          *reinterpret_cast<double2*>(p.out + ((size_t)blockIdx.x * TM + row) * TN + col) = make_double2(acc[i][j][2 * h], acc[i][j][2 * h + 1]);
        }
  }
}

// ------------------------------------------------------------------------------------------------ fp64 reference on the device
// OUT = A B^T on the fp64 tensor pipe (mma.sync.m8n8k4.f64), 4 warps x (32 rows x 64 columns): the arithmetic the product's
// DMMA updates perform (k ascending, one accumulator per entry)
__global__ void __launch_bounds__(128, 1) dmma_ref_kernel(const double* __restrict__ A, const double* __restrict__ B, int K, double* out, int reps) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
  for (int rep = 0; rep < reps; rep++) {
    double c[4][8][2];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 8; j++) c[i][j][0] = c[i][j][1] = 0.0;
    for (int k = 0; k < K; k += 4) {
      double a[4], b[8];
#pragma unroll
      for (int i = 0; i < 4; i++) a[i] = A[(size_t)(warp * 32 + i * 8 + g) * K + k + q];
#pragma unroll
      for (int j = 0; j < 8; j++) b[j] = B[(size_t)(j * 8 + g) * K + k + q];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 8; j++)
          asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[i][j][0]), "+d"(c[i][j][1]) : "d"(a[i]), "d"(b[j]));
    }
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 8; j++)
        *reinterpret_cast<double2*>(out + ((size_t)blockIdx.x * TM + warp * 32 + i * 8 + g) * TN + j * 8 + 2 * q) = make_double2(c[i][j][0], c[i][j][1]);
  }
}

// ------------------------------------------------------------------------------------------------ host
static double time_ms(cudaEvent_t a, cudaEvent_t b) {
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, a, b));
  return ms;
}

int main(int argc, char** argv) {
  if (argc < 4) {
    fprintf(stderr, "usage: %s panel.bin tcgen05|imma out_prefix [swapped] [version] [reps]\n", argv[0]);
    return 2;
  }
  const bool tc = !strcmp(argv[2], "tcgen05");
  const std::string prefix = argv[3];
  const uint32_t swapped = argc > 4 ? atoi(argv[4]) : 0, version = argc > 5 ? atoi(argv[5]) : 1;
  const int reps = argc > 6 ? atoi(argv[6]) : 20;
  FILE* f = fopen(argv[1], "rb");
  if (!f) { fprintf(stderr, "cannot open %s\n", argv[1]); return 2; }
  int hdr[4];
  if (fread(hdr, 4, 4, f) != 4) return 2;
  const int M = hdr[0], N = hdr[1], K = hdr[2], S = hdr[3];
  if (M != TM || N != TN || K % KB || S != NS) { fprintf(stderr, "panel %d x %d x %d, S = %d does not match the build\n", M, N, K, S); return 2; }
  const int nkb = K / KB;
  std::vector<double> A((size_t)M * K), B((size_t)N * K);
  if (fread(A.data(), 8, A.size(), f) != A.size() || fread(B.data(), 8, B.size(), f) != B.size()) return 2;
  fclose(f);

  int dev = 0, nsm = 0;
  CK(cudaSetDevice(dev));
  CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  double *dA, *dB, *dOut, *dRef;
  int8_t *qa, *qb;
  int *ea, *eb, *dump;
  CK(cudaMalloc(&dA, A.size() * 8));
  CK(cudaMalloc(&dB, B.size() * 8));
  CK(cudaMalloc(&qa, (size_t)nkb * A_BLOCK));
  CK(cudaMalloc(&qb, (size_t)nkb * B_BLOCK));
  CK(cudaMalloc(&ea, (size_t)nkb * TM * 4));
  CK(cudaMalloc(&eb, (size_t)nkb * TN * 4));
  CK(cudaMalloc(&dump, (size_t)nkb * NS * TM * TN * 4));
  CK(cudaMalloc(&dOut, (size_t)nsm * TM * TN * 8));
  CK(cudaMalloc(&dRef, (size_t)nsm * TM * TN * 8));
  CK(cudaMemcpy(dA, A.data(), A.size() * 8, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size() * 8, cudaMemcpyHostToDevice));
  CK(cudaMemset(dump, 0xff, (size_t)nkb * NS * TM * TN * 4));
  CK(cudaMemset(dOut, 0, (size_t)nsm * TM * TN * 8));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));

  auto run_split = [&]() {
    split_kernel<<<(TM * nkb * 32 + 255) / 256, 256>>>(dA, TM, K, qa, ea);
    split_kernel<<<(TN * nkb * 32 + 255) / 256, 256>>>(dB, TN, K, qb, eb);
  };
  run_split();
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int r = 0; r < 20; r++) run_split();
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  const double ms_split = time_ms(e0, e1) / 20;

  const size_t smem = A_BLOCK + B_BLOCK;
  auto kern = tc ? ozaki_tcgen05_kernel : ozaki_imma_kernel;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  Params p{qa, qb, ea, eb, dOut, dump, nkb, 1, 0, swapped, version, 64};

  // ---- check: one CTA, accumulators + result to files
  kern<<<1, 128, smem>>>(p);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  std::vector<int> hd((size_t)nkb * NS * TM * TN);
  std::vector<double> ho((size_t)TM * TN), hr((size_t)TM * TN);
  CK(cudaMemcpy(hd.data(), dump, hd.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(ho.data(), dOut, ho.size() * 8, cudaMemcpyDeviceToHost));
  dmma_ref_kernel<<<1, 128>>>(dA, dB, K, dRef, 1);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(hr.data(), dRef, hr.size() * 8, cudaMemcpyDeviceToHost));
  {
    FILE* o = fopen((prefix + "_D.bin").c_str(), "wb");
    fwrite(hd.data(), 4, hd.size(), o);
    fclose(o);
    o = fopen((prefix + "_out.bin").c_str(), "wb");
    fwrite(ho.data(), 8, ho.size(), o);
    fclose(o);
    o = fopen((prefix + "_dmma.bin").c_str(), "wb");
    fwrite(hr.data(), 8, hr.size(), o);
    fclose(o);
  }
  // quick on-box verdict: Ozaki result against the device's own DMMA result
  double mx = 0, scale = 0;
  for (size_t i = 0; i < ho.size(); i++) {
    mx = fmax(mx, fabs(ho[i] - hr[i]));
    scale = fmax(scale, fabs(hr[i]));
  }

  // ---- timings: one CTA per SM, every CTA the same update
  p.dump = nullptr;
  p.reps = reps;
  kern<<<nsm, 128, smem>>>(p);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  kern<<<nsm, 128, smem>>>(p);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  const double ms_full = time_ms(e0, e1) / reps;
  p.mode = 1;
  p.reps = reps * nkb;
  kern<<<nsm, 128, smem>>>(p);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  kern<<<nsm, 128, smem>>>(p);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  const double ms_mma = time_ms(e0, e1) / reps;  // per update-equivalent (nkb K blocks)
  if (tc) {  // MMA issue rate by shape (operand contents irrelevant)
    for (int n : {64, 128, 256}) {
      p.mode = 2;
      p.sweepN = n;
      p.reps = 200;
      kern<<<nsm, 128, smem>>>(p);
      CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(e0));
      kern<<<nsm, 128, smem>>>(p);
      CK(cudaEventRecord(e1));
      CK(cudaDeviceSynchronize());
      const double ms = time_ms(e0, e1);
      const double ops = 2.0 * TM * n * 32 * 144.0 * p.reps * nsm;
      printf("{\"engine\": \"tcgen05\", \"mma_shape\": \"128x%dx32 kind::i8 cta_group::1\", \"ctas\": %d, \"int8_tops\": %.1f, \"cycles_per_mma_at_1965MHz\": %.1f, "
             "\"tflops_equiv_at_36_gemms\": %.2f}\n", n, nsm, ops / ms * 1e-9, ms * 1e-3 * 1.965e9 / (144.0 * p.reps), ops / ms * 1e-9 / NPAIR);
    }
  }
  dmma_ref_kernel<<<nsm, 128>>>(dA, dB, K, dRef, reps);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  dmma_ref_kernel<<<nsm, 128>>>(dA, dB, K, dRef, reps);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  const double ms_dmma = time_ms(e0, e1) / reps;

  const double flop = 2.0 * TM * TN * K * nsm;                 // fp64-equivalent flop of one update on every SM
  const double iops = flop * NPAIR;                            // int8 multiply-adds x 2 actually issued
  printf("{\"engine\": \"%s\", \"M\": %d, \"N\": %d, \"K\": %d, \"slices\": %d, \"int8_gemms_per_fp64_gemm\": %d, \"ctas\": %d, "
         "\"lbo_sbo_swapped\": %u, \"desc_version\": %u, \"max_abs_diff_vs_device_dmma\": %.3e, \"max_abs_dmma\": %.3e, "
         "\"ms_split_both_panels\": %.5f, \"ms_update_unoverlapped\": %.5f, \"tflops_equiv_unoverlapped\": %.2f, "
         "\"ms_update_mma_only\": %.5f, \"tflops_equiv_mma_only\": %.2f, \"int8_tops_mma_only\": %.1f, "
         "\"ms_update_naive_dmma_from_global\": %.5f, \"tflops_naive_dmma\": %.2f}\n",
         argv[2], TM, TN, K, NS, NPAIR, nsm, swapped, version, mx, scale, ms_split, ms_full, flop / ms_full * 1e-9, ms_mma,
         flop / ms_mma * 1e-9, iops / ms_mma * 1e-9, ms_dmma, flop / ms_dmma * 1e-9);
  return 0;
}
