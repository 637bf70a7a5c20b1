// fp64 GEMM on the INT8 tensor cores (Ozaki split):  C[b] = A[b] B[b]^T  for fp64 A [M][K], B [N][K] (K contiguous).
//
// tcgen05.mma has no f64 kind, so fp64 contractions otherwise run on the DMMA pipe (37 TFLOP/s).  Here every operand row
// is split, per block of 256 columns, into S signed 7-bit slices under one exponent,
//     x = 2^e (q_0 2^-6 + q_1 2^-13 + ... + q_{S-1} 2^-(6 + 7 (S - 1))),   q_s in [-64, 64]   (oz_split_kernel)
// and     A B^T = sum_kb 2^(e_i + f_j) sum_{d < S} 2^-(12 + 7 d) D_d,   D_d = sum_{s + t = d} Q_s R_t^T
// with the D_d exact in int32: S (S + 1) / 2 int8 GEMMs on tcgen05.mma.cta_group::1.kind::i8 (M = 128, N = 64, K = 32),
// the S accumulators side by side in TMEM.  With S = 8 the result carries 55 bits below each block's row maximum: the
// error against the exact product is BELOW that of an fp64 FMA chain (1.2e-15 vs 2.2e-15 of |A||B|^T on rows spanning 12
// decades, profiles/r04j_ozaki_gemm.log; the accuracy study of the split: scratch/ozaki_study.py).
//
// Kernel (oz_gemm_kernel, 10 warps, one 128 x 64 output tile per CTA at a time, persistent grid):
//   * the split kernel writes the slices of a (tile, 64-column K stage) as ONE contiguous block in the canonical
//     no-swizzle K-major core-matrix order (8 rows x 16 bytes), so a stage arrives as flat cp.async.bulk copies completing
//     on an mbarrier; descriptors: LBO = K-direction stride of the core matrices, SBO = 8-row-group stride (= 128 B);
//   * warp 9: one elected thread is the TMA producer: two stages in flight, a stage buffer is refilled as soon as the
//     tcgen05.commit behind its MMAs arrives;  warp 8: one elected thread issues the MMAs (72 per stage; descriptors are
//     base + constant: one uniform add per operand per MMA);
//   * warps 0-7: epilogue per 256-column block: tcgen05.ld 32x32b, int64 combination of the diagonals, x 2^(e_i + f_j),
//     fp64 accumulation in registers (two warps per TMEM lane quarter).  TMEM is full, so the epilogue of a block and the
//     MMAs of the next do not overlap; the operands of the next block stream in underneath.
// Measured on B200 (296 x [512 x 1024 x 1024], S = 8): 61.3 TFLOP/s-equivalent against 35.7 for cuBLAS
// DgemmStridedBatched on the same operands and 32.7 for this repository's DMMA tile GEMM (profiles/r04j_ozaki_gemm.log;
// ncu: IMMA sub-pipe 48.7 % active, profiles/r04m_oz_gemm_ncu_raw.csv).  Held bit for bit to oracle/ozaki_emulation.py
// through dgwiener_gemm_ozaki (tests/test_gpu_ozaki_gemm.py).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dgw {
namespace oz {

constexpr int TM = 128, TN = 64, KS = 64;  // output tile, K stage (columns per TMA stage)
constexpr int BST = 4, KB = BST * KS;      // exponent block: 4 stages = 256 columns (the last block of a row may hold 2)
constexpr int NTHREADS = 320;              // 8 epilogue warps + MMA issuer warp + TMA producer warp
constexpr int S = 8;                                // slices: fp64-equivalent
// Exponent written for a row block that holds a NaN, an infinity or a magnitude >= 2^960 (whose products overflow fp64
// anyway): every output that touches the block is NaN -- a non-finite operand must stay loud (fmax() skips NaNs and the
// conversion of a NaN to int8 gives 0: without this a NaN would silently count as zero).  Row blocks below 2^-960 count as zero.
constexpr int E_NONFINITE = 1 << 20, E_LIMIT = 960;

// ------------------------------------------------------------------------------------------------ split
// One warp per (8 rows, exponent block): lane = (row % 8) + 8 * (16-column chunk of a stage).  A thread takes its row's 16
// columns of a stage at a time, so every slice leaves as one 16-byte store and a warp's stores are 4 x 128 contiguous
// bytes.  K % 128 == 0: a block holds 4 stages, the last one of a row possibly 2.
//   q: [b][tile = row / TR][stage = k / 64][s][k16 (4)][row group (TR / 8)][row % 8][k % 16],  ex: [b][block][rows]
template <int NS>
__global__ void __launch_bounds__(256, 3) oz_split_kernel(const double* __restrict__ X, long long sX, int ld, int rows, int K, int TR,
                                                       int8_t* __restrict__ q, long long sQ, int* __restrict__ ex, long long sE, int batch) {
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31, nks = K / KS, nkb = (nks + BST - 1) / BST, ng = rows >> 3;
  if (warp >= (long long)batch * ng * nkb) return;
  const int g8 = (int)(warp % ng), kb = (int)((warp / ng) % nkb), b = (int)(warp / ((long long)ng * nkb));
  const int r = g8 * 8 + (lane & 7), kc = lane >> 3;
  const int nst = min(BST, nks - kb * BST);
  const double* xrow = X + (size_t)b * sX + (size_t)r * ld + (size_t)kb * BST * KS + kc * 16;
  // pass 1: the block's row maximum (the values are read again from L1 / L2 below: keeping 4 x 16 doubles per thread
  // costs 176 registers and leaves one CTA per SM on a bandwidth-bound kernel)
  double amax = 0.0;
  int bad = 0;  // a NaN or an infinity in the block (x - x is 0 for every finite x)
  for (int st = 0; st < nst; st++) {
    const double4* src = reinterpret_cast<const double4*>(xrow + st * KS);
#pragma unroll
    for (int v = 0; v < 4; v++) {
      const double4 t = src[v];
      amax = fmax(amax, fmax(fmax(fabs(t.x), fabs(t.y)), fmax(fabs(t.z), fabs(t.w))));
      bad |= !((t.x - t.x) + (t.y - t.y) + (t.z - t.z) + (t.w - t.w) == 0.0);
    }
  }
  amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, 8));
  amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, 16));
  bad |= __shfl_xor_sync(0xffffffffu, bad, 8);
  bad |= __shfl_xor_sync(0xffffffffu, bad, 16);
  int e = 0;
  if (amax > 0.0) {
    const double f = frexp(amax, &e);  // amax = f 2^e, f in [0.5, 1)
    if (f == 0.5) e -= 1;              // exact power of two: |x 2^-e| <= 1 still holds
  }
  const bool zero = bad || e < -E_LIMIT;  // no slices: a non-finite block is carried by its exponent, a tiny one is 0
  if (bad || e > E_LIMIT) e = E_NONFINITE;
  else if (zero) e = 0;
  if (kc == 0) ex[(size_t)b * sE + (size_t)kb * rows + r] = e;
  const int tile = r / TR, rr = r % TR;
  const bool noslice = zero || e == E_NONFINITE;
  const double sc = noslice ? 0.0 : __longlong_as_double((long long)(6 - e + 1023) << 52);  // 2^(6 - e), |e| <= 960
  // pass 2: stage by stage, 16 columns per thread, S slices of 16 bytes each
  for (int st = 0; st < nst; st++) {
    const double4* src = reinterpret_cast<const double4*>(xrow + st * KS);
    double m[16];
#pragma unroll
    for (int v = 0; v < 4; v++) {
      const double4 t = src[v];
      m[4 * v] = noslice ? 0.0 : t.x * sc; m[4 * v + 1] = noslice ? 0.0 : t.y * sc;
      m[4 * v + 2] = noslice ? 0.0 : t.z * sc; m[4 * v + 3] = noslice ? 0.0 : t.w * sc;
    }
    int8_t* dst = q + (size_t)b * sQ + (((size_t)tile * nks + (kb * BST + st)) * NS) * TR * KS + (size_t)kc * TR * 16 +
                  (size_t)(rr >> 3) * 128 + (rr & 7) * 16;
#pragma unroll
    for (int s = 0; s < NS; s++) {
      uint32_t w[4];
#pragma unroll
      for (int v = 0; v < 4; v++) {
        w[v] = 0;
#pragma unroll
        for (int c = 0; c < 4; c++) {
          const double qv = rint(m[4 * v + c]);
          m[4 * v + c] = (m[4 * v + c] - qv) * 128.0;  // exact
          w[v] |= ((uint32_t)(uint8_t)(int8_t)(int)qv) << (8 * c);
        }
      }
      *reinterpret_cast<uint4*>(dst + (size_t)s * TR * KS) = make_uint4(w[0], w[1], w[2], w[3]);
    }
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
    // a legitimate wait is one K block (~10 us); a protocol error must surface as a CUDA error, not as a hung GPU
    if (!ok && ++spins > (1ll << 28)) __trap();
  }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared-memory matrix descriptor, no swizzle, K-major (in 16-byte units ((8, n), 2) : ((1, SBO), LBO)), sm_100 version bit
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// instruction descriptor: D = S32 (bits 4-5 = 2), A = B = signed int8 (bits 7-9, 10-12 = 1), both K-major, N >> 3, M >> 4
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

// sum_{d < NS} 2^-(12 + 7 d) D_d  =  hi 2^-33 + lo 2^-(12 + 7 (NS - 1)),  hi from D_0..D_3, lo from D_4..  (exact in int64 / fp64)
template <int NS>
__device__ __forceinline__ double combine(const int* D) {
  long long hi = D[0], lo = 0;
#pragma unroll
  for (int d = 1; d < 4; d++) hi = hi * 128 + D[d];
#pragma unroll
  for (int d = 4; d < NS; d++) lo = lo * 128 + D[d];
  return (double)hi * 0x1p-33 + (double)lo * pow2i(-(12 + 7 * (NS - 1)));
}

struct GemmParams {
  const int8_t* qa; long long sQA;  // [b][M / 128][K / 64][S][128 x 64]; batch strides may be 0 (shared operand)
  const int8_t* qb; long long sQB;  // [b][N / 64][K / 64][S][64 x 64]
  const int* ea; long long sEA;     // [b][K / 128][M]
  const int* eb; long long sEB;     // [b][K / 128][N]
  double* C; long long sC;          // [b][M][ldc]
  int ldc, M, N, K, batch;
  int lower_only;  // C holds complex entries interleaved along its rows: form only the tiles that touch its lower triangle
};

template <int NS>
__global__ void __launch_bounds__(NTHREADS, 1) oz_gemm_kernel(GemmParams p) {
  static_assert(NS >= 5 && NS <= 8, "slices");
  constexpr int A_ST = NS * TM * KS, B_ST = NS * TN * KS, STAGE = A_ST + B_ST;  // bytes
  constexpr int NCOL = 512;                                                     // TMEM columns (NS * 64 <= 512)
  extern __shared__ __align__(1024) uint8_t oz_smem[];
  __shared__ __align__(8) uint64_t bar_full[2], bar_free[2], bar_acc, bar_empty;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int mt = p.M / TM, nt = p.N / TN, nks = p.K / KS, nkb = (nks + BST - 1) / BST;
  const long long per = (long long)mt * nt, ntiles = (long long)p.batch * per;
  // this CTA's tiles: blockIdx.x, + gridDim.x, ... without the ones a lower-only result does not need (complex column
  // 32 in of tile column `in` against complex row 128 im + 127 of tile row `im`)
  auto next_tile = [&](long long t) {
    while (t < ntiles) {
      const int rem = (int)(t % per), im = rem / nt, in = rem % nt;
      if (!p.lower_only || in <= 4 * im + 3) break;
      t += gridDim.x;
    }
    return t;
  };
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

  if (warp == 9) {
    // ------------------------------------------------------------ TMA producer: the CTA's stream of stages, two in flight
    if (elect_one()) {
      unsigned g = 0;  // stages loaded so far; stage g lives in buffer g & 1, its k-th use being k = g >> 1
      for (long long tile = next_tile(blockIdx.x); tile < ntiles; tile = next_tile(tile + gridDim.x)) {
        const int b = (int)(tile / per), rem = (int)(tile % per), im = rem / nt, in = rem % nt;
        const int8_t* ga = p.qa + (size_t)b * p.sQA + (size_t)im * nks * A_ST;
        const int8_t* gb = p.qb + (size_t)b * p.sQB + (size_t)in * nks * B_ST;
        for (int ks = 0; ks < nks; ks++, g++) {
          const int buf = g & 1;
          if (g >= 2) mbar_wait(&bar_free[buf], ((g >> 1) - 1) & 1);  // the MMAs of the buffer's previous stage have read it
          uint8_t* sa = oz_smem + buf * STAGE;
          mbar_expect_tx(&bar_full[buf], STAGE);
          for (int c = 0; c < A_ST; c += 8192) bulk_g2s(sa + c, ga + (size_t)ks * A_ST + c, 8192, &bar_full[buf]);
          for (int c = 0; c < B_ST; c += 4096) bulk_g2s(sa + A_ST + c, gb + (size_t)ks * B_ST + c, 4096, &bar_full[buf]);
        }
      }
    }
    __syncwarp();
  } else if (warp == 8) {
    // ------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_i8(TM, TN);
      uint64_t da0[2], db0[2];
      for (int st = 0; st < 2; st++) {
        da0[st] = umma_desc(smem_u32(oz_smem + st * STAGE), TM * 16, 128);
        db0[st] = umma_desc(smem_u32(oz_smem + st * STAGE + A_ST), TN * 16, 128);
      }
      unsigned g = 0, n = 0;  // stages / exponent blocks issued so far by this CTA
      for (long long tile = next_tile(blockIdx.x); tile < ntiles; tile = next_tile(tile + gridDim.x)) {
        for (int kb = 0; kb < nkb; kb++, n++) {
          const int nst = min(BST, nks - kb * BST);
          if (n > 0) mbar_wait(&bar_empty, (n - 1) & 1);  // the epilogue has drained the accumulators of the previous block
          tc_fence_after();
          for (int q = 0; q < nst; q++, g++) {
            const int buf = g & 1;
            mbar_wait(&bar_full[buf], (g >> 1) & 1);
            tc_fence_after();
            const uint64_t dab = da0[buf], dbb = db0[buf];
#pragma unroll
            for (int j = 0; j < KS / 32; j++)
#pragma unroll
              for (int s = 0; s < NS; s++) {
                const uint64_t da = dab + (uint64_t)((s * TM * KS + j * 2 * TM * 16) >> 4);
#pragma unroll
                for (int t = 0; t < NS - s; t++) {
                  const uint64_t db = dbb + (uint64_t)((t * TN * KS + j * 2 * TN * 16) >> 4);
                  // every diagonal d is first touched by (stage 0, j = 0, s = 0, t = d) of its block: that MMA overwrites
                  umma_i8(tmem + (uint32_t)(s + t) * TN, da, db, idesc, (j > 0 || s > 0) ? 1u : (q > 0 ? 1u : 0u));
                }
              }
            umma_commit(&bar_free[buf]);  // arrives when the MMAs issued so far have read their operands
          }
          umma_commit(&bar_acc);
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue warps
    const int qd = warp & 3, h = warp >> 2;  // TMEM lane quarter, column half
    const int row = qd * 32 + lane;
    unsigned n = 0;
    for (long long tile = next_tile(blockIdx.x); tile < ntiles; tile = next_tile(tile + gridDim.x)) {
      const int b = (int)(tile / per), rem = (int)(tile % per), im = rem / nt, in = rem % nt;
      double acc[32];
#pragma unroll
      for (int c = 0; c < 32; c++) acc[c] = 0.0;
      for (int kb = 0; kb < nkb; kb++, n++) {
        const int ea_row = __ldg(p.ea + (size_t)b * p.sEA + (size_t)kb * p.M + im * TM + row);
        const int* ebp = p.eb + (size_t)b * p.sEB + (size_t)kb * p.N + in * TN + h * 32;
        mbar_wait(&bar_acc, n & 1);
        tc_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < 32; c0 += 8) {
          uint32_t D[NS][8];
#pragma unroll
          for (int d = 0; d < NS; d++) tmem_ld8(tmem + ((uint32_t)(qd * 32) << 16) + (uint32_t)(d * TN + h * 32 + c0), D[d]);
          tmem_wait_ld();
#pragma unroll
          for (int c = 0; c < 8; c++) {
            int Dd[NS];
#pragma unroll
            for (int d = 0; d < NS; d++) Dd[d] = (int)D[d][c];
            const int es = ea_row + __ldg(ebp + c0 + c);
            // (a non-finite row block on either side: NaN, and it stays NaN through the remaining blocks)
            acc[c0 + c] += es >= E_NONFINITE / 2 ? __longlong_as_double(0x7ff8000000000000ll) : combine<NS>(Dd) * pow2i(es);
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

// ------------------------------------------------------------------------------------------------ host
inline size_t slice_bytes(size_t rows, size_t K) { return rows * K * S; }     // per matrix
inline size_t exp_ints(size_t rows, size_t K) { return rows * ((K / KS + BST - 1) / BST); }  // per matrix
constexpr int gemm_smem() { return 2 * S * (TM + TN) * KS; }

// rows % TR == 0 (TR = 128 for an A operand, 64 for a B operand), K % 128 == 0, 16-byte aligned rows (ld % 2 == 0)
inline cudaError_t split(cudaStream_t st, const double* X, long long sX, int ld, int rows, int K, int TR, int8_t* q, long long sQ, int* ex,
                         long long sE, int batch) {
  const long long warps = (long long)batch * (rows / 8) * ((K / KS + BST - 1) / BST);
  if (warps <= 0) return cudaSuccess;
  oz_split_kernel<S><<<(unsigned)((warps * 32 + 255) / 256), 256, 0, st>>>(X, sX, ld, rows, K, TR, q, sQ, ex, sE, batch);
  return cudaGetLastError();
}

inline cudaError_t gemm_nt(cudaStream_t st, int sm_count, const GemmParams& p) {
  cudaError_t e = cudaFuncSetAttribute(oz_gemm_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm_smem());
  if (e != cudaSuccess) return e;
  const long long ntiles = (long long)p.batch * (p.M / TM) * (p.N / TN);
  if (ntiles <= 0) return cudaSuccess;
  const unsigned grid = (unsigned)(ntiles < sm_count ? ntiles : sm_count);
  oz_gemm_kernel<S><<<grid, NTHREADS, gemm_smem(), st>>>(p);
  return cudaGetLastError();
}

}  // namespace oz
}  // namespace dgw
