// TMA + mbarrier operand ring for the NT updates of the dense sampler (acc -= A B^T over row panels of the chain's
// workspace).
//
// Why.  The cp.async version (dense_core.cuh: gemm_nt) stages every operand with 16 per-thread 16-byte copies per
// 32-wide K chunk and needs two CTA barriers per chunk; ncu (profiles/r01o) put 5 % of the warp time on those barriers
// and 5 % on the copies' scoreboard.  Here ONE thread issues two `cp.async.bulk.tensor` boxes per chunk into a 4-stage
// ring; consumers wait on the stage's "full" mbarrier and release it through its "empty" mbarrier -- no CTA barrier
// inside the K loop, three chunks in flight instead of one.
//
// Layout.  A dense [128][KC] box has a row pitch of 0 mod 32 banks, and none of TMA's swizzles is conflict-free for
// the 8 x 4 DMMA fragments of doubles.  The tensor map therefore views the workspace as a 4-D tensor
//     (k mod 4, row, k div 4, slot)      strides  (8 B, pitch, 32 B, slot stride)
// and one box (4, 128, KC/4, 1) lands in shared memory as [k div 4][row][k mod 4]: the 32 lanes of a fragment load
// (row = 8 rows x k = 4) read 32 CONSECUTIVE doubles -- conflict-free with no padding at all.
#pragma once
#include <cuda.h>

#include "dense_core.cuh"

namespace dg {

#ifndef DGIBBS_TKC
#define DGIBBS_TKC 16
#endif
#ifndef DGIBBS_TNST
#define DGIBBS_TNST 4
#endif
constexpr int TKC = DGIBBS_TKC;                  // K chunk of one ring stage (TKC / 4 k-steps of the m8n8k4 DMMA)
constexpr int TNST = DGIBBS_TNST;                // ring stages
constexpr int TBOX_DOUBLES = TILE * TKC;         // one operand box: [TKC/4][128][4]
constexpr int TSTAGE_DOUBLES = 2 * TBOX_DOUBLES; // A + B
static_assert(TNST * TSTAGE_DOUBLES <= PIPE_DOUBLES, "the ring lives in the pipeline / staged-tile region");

__device__ __forceinline__ unsigned tp_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tp_mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tp_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void tp_mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tp_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tp_mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tp_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tp_mbar_wait(unsigned long long* bar, unsigned parity) {
  unsigned ok = 0;
  while (!ok) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(tp_smem_u32(bar)), "r"(parity)
        : "memory");
  }
}
// generic-proxy accesses of this thread (shared and global) are ordered before later async-proxy (TMA) accesses
__device__ __forceinline__ void tp_fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

__device__ __forceinline__ void tp_load_box(double* dst, const CUtensorMap* tm, unsigned long long* bar, int row, int kq,
                                            int slot) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
          tp_smem_u32(dst)),
      "l"(reinterpret_cast<unsigned long long>(tm)), "r"(tp_smem_u32(bar)), "r"(0), "r"(row), "r"(kq), "r"(slot)
      : "memory");
}

struct TmaRing {
  unsigned long long* full;   // [TNST] shared: count 1 (the producer's expect_tx) + the boxes' bytes
  unsigned long long* empty;  // [TNST] shared: count NWARP (one lane per consumer warp)
  unsigned g;                 // chunks that have gone through the ring since the kernel started (CTA-uniform)
};

// every thread of the CTA calls this once before the first gemm_nt_tma
__device__ __forceinline__ void tma_ring_init(TmaRing& R, unsigned long long* bars) {
  R.full = bars;
  R.empty = bars + TNST;
  R.g = 0;
  if (threadIdx.x == 0) {
    for (int s = 0; s < TNST; s++) {
      tp_mbar_init(&R.full[s], 1);
      tp_mbar_init(&R.empty[s], NWARP);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
}

// acc -= A * B^T over k-steps [KS0, KS1) of one ring stage; A, B boxes laid out [k div 4][128][4]
template <int NB0, int NB1, int KS0, int KS1>
__device__ __forceinline__ void mma_nt_box(const double* sA, const double* sB, Acc& acc, const Lane& ln) {
  const int a0 = (ln.rf[0] * 8 + ln.g) * 4 + ln.q;
  const int a1 = (ln.rf[1] * 8 + ln.g) * 4 + ln.q;
  const int bo = ln.g * 4 + ln.q;
  constexpr int NBM = NB0 > NB1 ? NB0 : NB1;
#pragma unroll
  for (int ks = KS0; ks < KS1; ks++) {
    double x0 = 0, x1 = 0;
    if (NB0 > 0) x0 = -sA[ks * (TILE * 4) + a0];
    if (NB1 > 0) x1 = -sA[ks * (TILE * 4) + a1];
#pragma unroll
    for (int nb = 0; nb < NBM; nb++) {
      const double b = sB[ks * (TILE * 4) + nb * 32 + bo];
      if (nb < NB0) dmma(acc[0][nb][0], acc[0][nb][1], x0, b);
      if (nb < NB1) dmma(acc[1][nb][0], acc[1][nb][1], x1, b);
    }
  }
}
template <int KS0, int KS1>
__device__ __forceinline__ void mma_nt_box_variant(const double* sA, const double* sB, Acc& acc, const Lane& ln, int variant) {
  switch (variant) {  // warp-uniform; predicated-off DMMAs would still occupy the pipe
    case V_FULL: mma_nt_box<16, 16, KS0, KS1>(sA, sB, acc, ln); break;
    case V_ROW0: mma_nt_box<16, 0, KS0, KS1>(sA, sB, acc, ln); break;
    case V_ROW1: mma_nt_box<0, 16, KS0, KS1>(sA, sB, acc, ln); break;
    case V_DIAG_LO: mma_nt_box<4, 16, KS0, KS1>(sA, sB, acc, ln); break;
    case V_DIAG_HI: mma_nt_box<8, 12, KS0, KS1>(sA, sB, acc, ln); break;
    default: break;
  }
}

// acc -= W[rowA : rowA+128, 0 : nchunks*TKC] * W[rowB : rowB+128, 0 : nchunks*TKC]^T, W = workspace `slot` of the map.
// Every thread of the CTA calls it; on return every warp is done with the ring (the region may be reused).
__device__ __forceinline__ void gemm_nt_tma(const CUtensorMap* tm, TmaRing& R, double* ring, Acc& acc, const Lane& ln, int slot,
                                            int rowA, int rowB, int nchunks, int variant) {
  // the ring region was last touched through the generic proxy (staged tile) and the panels in global memory were
  // written by generic stores: order them before the bulk copies
  tp_fence_proxy_async();
  __syncthreads();
  const bool producer = threadIdx.x == 0;
  const bool same = rowA == rowB;  // diagonal tile: one box is both operands
  const unsigned bytes = (same ? 1u : 2u) * TBOX_DOUBLES * (unsigned)sizeof(double);
  auto issue = [&](int c) {
    const unsigned G = R.g + (unsigned)c;
    const int s = (int)(G % TNST);
    const unsigned use = G / TNST;
    if (use > 0) tp_mbar_wait(&R.empty[s], (use - 1) & 1);  // every warp has released the stage's previous contents
    double* st = ring + s * TSTAGE_DOUBLES;
    tp_mbar_expect_tx(&R.full[s], bytes);
    tp_load_box(st, tm, &R.full[s], rowA, c * (TKC / 4), slot);
    if (!same) tp_load_box(st + TBOX_DOUBLES, tm, &R.full[s], rowB, c * (TKC / 4), slot);
  };
  if (producer)
    for (int c = 0; c < TNST - 1 && c < nchunks; c++) issue(c);
  for (int c = 0; c < nchunks; c++) {
    const unsigned G = R.g + (unsigned)c;
    const int s = (int)(G % TNST);
    tp_mbar_wait(&R.full[s], (G / TNST) & 1);
    const double* sA = ring + s * TSTAGE_DOUBLES;
    const double* sB = same ? sA : sA + TBOX_DOUBLES;
    mma_nt_box_variant<0, TKC / 8>(sA, sB, acc, ln, variant);
    // refill the stage consumed one chunk ago; by now the other warps have left it
    if (producer && c + TNST - 1 < nchunks) issue(c + TNST - 1);
    mma_nt_box_variant<TKC / 8, TKC / 4>(sA, sB, acc, ln, variant);
    __syncwarp();
    if ((threadIdx.x & 31) == 0) tp_mbar_arrive(&R.empty[s]);
  }
  R.g += (unsigned)nchunks;
  __syncthreads();
}

// NT-update policy of chol_factor_forward (dense_core.cuh) on the ring
struct NtTma {
  const CUtensorMap* tm;
  TmaRing* R;
  double* ring;
  int slot;
  __device__ __forceinline__ void operator()(Acc& acc, const Lane& ln, int rowA, int rowB, int kcols, int variant) {
    gemm_nt_tma(tm, *R, ring, acc, ln, slot, rowA, rowB, kcols / TKC, variant);
  }
};

}  // namespace dg
