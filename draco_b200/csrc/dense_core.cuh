// Generic pieces of the dense fp64 DMMA Cholesky machinery (tiles, pipelined K loops, blocked
// triangular solves, diagonal-tile factorisation, in-smem FFT) shared by the delay-spectrum sampler
// (sampler_dense.cu) and the cross-spectrum sampler (sampler_cross.cu).
#pragma once
#include "common.cuh"
#include "internal.h"

namespace dg {

// k-steps (of 4) of a chunk whose DMMAs are issued BEFORE the next chunk's cp.async copies (0: copies first).  The
// copies carry a memory clobber, so issuing them first holds back the fragment loads of the chunk that has just become
// ready and the tensor pipe idles behind every barrier; with 2 the pipe starts at once and the copies still have three
// quarters of a chunk (~3 us) to land: 31.46 -> 32.24 TFLOP/s at cfg2 (88.7 -> 90.9 % of cuBLAS DGEMM).
#ifndef DGIBBS_SPLIT_CHUNK
#define DGIBBS_SPLIT_CHUNK 2
#endif
constexpr int STAGES = 2;  // a KC = 32 chunk is ~4 us of DMMA work: one chunk in flight covers the L2 / HBM latency
constexpr int STAGE_DOUBLES = 2 * TILE * LDA_NT;      // A[128][20] + B[128][20]  (NN form needs less)
constexpr int TILE_DOUBLES = TILE * LDW;              // one staged 128x128 tile, pitch 132
constexpr int PIPE_DOUBLES = (STAGES * STAGE_DOUBLES > TILE_DOUBLES) ? STAGES * STAGE_DOUBLES : TILE_DOUBLES;
constexpr int WP_BLOCKS = 16 * 17 / 2;                // lower-triangular 8x8 blocks of a 128x128 tile
constexpr int WP_DOUBLES = WP_BLOCKS * 64;            // 8704 doubles = 68 KB
constexpr int CWIN = 2 * TILE;                        // window of the circulant column that one tile touches
constexpr int MISC_DOUBLES = 16 + CWIN;               // flag + reduction scratch, then the circulant window
static_assert(CWIN == THREADS, "one window entry per thread");
static_assert(TILE * LDA_NT + KC * LDB_NN <= STAGE_DOUBLES, "NN stage must fit");

typedef double Acc[2][16][2];

// optional phase profiler (DevBatch::prof != nullptr): thread 0 accumulates clock64 deltas per phase
#define PROF(idx)                                                                       \
  do {                                                                                  \
    if (d.prof && threadIdx.x == 0) {                                                   \
      long long t_ = clock64();                                                         \
      atomicAdd((unsigned long long*)&d.prof[idx], (unsigned long long)(t_ - tlast));   \
      tlast = t_;                                                                       \
    }                                                                                   \
  } while (0)

struct Lane {
  int warp, g, q;
  int rf[2];  // the two 8-row fragments of this warp: {w, 15-w} (balances triangular tiles)
  __device__ Lane() {
    warp = threadIdx.x >> 5;
    int l = threadIdx.x & 31;
    g = l >> 2;
    q = l & 3;
    rf[0] = warp;
    rf[1] = 15 - warp;
  }
};

// ----------------------------------------------------------------------------- thread-block cluster helpers
// A chain may be shared by the C CTAs of a cluster (sampler_dense.cu): they exchange data through the chain's global
// workspace (L2) and order it with the split-phase cluster barrier (release on arrive, acquire on wait, cluster scope).
__device__ __forceinline__ int cluster_ctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return (int)r;
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
  cluster_arrive();
  cluster_wait();
}

// which 8-column blocks each of the two row fragments of a warp computes (compile-time variants:
// predicated-off DMMAs still occupy the fp64 tensor pipe, so skipping must be a real branch)
enum Variant { V_FULL = 0, V_ROW0 = 1, V_ROW1 = 2, V_NONE = 3, V_DIAG_LO = 4, V_DIAG_HI = 5 };

__device__ __forceinline__ int rhs_variant(bool do0, bool do1) {
  return do0 ? (do1 ? V_FULL : V_ROW0) : (do1 ? V_ROW1 : V_NONE);
}

// packed W: block (bn, bk), bk <= bn, at ((bn*(bn+1))/2 + bk) * 64; element (r, c) of the block at
// r*8 + (c ^ swz(r)),  swz(r) = ((r >> 1) & 1) << 2  (conflict-free fragment reads in both forms)
__device__ __forceinline__ int wp_block(int bn, int bk) { return ((bn * (bn + 1)) / 2 + bk) * 64; }
__device__ __forceinline__ int wp_swz(int r) { return ((r >> 1) & 1) << 2; }

// ----------------------------------------------------------------------------- loaders
// NT chunk: A rows [0,128) x k0..k0+KC of a row-major panel (ld), same for B.
__device__ __forceinline__ void load_nt(double* stage, const double* A, const double* Bm, int ld, int k0) {
  double* sA = stage;
  double* sB = stage + TILE * LDA_NT;
  constexpr int PPR = KC / 2;  // 16-byte pieces per row
#pragma unroll
  for (int r = 0; r < TILE * PPR / THREADS; r++) {
    int idx = threadIdx.x + THREADS * r;
    int row = idx / PPR, c16 = idx % PPR;
    cp_async16(sA + row * LDA_NT + c16 * 2, A + (size_t)row * ld + k0 + c16 * 2);
    cp_async16(sB + row * LDA_NT + c16 * 2, Bm + (size_t)row * ld + k0 + c16 * 2);
  }
}
// NN chunk: A rows [0,128) x k0..k0+KC (lda);  B rows k0..k0+KC x 128 columns (ldb).
__device__ __forceinline__ void load_nn(double* stage, const double* A, int lda, const double* Bm, int ldb, int k0) {
  double* sA = stage;
  double* sB = stage + TILE * LDA_NT;
  constexpr int PPR = KC / 2;
#pragma unroll
  for (int r = 0; r < TILE * PPR / THREADS; r++) {
    int idx = threadIdx.x + THREADS * r;
    int row = idx / PPR, c16 = idx % PPR;
    cp_async16(sA + row * LDA_NT + c16 * 2, A + (size_t)row * lda + k0 + c16 * 2);
    int krow = idx >> 6, c = idx & 63;  // KC rows of 64 pieces
    cp_async16(sB + krow * LDB_NN + c * 2, Bm + (size_t)(k0 + krow) * ldb + c * 2);
  }
}
// packed W_j (contiguous WP_DOUBLES in global) -> smem
__device__ __forceinline__ void load_wp(double* sWp, const double* src) {
  for (int idx = threadIdx.x; idx < WP_DOUBLES / 2; idx += THREADS) cp_async16(sWp + idx * 2, src + idx * 2);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
}

// ----------------------------------------------------------------------------- chunk math
// acc -= A * B^T over one KC chunk (A, B both [row][k]); NB0/NB1 = column blocks done for rf0/rf1.
template <int NB0, int NB1, int KS0 = 0, int KS1 = KC / 4>
__device__ __forceinline__ void mma_nt(const double* stage, Acc& acc, const Lane& ln) {
  const double* sA = stage;
  const double* sB = stage + TILE * LDA_NT;
  const int r0 = (ln.rf[0] * 8 + ln.g) * LDA_NT + ln.q;
  const int r1 = (ln.rf[1] * 8 + ln.g) * LDA_NT + ln.q;
  const int rb = ln.g * LDA_NT + ln.q;
  constexpr int NBM = NB0 > NB1 ? NB0 : NB1;
#pragma unroll
  for (int ks = KS0; ks < KS1; ks++) {
    double a0 = 0, a1 = 0;
    if (NB0 > 0) a0 = -sA[r0 + ks * 4];
    if (NB1 > 0) a1 = -sA[r1 + ks * 4];
#pragma unroll
    for (int nb = 0; nb < NBM; nb++) {
      double b = sB[rb + nb * 8 * LDA_NT + ks * 4];
      if (nb < NB0) dmma(acc[0][nb][0], acc[0][nb][1], a0, b);
      if (nb < NB1) dmma(acc[1][nb][0], acc[1][nb][1], a1, b);
    }
  }
}
// acc += sign * A * B over one KC chunk (A [row][k], B [k][col]).
template <bool NEG, bool D0, bool D1, int KS0 = 0, int KS1 = KC / 4>
__device__ __forceinline__ void mma_nn(const double* stage, Acc& acc, const Lane& ln) {
  const double* sA = stage;
  const double* sB = stage + TILE * LDA_NT;
  const int r0 = (ln.rf[0] * 8 + ln.g) * LDA_NT + ln.q;
  const int r1 = (ln.rf[1] * 8 + ln.g) * LDA_NT + ln.q;
  const int rb = ln.q * LDB_NN + ln.g;
#pragma unroll
  for (int ks = KS0; ks < KS1; ks++) {
    double a0 = 0, a1 = 0;
    if (D0) a0 = NEG ? -sA[r0 + ks * 4] : sA[r0 + ks * 4];
    if (D1) a1 = NEG ? -sA[r1 + ks * 4] : sA[r1 + ks * 4];
#pragma unroll
    for (int nb = 0; nb < 16; nb++) {
      double b = sB[rb + ks * 4 * LDB_NN + nb * 8];
      if (D0) dmma(acc[0][nb][0], acc[0][nb][1], a0, b);
      if (D1) dmma(acc[1][nb][0], acc[1][nb][1], a1, b);
    }
  }
}

// ----------------------------------------------------------------------------- pipelined K loops
// acc -= A[0:128, k] * B[0:128, k]^T  for k in [0, nchunks*KC)
__device__ __forceinline__ void gemm_nt(double* pipe, Acc& acc, const Lane& ln, const double* A, const double* Bm,
                                        int ld, int nchunks, int variant) {
#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < nchunks) load_nt(pipe + s * STAGE_DOUBLES, A, Bm, ld, s * KC);
    cp_async_commit();
  }
  for (int c = 0; c < nchunks; c++) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    int nxt = c + STAGES - 1;
    const double* st = pipe + (c % STAGES) * STAGE_DOUBLES;
#if DGIBBS_SPLIT_CHUNK > 0
    // the tensor pipe starts on this chunk right after the barrier; the next chunk's copies are issued behind the
    // first quarter of its DMMAs (their memory clobber would otherwise hold the fragment loads back)
    constexpr int KH = DGIBBS_SPLIT_CHUNK;  // k-steps (of 4) issued before the copies
    switch (variant) {  // warp-uniform
      case V_FULL: mma_nt<16, 16, 0, KH>(st, acc, ln); break;
      case V_ROW0: mma_nt<16, 0, 0, KH>(st, acc, ln); break;
      case V_ROW1: mma_nt<0, 16, 0, KH>(st, acc, ln); break;
      case V_DIAG_LO: mma_nt<4, 16, 0, KH>(st, acc, ln); break;
      case V_DIAG_HI: mma_nt<8, 12, 0, KH>(st, acc, ln); break;
      default: break;
    }
    if (nxt < nchunks) load_nt(pipe + (nxt % STAGES) * STAGE_DOUBLES, A, Bm, ld, nxt * KC);
    cp_async_commit();
    switch (variant) {
      case V_FULL: mma_nt<16, 16, KH>(st, acc, ln); break;
      case V_ROW0: mma_nt<16, 0, KH>(st, acc, ln); break;
      case V_ROW1: mma_nt<0, 16, KH>(st, acc, ln); break;
      case V_DIAG_LO: mma_nt<4, 16, KH>(st, acc, ln); break;
      case V_DIAG_HI: mma_nt<8, 12, KH>(st, acc, ln); break;
      default: break;
    }
#else
    if (nxt < nchunks) load_nt(pipe + (nxt % STAGES) * STAGE_DOUBLES, A, Bm, ld, nxt * KC);
    cp_async_commit();
    switch (variant) {  // warp-uniform
      case V_FULL: mma_nt<16, 16>(st, acc, ln); break;
      case V_ROW0: mma_nt<16, 0>(st, acc, ln); break;
      case V_ROW1: mma_nt<0, 16>(st, acc, ln); break;
      case V_DIAG_LO: mma_nt<4, 16>(st, acc, ln); break;
      case V_DIAG_HI: mma_nt<8, 12>(st, acc, ln); break;
      default: break;
    }
#endif
  }
  cp_async_wait<0>();
  __syncthreads();
}
// acc (+/-)= A[0:128, kbase + k] * B[kbase + k, 0:128]  for k in [0, nchunks*KC)
template <bool NEG>
__device__ __forceinline__ void gemm_nn(double* pipe, Acc& acc, const Lane& ln, const double* A, int lda,
                                        const double* Bm, int ldb, int kbase, int nchunks, int variant) {
#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < nchunks) load_nn(pipe + s * STAGE_DOUBLES, A, lda, Bm, ldb, kbase + s * KC);
    cp_async_commit();
  }
  for (int c = 0; c < nchunks; c++) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    int nxt = c + STAGES - 1;
    const double* st = pipe + (c % STAGES) * STAGE_DOUBLES;
#if DGIBBS_SPLIT_CHUNK > 0
    constexpr int KH = DGIBBS_SPLIT_CHUNK;
    switch (variant) {
      case V_FULL: mma_nn<NEG, true, true, 0, KH>(st, acc, ln); break;
      case V_ROW0: mma_nn<NEG, true, false, 0, KH>(st, acc, ln); break;
      case V_ROW1: mma_nn<NEG, false, true, 0, KH>(st, acc, ln); break;
      default: break;
    }
    if (nxt < nchunks) load_nn(pipe + (nxt % STAGES) * STAGE_DOUBLES, A, lda, Bm, ldb, kbase + nxt * KC);
    cp_async_commit();
    switch (variant) {
      case V_FULL: mma_nn<NEG, true, true, KH>(st, acc, ln); break;
      case V_ROW0: mma_nn<NEG, true, false, KH>(st, acc, ln); break;
      case V_ROW1: mma_nn<NEG, false, true, KH>(st, acc, ln); break;
      default: break;
    }
#else
    if (nxt < nchunks) load_nn(pipe + (nxt % STAGES) * STAGE_DOUBLES, A, lda, Bm, ldb, kbase + nxt * KC);
    cp_async_commit();
    switch (variant) {
      case V_FULL: mma_nn<NEG, true, true>(st, acc, ln); break;
      case V_ROW0: mma_nn<NEG, true, false>(st, acc, ln); break;
      case V_ROW1: mma_nn<NEG, false, true>(st, acc, ln); break;
      default: break;
    }
#endif
  }
  cp_async_wait<0>();
  __syncthreads();
}

// ----------------------------------------------------------------------------- tile staging
__device__ __forceinline__ void acc_zero(Acc& acc) {
#pragma unroll
  for (int f = 0; f < 2; f++)
#pragma unroll
    for (int nb = 0; nb < 16; nb++) acc[f][nb][0] = acc[f][nb][1] = 0.0;
}
// registers -> staged tile sT[128][LDW]
__device__ __forceinline__ void stage_acc(const Acc& acc, double* sT, const Lane& ln, bool do0, bool do1) {
#pragma unroll
  for (int f = 0; f < 2; f++) {
    if (f == 0 ? !do0 : !do1) continue;
    double* row = sT + (ln.rf[f] * 8 + ln.g) * LDW + 2 * ln.q;
#pragma unroll
    for (int nb = 0; nb < 16; nb++) *reinterpret_cast<double2*>(row + nb * 8) = make_double2(acc[f][nb][0], acc[f][nb][1]);
  }
}
// Interior tiles (real mode, no padding) only touch c[(ti - tj) 128 - 127 .. (ti - tj) 128 + 127] of the circulant's first
// column: cwin[m] = c[(ti - tj) 128 - 128 + m] is fetched (one entry per thread, from L2) before the tile's GEMM and is
// visible after the GEMM's barriers.  Entry (r, c) of the tile is cwin[128 + r - c].
__device__ __forceinline__ void load_circ_window(double* cwin, const double* ct, int N, int dt) {
  int k = dt * TILE - TILE + (int)threadIdx.x;
  if (k < 0) k += N;
  cwin[threadIdx.x] = k < N ? ct[k] : 0.0;
}
// stage_acc fused with add_generated for an interior tile strictly below the diagonal: the circulant entry is added in
// registers on the way to shared memory
__device__ __forceinline__ void stage_acc_circ(const Acc& acc, double* sT, const Lane& ln, const double* cwin) {
#pragma unroll
  for (int f = 0; f < 2; f++) {
    const int r = ln.rf[f] * 8 + ln.g;
    double* row = sT + r * LDW + 2 * ln.q;
    const double* cp = cwin + TILE + r - 2 * ln.q;
#pragma unroll
    for (int nb = 0; nb < 16; nb++)
      *reinterpret_cast<double2*>(row + nb * 8) = make_double2(acc[f][nb][0] + cp[-nb * 8], acc[f][nb][1] + cp[-nb * 8 - 1]);
  }
}
// the same on the accumulator registers, before the tile is staged (lower-triangle blocks of this warp's two row
// fragments; one pass over shared memory less than add_generated_diag_win on the staged tile)
__device__ __forceinline__ void acc_add_diag_win(Acc& acc, const Lane& ln, const double* cwin, const double* Si, int R0) {
#pragma unroll
  for (int f = 0; f < 2; f++) {
    const int rb = ln.rf[f], r = rb * 8 + ln.g;
    const double* cp = cwin + TILE + r - 2 * ln.q;
#pragma unroll
    for (int nb = 0; nb < (f == 0 ? 8 : 16); nb++) {
      if (nb > rb) continue;
      double v0 = cp[-nb * 8], v1 = cp[-nb * 8 - 1];
      if (nb == rb) {
        if (2 * ln.q == ln.g) v0 += Si[R0 + r];
        if (2 * ln.q + 1 == ln.g) v1 += Si[R0 + r];
      }
      acc[f][nb][0] += v0;
      acc[f][nb][1] += v1;
    }
  }
}
// lower triangle of an interior diagonal tile: sT += circ window + diag(Si)
__device__ __forceinline__ void add_generated_diag_win(double* sT, const double* cwin, const double* Si, int R0) {
#pragma unroll 2
  for (int e = threadIdx.x; e < TILE * TILE; e += THREADS) {
    const int r = e >> 7, c = e & 127;
    if (c > r) continue;
    double v = cwin[TILE + r - c];
    if (r == c) v += Si[R0 + r];
    sT[r * LDW + c] += v;
  }
}
__device__ __forceinline__ void tile_store_global(const Acc& acc, double* base, int ld, const Lane& ln, bool do0,
                                                  bool do1) {
#pragma unroll
  for (int f = 0; f < 2; f++) {
    if (f == 0 ? !do0 : !do1) continue;
    double* row = base + (size_t)(ln.rf[f] * 8 + ln.g) * ld + 2 * ln.q;
#pragma unroll
    for (int nb = 0; nb < 16; nb++) *reinterpret_cast<double2*>(row + nb * 8) = make_double2(acc[f][nb][0], acc[f][nb][1]);
  }
}
__device__ __forceinline__ void tile_load_global(Acc& acc, const double* base, int ld, const Lane& ln) {
#pragma unroll
  for (int f = 0; f < 2; f++) {
    const double* row = base + (size_t)(ln.rf[f] * 8 + ln.g) * ld + 2 * ln.q;
#pragma unroll
    for (int nb = 0; nb < 16; nb++) {
      double2 v = *reinterpret_cast<const double2*>(row + nb * 8);
      acc[f][nb][0] = v.x;
      acc[f][nb][1] = v.y;
    }
  }
}
// sT += tile (ti, tj) of Ci = circ(c) + diag(Si) (identity in the padding).  Rolled, coalesced.
__device__ __forceinline__ void add_generated(double* sT, const double* ct, const double* Si, int N, int nsys, int cplx,
                                              int ti, int tj, bool lower_only) {
  const bool interior = !cplx && (ti + 1) * TILE <= nsys && (tj + 1) * TILE <= nsys;
  const int R0 = ti * TILE, C0 = tj * TILE;
#pragma unroll 2
  for (int e = threadIdx.x; e < TILE * TILE; e += THREADS) {
    int r = e >> 7, c = e & 127;
    if (lower_only && c > r) continue;
    int R = R0 + r, C = C0 + c;
    double v;
    if (interior) {
      int k = R - C;
      if (k < 0) k += N;
      v = ct[k];
      if (R == C) v += Si[R];
    } else if (R < nsys && C < nsys) {
      v = circ_entry(ct, N, cplx, R, C);
      if (R == C) v += Si[R];
    } else {
      v = (R == C) ? 1.0 : 0.0;
    }
    sT[r * LDW + c] += v;
  }
}
// sT[r0:r1] += src[r0:r1][0:128]  (row pitch ld in global)
__device__ __forceinline__ void add_global(double* sT, const double* src, int ld, int r1, int r0 = 0) {
  for (int e = r0 * (TILE / 2) + threadIdx.x; e < r1 * (TILE / 2); e += THREADS) {
    int r = e >> 6, c = (e & 63) * 2;
    double2 v = *reinterpret_cast<const double2*>(src + (size_t)r * ld + c);
    double2* p = reinterpret_cast<double2*>(sT + r * LDW + c);
    double2 o = *p;
    o.x += v.x;
    o.y += v.y;
    *p = o;
  }
}

// In place  sT <- sT * inv(L_jj)^T  by blocked forward substitution over the 16 column blocks:
//   X_p = (C_p - sum_{k<p} X_k L(p,k)^T) W8_p^T ,  W8_p = inv(L(p,p))  (8x8, well conditioned)
// sWp holds L_jj packed by 8x8 blocks with W8_p in the diagonal block slots.  Each warp owns its 16
// rows, so only __syncwarp is needed.  Same accuracy class as LAPACK's dtrsm (explicit inverses only
// at the 8x8 level): a full 128x128 inverse costs 20x in accuracy at cond(Ci) ~ 4e8.
template <bool D0, bool D1>
__device__ __forceinline__ void trisolve_nt_body(double* sT, const double* sWp, const Lane& ln) {
  double* r0 = sT + (ln.rf[0] * 8 + ln.g) * LDW;
  double* r1 = sT + (ln.rf[1] * 8 + ln.g) * LDW;
  const int q = ln.q, sw = wp_swz(ln.g);
  const int bo0 = ln.g * 8 + (q ^ sw), bo1 = ln.g * 8 + ((4 + q) ^ sw);
#pragma unroll 1
  for (int p = 0; p < 16; p++) {
    double c00 = 0, c01 = 0, c10 = 0, c11 = 0, e00 = 0, e01 = 0, e10 = 0, e11 = 0;
    if (D0) { c00 = r0[8 * p + 2 * q]; c01 = r0[8 * p + 2 * q + 1]; }
    if (D1) { c10 = r1[8 * p + 2 * q]; c11 = r1[8 * p + 2 * q + 1]; }
    const double* wb = sWp + wp_block(p, 0);
#pragma unroll 1
    for (int k = 0; k < p; k++) {
      double b0 = wb[k * 64 + bo0], b1 = wb[k * 64 + bo1];
      if (D0) { dmma(c00, c01, -r0[8 * k + q], b0); dmma(e00, e01, -r0[8 * k + 4 + q], b1); }
      if (D1) { dmma(c10, c11, -r1[8 * k + q], b0); dmma(e10, e11, -r1[8 * k + 4 + q], b1); }
    }
    __syncwarp();
    if (D0) { r0[8 * p + 2 * q] = c00 + e00; r0[8 * p + 2 * q + 1] = c01 + e01; }
    if (D1) { r1[8 * p + 2 * q] = c10 + e10; r1[8 * p + 2 * q + 1] = c11 + e11; }
    __syncwarp();
    const double* w8 = sWp + wp_block(p, p);
    double b0 = w8[bo0], b1 = w8[bo1];
    double o00 = 0, o01 = 0, o10 = 0, o11 = 0;
    if (D0) { dmma(o00, o01, r0[8 * p + q], b0); dmma(o00, o01, r0[8 * p + 4 + q], b1); }
    if (D1) { dmma(o10, o11, r1[8 * p + q], b0); dmma(o10, o11, r1[8 * p + 4 + q], b1); }
    __syncwarp();
    if (D0) { r0[8 * p + 2 * q] = o00; r0[8 * p + 2 * q + 1] = o01; }
    if (D1) { r1[8 * p + 2 * q] = o10; r1[8 * p + 2 * q + 1] = o11; }
    __syncwarp();
  }
}
// In place  sT <- sT * inv(L_jj)  by blocked backward substitution:
//   X_p = (Z_p - sum_{k>p} X_k L(k,p)) W8_p
template <bool D0, bool D1>
__device__ __forceinline__ void trisolve_nn_body(double* sT, const double* sWp, const Lane& ln) {
  double* r0 = sT + (ln.rf[0] * 8 + ln.g) * LDW;
  double* r1 = sT + (ln.rf[1] * 8 + ln.g) * LDW;
  const int q = ln.q;
  const int bo0 = q * 8 + (ln.g ^ wp_swz(q)), bo1 = (4 + q) * 8 + (ln.g ^ wp_swz(4 + q));
#pragma unroll 1
  for (int p = 15; p >= 0; p--) {
    double c00 = 0, c01 = 0, c10 = 0, c11 = 0, e00 = 0, e01 = 0, e10 = 0, e11 = 0;
    if (D0) { c00 = r0[8 * p + 2 * q]; c01 = r0[8 * p + 2 * q + 1]; }
    if (D1) { c10 = r1[8 * p + 2 * q]; c11 = r1[8 * p + 2 * q + 1]; }
#pragma unroll 1
    for (int k = p + 1; k < 16; k++) {
      const double* wb = sWp + wp_block(k, p);
      double b0 = wb[bo0], b1 = wb[bo1];
      if (D0) { dmma(c00, c01, -r0[8 * k + q], b0); dmma(e00, e01, -r0[8 * k + 4 + q], b1); }
      if (D1) { dmma(c10, c11, -r1[8 * k + q], b0); dmma(e10, e11, -r1[8 * k + 4 + q], b1); }
    }
    __syncwarp();
    if (D0) { r0[8 * p + 2 * q] = c00 + e00; r0[8 * p + 2 * q + 1] = c01 + e01; }
    if (D1) { r1[8 * p + 2 * q] = c10 + e10; r1[8 * p + 2 * q + 1] = c11 + e11; }
    __syncwarp();
    const double* w8 = sWp + wp_block(p, p);
    double b0 = w8[bo0], b1 = w8[bo1];
    double o00 = 0, o01 = 0, o10 = 0, o11 = 0;
    if (D0) { dmma(o00, o01, r0[8 * p + q], b0); dmma(o00, o01, r0[8 * p + 4 + q], b1); }
    if (D1) { dmma(o10, o11, r1[8 * p + q], b0); dmma(o10, o11, r1[8 * p + 4 + q], b1); }
    __syncwarp();
    if (D0) { r0[8 * p + 2 * q] = o00; r0[8 * p + 2 * q + 1] = o01; }
    if (D1) { r1[8 * p + 2 * q] = o10; r1[8 * p + 2 * q + 1] = o11; }
    __syncwarp();
  }
}
template <bool NN>
__device__ __forceinline__ void trisolve(double* sT, const double* sWp, const Lane& ln, int variant) {
  switch (variant) {
    case V_FULL: NN ? trisolve_nn_body<true, true>(sT, sWp, ln) : trisolve_nt_body<true, true>(sT, sWp, ln); break;
    case V_ROW0: NN ? trisolve_nn_body<true, false>(sT, sWp, ln) : trisolve_nt_body<true, false>(sT, sWp, ln); break;
    case V_ROW1: NN ? trisolve_nn_body<false, true>(sT, sWp, ln) : trisolve_nt_body<false, true>(sT, sWp, ln); break;
    default: break;
  }
}
// ----------------------------------------------------------------------------- triangular solve in registers
// The same solves on the ACCUMULATOR registers (the tile as the GEMM left it: lane (g, q) holds columns 2q, 2q+1 of row g
// of every 8x8 block), right-looking: once block column p is solved it is applied to the blocks behind it, which are
// independent of each other -- the tile makes no round trip through shared memory, there is no __syncwarp, and the only
// dependent chain is  quad shuffle -> 2 DMMA (W8_p) -> quad shuffle -> 2 DMMA  per block column.  The staged form above
// ran at ~10 cycles per DMMA (shared-memory read-modify-write of the tile, 30-cycle load latency in front of every
// k-step) against 4.3 in the GEMM loops (profiles/r02x_upd_probe.log).
#ifndef DGIBBS_TRISOLVE_REG
#define DGIBBS_TRISOLVE_REG 1
#endif
// accumulator layout -> A-operand layout of one 8x8 block: lane (g, q) gets entries (g, q) and (g, q + 4)
struct QuadXpose {
  int s0, s1;
  bool odd;
  __device__ QuadXpose() {
    const int lane = threadIdx.x & 31, q = lane & 3;
    s0 = (lane & ~3) | (q >> 1);
    s1 = s0 + 2;
    odd = q & 1;
  }
  __device__ __forceinline__ void operator()(double c0, double c1, double& a0, double& a1) const {
    const double u0 = __shfl_sync(0xffffffffu, c0, s0), u1 = __shfl_sync(0xffffffffu, c1, s0);
    const double v0 = __shfl_sync(0xffffffffu, c0, s1), v1 = __shfl_sync(0xffffffffu, c1, s1);
    a0 = odd ? u1 : u0;
    a1 = odd ? v1 : v0;
  }
};
// acc <- acc * inv(L_jj)^T   (forward; NN = false)    or    acc <- acc * inv(L_jj)   (backward; NN = true)
template <bool NN, bool D0, bool D1>
__device__ __forceinline__ void trisolve_reg_body(Acc& acc, const double* sWp, const Lane& ln) {
  const int g = ln.g, q = ln.q;
  const int bo0 = NN ? q * 8 + (g ^ wp_swz(q)) : g * 8 + (q ^ wp_swz(g));
  const int bo1 = NN ? (4 + q) * 8 + (g ^ wp_swz(4 + q)) : g * 8 + ((4 + q) ^ wp_swz(g));
  const QuadXpose xp;
#pragma unroll
  for (int pp = 0; pp < 16; pp++) {
    const int p = NN ? 15 - pp : pp;
    const double* w8 = sWp + wp_block(p, p);
    const double wb0 = w8[bo0], wb1 = w8[bo1];
    double a[2][2];
#pragma unroll
    for (int f = 0; f < 2; f++) {
      if (f == 0 ? !D0 : !D1) continue;
      double x0, x1;
      xp(acc[f][p][0], acc[f][p][1], x0, x1);
      double o0 = 0.0, o1 = 0.0, e0 = 0.0, e1 = 0.0;
      dmma(o0, o1, x0, wb0);
      dmma(e0, e1, x1, wb1);
      o0 += e0;
      o1 += e1;
      acc[f][p][0] = o0;
      acc[f][p][1] = o1;
      xp(o0, o1, a[f][0], a[f][1]);
      a[f][0] = -a[f][0];
      a[f][1] = -a[f][1];
    }
#pragma unroll
    for (int nn = 1; nn < 16; nn++) {
      if (nn > 15 - pp) continue;  // blocks behind column p: p+1 .. 15 (forward), p-1 .. 0 (backward)
      const int nb = NN ? p - nn : p + nn;
      const double* lb = sWp + (NN ? wp_block(p, nb) : wp_block(nb, p));
      const double b0 = lb[bo0], b1 = lb[bo1];
      if (D0) {
        dmma(acc[0][nb][0], acc[0][nb][1], a[0][0], b0);
        dmma(acc[0][nb][0], acc[0][nb][1], a[0][1], b1);
      }
      if (D1) {
        dmma(acc[1][nb][0], acc[1][nb][1], a[1][0], b0);
        dmma(acc[1][nb][0], acc[1][nb][1], a[1][1], b1);
      }
    }
  }
}
template <bool NN>
__device__ __forceinline__ void trisolve_reg(Acc& acc, const double* sWp, const Lane& ln, int variant) {
  switch (variant) {  // warp-uniform
    case V_FULL: trisolve_reg_body<NN, true, true>(acc, sWp, ln); break;
    case V_ROW0: trisolve_reg_body<NN, true, false>(acc, sWp, ln); break;
    case V_ROW1: trisolve_reg_body<NN, false, true>(acc, sWp, ln); break;
    default: break;
  }
}
// acc += the circulant entries of an interior tile strictly below the diagonal (stage_acc_circ without the staging)
__device__ __forceinline__ void acc_add_circ(Acc& acc, const Lane& ln, const double* cwin) {
#pragma unroll
  for (int f = 0; f < 2; f++) {
    const int r = ln.rf[f] * 8 + ln.g;
    const double* cp = cwin + TILE + r - 2 * ln.q;
#pragma unroll
    for (int nb = 0; nb < 16; nb++) {
      acc[f][nb][0] += cp[-nb * 8];
      acc[f][nb][1] += cp[-nb * 8 - 1];
    }
  }
}
// acc += tile at `base` (row pitch ld) for the fragments that are in use
__device__ __forceinline__ void acc_add_global(Acc& acc, const double* base, int ld, const Lane& ln, bool do0, bool do1) {
#pragma unroll
  for (int f = 0; f < 2; f++) {
    if (f == 0 ? !do0 : !do1) continue;
    const double* row = base + (size_t)(ln.rf[f] * 8 + ln.g) * ld + 2 * ln.q;
#pragma unroll
    for (int nb = 0; nb < 16; nb++) {
      const double2 v = *reinterpret_cast<const double2*>(row + nb * 8);
      acc[f][nb][0] += v.x;
      acc[f][nb][1] += v.y;
    }
  }
}

// staged tile rows [r0, r1) -> global (row pitch ld), coalesced
__device__ __forceinline__ void copy_tile_out(const double* sT, double* dst, int ld, int r1, int r0 = 0) {
  for (int e = r0 * (TILE / 2) + threadIdx.x; e < r1 * (TILE / 2); e += THREADS) {
    int r = e >> 6, c = (e & 63) * 2;
    *reinterpret_cast<double2*>(dst + (size_t)r * ld + c) = *reinterpret_cast<const double2*>(sT + r * LDW + c);
  }
}


// ----------------------------------------------------------------------------- masked inverse DFT as an FFT
// y[t, n] = sum_i u[t, 2i] cos(2 pi i n / N) - u[t, 2i+1] sin(2 pi i n / N) = Re sum_i U_i e^{+2 pi i (i n) / N}
// (the rows of fourier_matrix_r2c / _c2c, delay.py:1480-1510, 1549-1585, applied transposed) is an
// unnormalised inverse DFT of the zero-filled channel grid: O(N log N) per sample instead of the
// 4 n M flop of the dense projection, and it needs no projection table.  Stockham autosort, radix 4
// (+ one radix-2 stage), ping-pong between two shared-memory buffers, exact twiddles from a table.
__device__ __forceinline__ double2 tw_get(const double2* tw, int m, int N) {  // e^{2 pi i m / N}, 0 <= m < N
  const int half = N >> 1;
  double2 w = tw[m & (half - 1)];
  if (m & half) { w.x = -w.x; w.y = -w.y; }
  return w;
}
__device__ __forceinline__ double2 cmul(double2 a, double2 b) {
  return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ double2 cadd(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 csub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }
// 4-point inverse DFT (w4 = +i) in place: (b0, b1, b2, b3) -> (X0, X1, X2, X3)
__device__ __forceinline__ void dft4_inv(double2& b0, double2& b1, double2& b2, double2& b3) {
  const double2 s02 = cadd(b0, b2), d02 = csub(b0, b2), s13 = cadd(b1, b3), d13 = csub(b1, b3);
  b0 = cadd(s02, s13);
  b1 = make_double2(d02.x - d13.y, d02.y + d13.x);  // d02 + i d13
  b2 = csub(s02, s13);
  b3 = make_double2(d02.x + d13.y, d02.y - d13.x);  // d02 - i d13
}
// The buffer written by the first radix-8 stage (element 8 j + q from thread j) and read by the second one is
// XOR-swizzled inside every aligned group of 8 elements so that the 128-byte-strided stores of a warp spread over
// the banks; reads of consecutive elements stay a permutation of consecutive elements.
__device__ __forceinline__ int fft_swz(int p, bool on) { return on ? (p ^ ((p >> 3) & 7)) : p; }

// Inverse (sign +) DFT of x[0:N) (N = 2^m >= 8); returns the buffer that holds the result (x or y).
// Stockham autosort: radix-8 stages (8-point butterflies in registers: 4 passes over shared memory at N = 2048
// instead of the 6 of a radix-4 chain), then one radix-4 or radix-2 stage; exact twiddles from the table.
// All NT threads of the CTA must call it; the inputs must be visible (caller synchronises before).
// Batched form: G independent transforms in lockstep, transform g in x[g N .. (g+1) N) / y[g N ..): for small N one
// transform keeps only N / 8 threads busy per pass, G of them fill the CTA (per-element arithmetic unchanged).
template <int NT = THREADS>
static __device__ double2* fft_inverse_smem(double2* x, double2* y, const double2* tw, int N, int G = 1) {
  int Ns = 1;
  bool swz = false;  // is the current input buffer swizzled?
  const int e8 = N >> 3, l8 = __ffs(e8) - 1;
  while (Ns * 8 <= N) {
    const int s = N / (8 * Ns);
    const bool swz_out = Ns == 1 && N > 8;
    for (int jj = threadIdx.x; jj < G * e8; jj += NT) {
      const int j = jj & (e8 - 1), go = (jj >> l8) * N;
      const int k = j & (Ns - 1);
      double2 v[8];
#pragma unroll
      for (int r = 0; r < 8; r++) v[r] = x[go + fft_swz(j + r * e8, swz)];
      if (Ns > 1) {
        const int ks = k * s;  // r k s < N for r < 8
#pragma unroll
        for (int r = 1; r < 8; r++) v[r] = cmul(v[r], tw_get(tw, r * ks, N));
      }
      dft4_inv(v[0], v[2], v[4], v[6]);  // E_q
      dft4_inv(v[1], v[3], v[5], v[7]);  // O_q
      const double h = 0.70710678118654752440;
      const double2 o1 = make_double2((v[3].x - v[3].y) * h, (v[3].x + v[3].y) * h);   // w8 O_1
      const double2 o2 = make_double2(-v[5].y, v[5].x);                                  // i O_2
      const double2 o3 = make_double2(-(v[7].x + v[7].y) * h, (v[7].x - v[7].y) * h);  // w8^3 O_3
      const int j0 = ((j - k) << 3) + k;
      double2* yg = y + go;
      yg[fft_swz(j0, swz_out)] = cadd(v[0], v[1]);
      yg[fft_swz(j0 + Ns, swz_out)] = cadd(v[2], o1);
      yg[fft_swz(j0 + 2 * Ns, swz_out)] = cadd(v[4], o2);
      yg[fft_swz(j0 + 3 * Ns, swz_out)] = cadd(v[6], o3);
      yg[fft_swz(j0 + 4 * Ns, swz_out)] = csub(v[0], v[1]);
      yg[fft_swz(j0 + 5 * Ns, swz_out)] = csub(v[2], o1);
      yg[fft_swz(j0 + 6 * Ns, swz_out)] = csub(v[4], o2);
      yg[fft_swz(j0 + 7 * Ns, swz_out)] = csub(v[6], o3);
    }
    __syncthreads();
    double2* tmp = x; x = y; y = tmp;
    Ns <<= 3;
    swz = swz_out;
  }
  if (Ns * 4 == N) {  // one radix-4 stage
    const int quarter = N >> 2, s = 1, lq = __ffs(quarter) - 1;
    for (int jj = threadIdx.x; jj < G * quarter; jj += NT) {  // k = j (Ns = N / 4)
      const int j = jj & (quarter - 1), go = (jj >> lq) * N;
      double2 v0 = x[go + fft_swz(j, swz)], v1 = x[go + fft_swz(j + quarter, swz)], v2 = x[go + fft_swz(j + 2 * quarter, swz)],
              v3 = x[go + fft_swz(j + 3 * quarter, swz)];
      v1 = cmul(v1, tw_get(tw, j * s, N));
      v2 = cmul(v2, tw_get(tw, 2 * j * s, N));
      v3 = cmul(v3, tw_get(tw, 3 * j * s, N));
      dft4_inv(v0, v1, v2, v3);
      double2* yg = y + go;
      yg[j] = v0;
      yg[j + quarter] = v1;
      yg[j + 2 * quarter] = v2;
      yg[j + 3 * quarter] = v3;
    }
    __syncthreads();
    double2* tmp = x; x = y; y = tmp;
  } else if (Ns * 2 == N) {  // one radix-2 stage, twiddle stride 1
    const int half = N >> 1, lh = __ffs(half) - 1;
    for (int jj = threadIdx.x; jj < G * half; jj += NT) {
      const int j = jj & (half - 1), go = (jj >> lh) * N;
      const double2 v0 = x[go + fft_swz(j, swz)], v1 = cmul(x[go + fft_swz(j + half, swz)], tw[j]);
      y[go + j] = cadd(v0, v1);
      y[go + j + half] = csub(v0, v1);
    }
    __syncthreads();
    double2* tmp = x; x = y; y = tmp;
  }
  return x;
}

// u_i = h_i (z_i + w2_i) of sample t, channel i of the full grid (delay.py:1814,1817); 0 where masked
struct DrawCtx {
  const DevBatch* d;
  const Philox* ph;
  const int* cidx;
  const double* h;
  const double* zT;
  const double* w1;  // injected draws of this iteration (or null)
  const double* w2;
  int Tb, it, half;
};
// Device-RNG layout of the FFT path: one Philox block serves the sample pair (t & ~1, t | 1): words (x, y) give
// the two normals of the even sample, (z, w) those of the odd one (normal_quad, common.cuh: fp64 Box-Muller by default).
__device__ __forceinline__ void u_value2(const DrawCtx& c, int t0, int i, bool pair, double2& U0, double2& U1) {
  U0 = U1 = make_double2(0.0, 0.0);
  const int ci = c.cidx[i];
  if (ci < 0) return;
  double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
  const bool has1 = pair && t0 + 1 < c.Tb;
  if (c.d->rng_mode == 1) {
    unsigned long long blk = (unsigned long long)(t0 >> 1) * c.half + i;
    double e0, e1, e2, e3;
    normal_quad(*c.ph, (unsigned)blk, (unsigned)(blk >> 32), (unsigned)c.it, STREAM_W2, STREAM_W2_ODD, pair || !(t0 & 1),
                has1 || (!pair && (t0 & 1)), e0, e1, e2, e3);
    if (pair || !(t0 & 1)) { a0 = e0; a1 = e1; b0 = e2; b1 = e3; }
    else { a0 = e2; a1 = e3; }
  } else if (c.d->rng_mode == 0) {
    a0 = c.w2[(long long)(2 * ci) * c.Tb + t0];
    a1 = c.w2[(long long)(2 * ci + 1) * c.Tb + t0];
    if (has1) {
      b0 = c.w2[(long long)(2 * ci) * c.Tb + t0 + 1];
      b1 = c.w2[(long long)(2 * ci + 1) * c.Tb + t0 + 1];
    }
  }
  const double h0 = c.h[2 * i], h1 = c.h[2 * i + 1];
  const double2 z = *reinterpret_cast<const double2*>(c.zT + (size_t)t0 * c.d->Kpad + 2 * i);
  U0 = make_double2(h0 * (z.x + a0), h1 * (z.y + a1));
  if (has1) {
    const double2 y = *reinterpret_cast<const double2*>(c.zT + (size_t)(t0 + 1) * c.d->Kpad + 2 * i);
    U1 = make_double2(h0 * (y.x + b0), h1 * (y.y + b1));
  }
}
// the standard normals w1[t0, k], w1[t0, k+1] (k even) and, if `pair`, w1[t0+1, k], w1[t0+1, k+1]
__device__ __forceinline__ void w1_pair2(const DrawCtx& c, int t0, int k, int nsys, bool pair, double& a0, double& a1,
                                         double& b0, double& b1) {
  a0 = a1 = b0 = b1 = 0.0;
  if (c.d->rng_mode == 1) {
    unsigned long long blk = ((unsigned long long)(t0 >> 1) * nsys + k) >> 1;
    double e0, e1, e2, e3;
    normal_quad(*c.ph, (unsigned)blk, (unsigned)(blk >> 32), (unsigned)c.it, STREAM_W1, STREAM_W1_ODD, pair || !(t0 & 1),
                (pair && t0 + 1 < c.Tb) || (!pair && (t0 & 1)), e0, e1, e2, e3);
    if (pair || !(t0 & 1)) { a0 = e0; a1 = e1; b0 = e2; b1 = e3; }
    else { a0 = e2; a1 = e3; }
  } else if (c.d->rng_mode == 0) {
    a0 = c.w1[(long long)k * c.Tb + t0];
    a1 = c.w1[(long long)(k + 1) * c.Tb + t0];
    if (pair && t0 + 1 < c.Tb) {
      b0 = c.w1[(long long)k * c.Tb + t0 + 1];
      b1 = c.w1[(long long)(k + 1) * c.Tb + t0 + 1];
    }
  }
}

// Fused prepass + phase 0:  Rh[t, :] = sqrt(Si) w1[t] + F^T (h (z[t] + w2[t]))   for every sample row,
// zero in the padding rows / columns the factorisation sweeps over.
//   real mode    : two samples per complex FFT (their spectra are Hermitian-extended and combined as A + iB)
//   complex mode : one sample per FFT, output interleaved re/im
// A cluster of `nrank` CTAs shares the rows: CTA `rank` takes every nrank-th sample (pair).
// `fbuf` holds PIPE_DOUBLES / 2 complex entries: G = min(THREADS / (N / 8), what fits, 16) transforms run in lockstep
// (fft_inverse_smem's batched form), i.e. 2 G real samples per pass: at N = 256 the whole prepass of T = 32 samples is
// 2 passes instead of 16 (cfg1: 29 % of the kernel before).  Per-sample arithmetic does not depend on G.
static __device__ void project_rows_fft(const DevBatch& d, const DrawCtx& c, double* Rh, const double* Si, double2* fbuf,
                                 double2* tw_s, int rows, int rank = 0, int nrank = 1) {
  const int N = d.N, nsys = d.nsys, npad = d.npad, nch = d.nch;
  for (int m = threadIdx.x; m < N / 2; m += THREADS) tw_s[m] = d.twtab[m];
  const int step = d.cplx ? 1 : 2;
  int G = THREADS / max(1, N >> 3);
  G = max(1, min(min(G, 16), (PIPE_DOUBLES / 2) / (2 * N)));
  double2* bufA = fbuf;
  double2* bufB = fbuf + (size_t)G * N;
  const int hp = npad / 2;
  long long tp = clock64();
#define PROFX(idx)                                                                       \
  do {                                                                                   \
    if (d.prof && threadIdx.x == 0) {                                                    \
      long long t_ = clock64();                                                          \
      atomicAdd((unsigned long long*)&d.prof[idx], (unsigned long long)(t_ - tp));       \
      tp = t_;                                                                           \
    }                                                                                    \
  } while (0)
  for (int tb = rank * step * G; tb < rows; tb += nrank * step * G) {
    __syncthreads();  // previous epilogue finished with the buffers (and tw_s is loaded)
    PROFX(14);
    // ---- stage u of the samples tb + g step (g < G) that exist; rows beyond Tb are zero padding
    if (!d.cplx) {
      for (int e = threadIdx.x; e < G * nch; e += THREADS) {
        const int g = e / nch, i = e - g * nch, t0 = tb + g * step;
        if (t0 >= c.Tb || t0 >= rows) continue;
        double2* bA = bufA + (size_t)g * N;
        double2 U0, U1;
        u_value2(c, t0, i, true, U0, U1);
        if (i == 0 || i == N / 2) {
          bA[i] = make_double2(U0.x, U1.x);
        } else {
          bA[i] = make_double2(0.5 * (U0.x - U1.y), 0.5 * (U0.y + U1.x));       // (U0 + i U1) / 2
          bA[N - i] = make_double2(0.5 * (U0.x + U1.y), 0.5 * (U1.x - U0.y));   // (conj U0 + i conj U1) / 2
        }
      }
    } else {
      for (int e = threadIdx.x; e < G * N; e += THREADS) {
        const int g = e / N, i = e - g * N, t0 = tb + g;
        if (t0 >= c.Tb || t0 >= rows) continue;
        double2 U0, U1;
        u_value2(c, t0, i, false, U0, U1);
        bufA[(size_t)g * N + i] = U0;
      }
    }
    __syncthreads();
    PROFX(12);
    const double2* res = fft_inverse_smem(bufA, bufB, tw_s, N, G);
    PROFX(13);
    // ---- y = sqrt(Si) w1 + F^T u  (delay.py:1815-1817); zeros in the padding rows / columns
    for (int e = threadIdx.x; e < G * hp; e += THREADS) {
      const int g = e / hp, p = e - g * hp, k = 2 * p, t0 = tb + g * step;
      if (t0 >= rows) continue;
      const double2* rg = res + (size_t)g * N;
      if (!d.cplx) {
        const bool two = t0 + 1 < rows;
        double2 ya = make_double2(0.0, 0.0), yb = make_double2(0.0, 0.0);
        if (k < nsys && t0 < c.Tb) {
          const double2 r0 = rg[k], r1 = rg[k + 1];
          const double s0 = sqrt(Si[k]), s1 = sqrt(Si[k + 1]);
          double a0, a1, b0, b1;
          w1_pair2(c, t0, k, nsys, true, a0, a1, b0, b1);
          ya = make_double2(s0 * a0 + r0.x, s1 * a1 + r1.x);
          if (t0 + 1 < c.Tb) yb = make_double2(s0 * b0 + r0.y, s1 * b1 + r1.y);
        }
        *reinterpret_cast<double2*>(Rh + (size_t)t0 * npad + k) = ya;
        if (two) *reinterpret_cast<double2*>(Rh + (size_t)(t0 + 1) * npad + k) = yb;
      } else {
        double2 ya = make_double2(0.0, 0.0);
        if (k < nsys && t0 < c.Tb) {
          const double2 r0 = rg[p];
          double n0, n1, b0, b1;
          w1_pair2(c, t0, k, nsys, false, n0, n1, b0, b1);
          ya = make_double2(sqrt(Si[k]) * n0 + r0.x, sqrt(Si[k + 1]) * n1 + r0.y);
        }
        *reinterpret_cast<double2*>(Rh + (size_t)t0 * npad + k) = ya;
      }
    }
  }
  __syncthreads();
}

// ----------------------------------------------------------------------------- diagonal tile
#ifndef DGIBBS_FACTOR_LOOKAHEAD
#define DGIBBS_FACTOR_LOOKAHEAD 1
#endif
// 8x8 Cholesky of block (p, p) of the staged tile and its inverse W8_p, in registers: every lane of the calling warp
// computes the same values (identical-value stores).  W8_p goes to the diagonal slot p of sWp (bank-swizzled).
// Returns false on a non-positive pivot (same test as LAPACK dpotrf: ajj <= 0 or NaN).
__device__ __forceinline__ bool pivot_block(const double* sT, double* sWp, int p) {
  double a[8][8], w[8][8], rinv[8];
  bool ok = true;
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int c = 0; c <= i; c++) a[i][c] = sT[(8 * p + i) * LDW + 8 * p + c];
#pragma unroll
  for (int k = 0; k < 8; k++) {
    double dk = a[k][k];
    if (!(dk > 0.0)) ok = false;
    const double r = rsqrt(dk);  // one reciprocal square root instead of sqrt + division on the serial chain
    const double piv = dk * r;
    rinv[k] = r;
    a[k][k] = piv;
#pragma unroll
    for (int i = k + 1; i < 8; i++) a[i][k] *= r;
#pragma unroll
    for (int c = k + 1; c < 8; c++)
#pragma unroll
      for (int i = c; i < 8; i++) a[i][c] -= a[i][k] * a[c][k];
  }
#pragma unroll
  for (int c = 0; c < 8; c++) {
    w[c][c] = rinv[c];
#pragma unroll
    for (int i = c + 1; i < 8; i++) {
      double s = 0.0;
#pragma unroll
      for (int k = c; k < i; k++) s += a[i][k] * w[k][c];
      w[i][c] = -s * rinv[i];
    }
  }
  double* wpp = sWp + wp_block(p, p);
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int c = 0; c < 8; c++) wpp[i * 8 + (c ^ wp_swz(i))] = (c <= i) ? w[i][c] : 0.0;
  return ok;
}

// named barrier 1 of the CTA (barrier 0 is __syncthreads): warp 0 arrives, warps 1.. wait
__device__ __forceinline__ void bar1_arrive() { asm volatile("bar.arrive 1, %0;" ::"n"(THREADS) : "memory"); }
__device__ __forceinline__ void bar1_sync() { asm volatile("bar.sync 1, %0;" ::"n"(THREADS) : "memory"); }

#if DGIBBS_FACTOR_LOOKAHEAD
// In-place blocked Cholesky of the staged 128x128 tile (lower) by 8-column panels: LEFT-looking updates (a block is
// read and written once, its sum runs in registers) with a LOOK-AHEAD of one panel.  The 8x8 pivot factorisation is a
// serial chain of 8 reciprocal square roots that one warp computes (~1.2 k cycles), so warp 0 runs ahead on the pivot
// path while warps 1..7 prepare the next panel column.  Per panel p (W8_p known, column p updated but not solved):
//   warp 0      : L(p+1,p) = C(p+1,p) W8_p^T;  [arrive]  C(p+1,p+1) -= L(p+1,0:p+1) L(p+1,0:p+1)^T;
//                 W8_{p+1} = inv(chol(C(p+1,p+1)))
//   warps 1..7  : L(rb,p) = C(rb,p) W8_p^T (rb > p+1);  [wait]  C(rb,p+1) -= L(rb,0:p+1) L(p+1,0:p+1)^T  (rb > p+1)
// Earlier forms, cycles per tile in isolation (scratch/factor_probe.cu, profiles/r02x_upd_probe.log, r03i): left-looking
// without look-ahead (three CTA barriers and seven idle warps per pivot block; DGIBBS_FACTOR_LOOKAHEAD=0) 50.5 k;
// right-looking with look-ahead 49.5 k (every trailing block read-modify-written in shared memory per panel: ~10 cycles
// per DMMA).
// On exit sWp holds L_jj packed by 8x8 blocks (bank-swizzled; written as the blocks are solved) with W8_p in the
// diagonal slots.  Returns false on a non-positive pivot.
static __device__ bool factor_diag(double* sT, double* sWp, double* misc, const Lane& ln) {
  (void)misc;
  const int g = ln.g, q = ln.q;
  const int bo0 = g * 8 + (q ^ wp_swz(g)), bo1 = g * 8 + ((4 + q) ^ wp_swz(g));
  // block (rb, p) <- block (rb, p) W8_p^T
  // (and its packed copy in sWp: entry (n, k) of the block at n * 8 + (k ^ swz(n)))
  const int po = g * 8 + ((2 * q) ^ wp_swz(g));
  auto solve_block = [&](int rb, int p, double b0, double b1) {
    double* cp = sT + (rb * 8 + g) * LDW + 8 * p;
    const double a0 = cp[q], a1 = cp[4 + q];
    double o0 = 0.0, o1 = 0.0;
    dmma(o0, o1, a0, b0);
    dmma(o0, o1, a1, b1);
    __syncwarp();
    *reinterpret_cast<double2*>(cp + 2 * q) = make_double2(o0, o1);
    *reinterpret_cast<double2*>(sWp + wp_block(rb, p) + po) = make_double2(o0, o1);
  };
  // blocks (rbA, cb) and, if rbB < 16, (rbB, cb) -= L(rb, 0:nk) L(cb, 0:nk)^T; two accumulators per block (k4 halves)
  auto update_pair = [&](int rbA, int rbB, int cb, int nk) {
    const bool two = rbB < 16;
    double2* cpA = reinterpret_cast<double2*>(sT + (rbA * 8 + g) * LDW + 8 * cb + 2 * q);
    double2* cpB = reinterpret_cast<double2*>(sT + ((two ? rbB : rbA) * 8 + g) * LDW + 8 * cb + 2 * q);
    const double* apA = sT + (rbA * 8 + g) * LDW + q;
    const double* apB = sT + ((two ? rbB : rbA) * 8 + g) * LDW + q;
    const double* bp = sT + (cb * 8 + g) * LDW + q;
    double2 cA = *cpA, cB = *cpB;
    double eA0 = 0.0, eA1 = 0.0, eB0 = 0.0, eB1 = 0.0;
    if (two) {
#pragma unroll 2
      for (int k = 0; k < nk; k++) {
        const double b0 = bp[8 * k], b1 = bp[8 * k + 4];
        dmma(cA.x, cA.y, -apA[8 * k], b0);
        dmma(eA0, eA1, -apA[8 * k + 4], b1);
        dmma(cB.x, cB.y, -apB[8 * k], b0);
        dmma(eB0, eB1, -apB[8 * k + 4], b1);
      }
      *cpB = make_double2(cB.x + eB0, cB.y + eB1);
    } else {
#pragma unroll 2
      for (int k = 0; k < nk; k++) {
        dmma(cA.x, cA.y, -apA[8 * k], bp[8 * k]);
        dmma(eA0, eA1, -apA[8 * k + 4], bp[8 * k + 4]);
      }
    }
    *cpA = make_double2(cA.x + eA0, cA.y + eA1);
  };
  // the verdict on a pivot block travels through the panel's closing barrier (warp 0 runs ahead of the others: a flag
  // in shared memory could be read on both sides of its next store)
  int bad = 0;
  if (ln.warp == 0) bad = pivot_block(sT, sWp, 0) ? 0 : 1;
  if (__syncthreads_or(bad)) return false;
  for (int p = 0; p < 15; p++) {
    const double* w8 = sWp + wp_block(p, p);
    const double b0 = w8[bo0], b1 = w8[bo1];
    if (ln.warp == 0) {
      solve_block(p + 1, p, b0, b1);
      __threadfence_block();
      __syncwarp();
      bar1_arrive();  // L(p+1, p) is in place
      update_pair(p + 1, 16, p + 1, p + 1);
      __syncwarp();
      bad = pivot_block(sT, sWp, p + 1) ? 0 : 1;
    } else {
      const int rbA = p + 1 + ln.warp, rbB = rbA + NWARP - 1;  // at most 14 blocks below (p+1, p+1) for 7 warps
      if (rbA < 16) solve_block(rbA, p, b0, b1);
      if (rbB < 16) solve_block(rbB, p, b0, b1);
      bar1_sync();
      if (rbA < 16) update_pair(rbA, rbB, p + 1, p + 1);
    }
    if (__syncthreads_or(bad)) return false;
  }
  return true;
}
#else
// In-place blocked Cholesky of the staged 128x128 tile (lower) by 8-column panels, left-looking.  Per panel p:
//   S1  all warps : panel blocks  C(rb,p) -= L(rb,0:p) L(p,0:p)^T                      (DMMA)
//   S2  warp 0    : 8x8 Cholesky of C(p,p) and its inverse W8_p, in registers
//   S3  all warps : L(rb,p) = C(rb,p) W8_p^T  (rb > p)                                 (DMMA)
// On exit sWp holds L_jj packed by 8x8 blocks (bank-swizzled) with W8_p in the diagonal slots.
// Returns false on a non-positive pivot (same test as LAPACK dpotrf: ajj <= 0 or NaN).
static __device__ bool factor_diag(double* sT, double* sWp, double* misc, const Lane& ln) {
  int* flag = reinterpret_cast<int*>(misc);
  const int g = ln.g, q = ln.q;
  const int bo0 = g * 8 + (q ^ wp_swz(g)), bo1 = g * 8 + ((4 + q) ^ wp_swz(g));
  if (threadIdx.x == 0) *flag = 0;
  for (int p = 0; p < 16; p++) {
    // ---- S1
    for (int rb = p + ln.warp; rb < 16; rb += NWARP) {
      double* cp = sT + (rb * 8 + g) * LDW + 8 * p + 2 * q;
      double c0 = cp[0], c1 = cp[1], e0 = 0.0, e1 = 0.0;
      const double* ap = sT + (rb * 8 + g) * LDW + q;
      const double* bp = sT + (8 * p + g) * LDW + q;
#pragma unroll 1
      for (int k = 0; k < p; k++) {
        dmma(c0, c1, -ap[8 * k], bp[8 * k]);
        dmma(e0, e1, -ap[8 * k + 4], bp[8 * k + 4]);
      }
      cp[0] = c0 + e0;
      cp[1] = c1 + e1;
    }
    __syncthreads();
    // ---- S2 (every lane of warp 0 computes the same 8x8 factor; identical-value stores)
    if (ln.warp == 0) {
      if (!pivot_block(sT, sWp, p)) *flag = 1;
    }
    __syncthreads();
    if (*flag) return false;
    // ---- S3
    {
      const double* w8 = sWp + wp_block(p, p);
      const double b0 = w8[bo0], b1 = w8[bo1];
      for (int rb = p + 1 + ln.warp; rb < 16; rb += NWARP) {
        double* cp = sT + (rb * 8 + g) * LDW + 8 * p;
        double a0 = cp[q], a1 = cp[4 + q];
        double o0 = 0.0, o1 = 0.0;
        dmma(o0, o1, a0, b0);
        dmma(o0, o1, a1, b1);
        __syncwarp();
        cp[2 * q] = o0;
        cp[2 * q + 1] = o1;
      }
    }
    __syncthreads();
  }
  // pack the strictly-lower 8x8 blocks of L_jj next to the W8 diagonal blocks
  for (int e = threadIdx.x; e < TILE * TILE; e += THREADS) {
    int n = e >> 7, k = e & 127;
    int bn = n >> 3, bk = k >> 3;
    if (bk < bn) sWp[wp_block(bn, bk) + (n & 7) * 8 + ((k & 7) ^ wp_swz(n & 7))] = sT[n * LDW + k];
  }
  __syncthreads();
  return true;
}
#endif

// ----------------------------------------------------------------------------- diagonal branch
// Ci = D + b u u^T with D = diag(a + Si), u_k = (-1)^k: solve every RHS row in place by
// Sherman-Morrison.  Returns false when the system is not positive definite (some D_k <= 0 or
// 1 + b u^T D^-1 u <= 0), which is when the reference's Cholesky would fail.
// Every CTA of a cluster evaluates the (identical) positivity test; CTA `rank` solves every nrank-th group of rows.
static __device__ bool diag_solve_rows(double* Rh, const double* Si, double a, double bc, int nsys, int npad, int Tb,
                                double* misc, int rank = 0, int nrank = 1) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double ql = 0.0;
  bool pos = true;
  for (int k = threadIdx.x; k < nsys; k += THREADS) {
    double D = a + Si[k];
    if (!(D > 0.0)) pos = false;
    ql += 1.0 / D;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ql += __shfl_xor_sync(0xffffffffu, ql, o);
  __syncthreads();
  if (lane == 0) misc[2 + warp] = ql;
  int allpos = __syncthreads_and(pos ? 1 : 0);
  double q = 0.0;
#pragma unroll
  for (int w = 0; w < NWARP; w++) q += misc[2 + w];
  const double den = 1.0 + bc * q;
  if (!allpos || !(den > 0.0)) return false;
  const double beta = bc / den;
  for (int t = warp + rank * NWARP; t < Tb; t += NWARP * nrank) {
    double* row = Rh + (size_t)t * npad;
    double s = 0.0;
    if (bc != 0.0) {
      for (int k = lane; k < nsys; k += 32) s += ((k & 1) ? -row[k] : row[k]) / (a + Si[k]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    const double bs = beta * s;
    for (int k = lane; k < nsys; k += 32) row[k] = (row[k] - ((k & 1) ? -bs : bs)) / (a + Si[k]);
  }
  __syncthreads();
  return true;
}

// ----------------------------------------------------------------------------- factor + solve driver
// Workspace of one chain: L rows then the RHS rows (sample-major), pitch npad; the packed inverse
// diagonal blocks of every block column (for the backward solve); shared-memory scratch.
struct CholWS {
  double* Wk;    // [npad][npad]  L (lower), row-major
  double* Rh;    // [RT*TILE][npad]  right-hand sides, one row per sample
  double* Wpg;   // [nblk][WP_DOUBLES]
  double* sWp;   // shared: WP_DOUBLES
  double* pipe;  // shared: PIPE_DOUBLES
  double* misc;  // shared: MISC_DOUBLES
  int npad, nblk, RT, rhs_frags;
};

// The NT update of one tile, acc -= W[rowA : rowA+128, 0 : kcols] W[rowB : rowB+128, 0 : kcols]^T with W = the slot's
// workspace (the RHS rows follow the npad rows of L), is a policy: per-thread cp.async staging here, the TMA + mbarrier
// ring in tma_pipe.cuh (NtTma).
struct NtCpAsync {
  const double* Wk;
  double* pipe;
  int npad;
  __device__ __forceinline__ void operator()(Acc& acc, const Lane& ln, int rowA, int rowB, int kcols, int variant) {
    gemm_nt(pipe, acc, ln, Wk + (size_t)rowA * npad, Wk + (size_t)rowB * npad, npad, kcols / KC, variant);
  }
};

// Left-looking blocked Cholesky of the matrix whose 128x128 tiles `gen(sT, ti, tj, lower_only)` ADDS
// into a staged tile, with the RHS rows riding along as extra row tiles (forward solve).  Returns false
// on a non-positive pivot (LAPACK dpotrf semantics).  Every thread of the CTA must call it.  w.Rh must be
// w.Wk + npad * npad (one contiguous workspace).
template <class Gen, class NT>
__device__ __forceinline__ bool chol_factor_forward(const CholWS& w, const Lane& ln, const Gen& gen, NT& nt) {
  Acc acc;
  const int npad = w.npad, nblk = w.nblk;
  for (int j = 0; j < nblk; j++) {
    acc_zero(acc);
    nt(acc, ln, j * TILE, j * TILE, j * TILE, ln.warp < 4 ? V_DIAG_LO : V_DIAG_HI);
    stage_acc(acc, w.pipe, ln, true, true);
    __syncthreads();
    gen(w.pipe, j, j, true);
    __syncthreads();
    if (!factor_diag(w.pipe, w.sWp, w.misc, ln)) return false;
    for (int idx = threadIdx.x; idx < WP_DOUBLES / 2; idx += THREADS)
      reinterpret_cast<double2*>(w.Wpg + (size_t)j * WP_DOUBLES)[idx] = reinterpret_cast<const double2*>(w.sWp)[idx];
    __syncthreads();
    for (int i = j + 1; i < nblk + w.RT; i++) {
      bool do0 = true, do1 = true;
      int nrows = TILE;
      double* out;
      if (i < nblk) {
        out = w.Wk + (size_t)i * TILE * npad + j * TILE;
      } else {
        const int rt = i - nblk;
        do0 = rt * 16 + ln.rf[0] < w.rhs_frags;
        do1 = rt * 16 + ln.rf[1] < w.rhs_frags;
        nrows = min(TILE, max(0, w.rhs_frags * 8 - rt * TILE));
        out = w.Rh + (size_t)rt * TILE * npad + j * TILE;
      }
      const int var = rhs_variant(do0, do1);
      acc_zero(acc);
      nt(acc, ln, i * TILE, j * TILE, j * TILE, var);
#if DGIBBS_TRISOLVE_REG
      if (i >= nblk) {  // right-hand-side tile: solved on the accumulator registers
        acc_add_global(acc, out, npad, ln, do0, do1);
        trisolve_reg<false>(acc, w.sWp, ln, var);
        tile_store_global(acc, out, npad, ln, do0, do1);
        __syncthreads();
        continue;
      }
#endif
      stage_acc(acc, w.pipe, ln, do0, do1);
      __syncthreads();
      if (i < nblk) gen(w.pipe, i, j, false);
      else add_global(w.pipe, out, npad, nrows);
      __syncthreads();
      trisolve<false>(w.pipe, w.sWp, ln, var);
      __syncthreads();
      copy_tile_out(w.pipe, out, npad, nrows);
      __syncthreads();
    }
  }
  return true;
}
template <class Gen>
__device__ __forceinline__ bool chol_factor_forward(const CholWS& w, const Lane& ln, const Gen& gen) {
  NtCpAsync nt{w.Wk, w.pipe, w.npad};
  return chol_factor_forward(w, ln, gen, nt);
}

// Backward solve  x^T = z^T L^-1  on the RHS rows, in place.
__device__ __forceinline__ void chol_backward(const CholWS& w, const Lane& ln) {
  Acc acc;
  const int npad = w.npad, nblk = w.nblk;
  for (int rt = 0; rt < w.RT; rt++) {
    const bool do0 = rt * 16 + ln.rf[0] < w.rhs_frags, do1 = rt * 16 + ln.rf[1] < w.rhs_frags;
    const int var = rhs_variant(do0, do1);
    const int nrows = min(TILE, max(0, w.rhs_frags * 8 - rt * TILE));
    double* rrow = w.Rh + (size_t)rt * TILE * npad;
    for (int j = nblk - 1; j >= 0; j--) {
      double* xt = rrow + j * TILE;
      for (int idx = threadIdx.x; idx < WP_DOUBLES / 2; idx += THREADS)
        cp_async16(w.sWp + idx * 2, w.Wpg + (size_t)j * WP_DOUBLES + idx * 2);
      cp_async_commit();
      acc_zero(acc);
      gemm_nn<true>(w.pipe, acc, ln, rrow, npad, w.Wk + j * TILE, npad, (j + 1) * TILE, (nblk - 1 - j) * (TILE / KC), var);
#if DGIBBS_TRISOLVE_REG
      acc_add_global(acc, xt, npad, ln, do0, do1);
      trisolve_reg<true>(acc, w.sWp, ln, var);
      tile_store_global(acc, xt, npad, ln, do0, do1);
      __syncthreads();  // x_j is read by the next update; sWp is reloaded
      (void)nrows;
#else
      stage_acc(acc, w.pipe, ln, do0, do1);
      __syncthreads();
      add_global(w.pipe, xt, npad, nrows);
      __syncthreads();
      trisolve<true>(w.pipe, w.sWp, ln, var);
      __syncthreads();
      copy_tile_out(w.pipe, xt, npad, nrows);
      __syncthreads();
#endif
    }
  }
}

}  // namespace dg
