// Dense (delay-space) branch of the Gibbs sampler: one persistent CTA per baseline chain.
//
// Per iteration and baseline (reference: _draw_signal_sample_f + _draw_ps_sample,
// draco/analysis/delay.py:1778-1819, 1862-1880):
//   prepass   Si = 1/S ;  Y = sqrt(Si) w1 ;  u = h (z + w2)
//   phase 0   Y += u F                      masked inverse DFT as a DMMA GEMM against the shared table
//   factor    Ci = circ(c) + diag(Si) = L L^T, left-looking, 128-wide block columns, DMMA updates;
//             Ci is never stored: its tiles are generated from the circulant first column.
//             The RHS rows ride along as extra row tiles, so the forward solve is free.
//   backward  x = L^-T z  (DMMA), sample-major
//   draw      S = var_t(x) T / chi2
// The diagonal blocks hold W_j = inv(L_jj) so that every triangular solve is a DMMA multiply.
#include "common.cuh"
#include "internal.h"

namespace dg {

constexpr int STAGES = 4;
constexpr int STAGE_DOUBLES = 2 * TILE * LDA_NT;        // A[128][20] + B[128][20]  (NN form needs less)
constexpr int PIPE_DOUBLES = STAGES * STAGE_DOUBLES;    // 20480 doubles = 160 KB >= one 128x132 tile
static_assert(PIPE_DOUBLES >= TILE * LDW, "pipeline region must hold a full tile");
static_assert(TILE * LDA_NT + KC * LDB_NN <= STAGE_DOUBLES, "NN stage must fit");

typedef double Acc[2][16][2];

// optional phase profiler (DevBatch::prof != nullptr): thread 0 accumulates clock64 deltas per phase
#define PROF(idx)                                   \
  do {                                              \
    if (d.prof && threadIdx.x == 0) {               \
      long long t_ = clock64();                     \
      atomicAdd((unsigned long long*)&d.prof[idx], (unsigned long long)(t_ - tlast)); \
      tlast = t_;                                   \
    }                                               \
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

// ----------------------------------------------------------------------------- loaders
// NT chunk: A rows [0,128) x k0..k0+KC of a row-major panel (ld), same for B.
__device__ __forceinline__ void load_nt(double* stage, const double* A, const double* Bm, int ld, int k0) {
  double* sA = stage;
  double* sB = stage + TILE * LDA_NT;
#pragma unroll
  for (int r = 0; r < 4; r++) {
    int idx = threadIdx.x + THREADS * r;  // 0..1023
    int row = idx >> 3, c16 = idx & 7;
    cp_async16(sA + row * LDA_NT + c16 * 2, A + (size_t)row * ld + k0 + c16 * 2);
    cp_async16(sB + row * LDA_NT + c16 * 2, Bm + (size_t)row * ld + k0 + c16 * 2);
  }
}
// NN chunk: A rows [0,128) x k0..k0+KC (lda);  B rows k0..k0+KC x 128 columns (ldb).
__device__ __forceinline__ void load_nn(double* stage, const double* A, int lda, const double* Bm, int ldb, int k0) {
  double* sA = stage;
  double* sB = stage + TILE * LDA_NT;
#pragma unroll
  for (int r = 0; r < 4; r++) {
    int idx = threadIdx.x + THREADS * r;
    int row = idx >> 3, c16 = idx & 7;
    cp_async16(sA + row * LDA_NT + c16 * 2, A + (size_t)row * lda + k0 + c16 * 2);
    int krow = idx >> 6, c = idx & 63;
    cp_async16(sB + krow * LDB_NN + c * 2, Bm + (size_t)(k0 + krow) * ldb + c * 2);
  }
}
// full 128x128 tile (row pitch ld in global) -> smem pitch LDW
__device__ __forceinline__ void load_tile(double* sW, const double* src, int ld) {
#pragma unroll 4
  for (int r = 0; r < 32; r++) {
    int idx = threadIdx.x + THREADS * r;  // 0..8191
    int row = idx >> 6, c = idx & 63;
    cp_async16(sW + row * LDW + c * 2, src + (size_t)row * ld + c * 2);
  }
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
}

// ----------------------------------------------------------------------------- chunk math
// acc -= A * B^T over one KC chunk (A, B both [row][k]).  nbmax: last column block to touch.
__device__ __forceinline__ void mma_nt(const double* stage, Acc& acc, const Lane& ln, bool do0, bool do1,
                                       int nbmax0, int nbmax1) {
  const double* sA = stage;
  const double* sB = stage + TILE * LDA_NT;
  const int r0 = (ln.rf[0] * 8 + ln.g) * LDA_NT + ln.q;
  const int r1 = (ln.rf[1] * 8 + ln.g) * LDA_NT + ln.q;
  const int rb = ln.g * LDA_NT + ln.q;
#pragma unroll
  for (int ks = 0; ks < KC / 4; ks++) {
    double a0 = -sA[r0 + ks * 4], a1 = -sA[r1 + ks * 4];
#pragma unroll
    for (int nb = 0; nb < 16; nb++) {
      double b = sB[rb + nb * 8 * LDA_NT + ks * 4];
      if (do0 && nb <= nbmax0) dmma(acc[0][nb][0], acc[0][nb][1], a0, b);
      if (do1 && nb <= nbmax1) dmma(acc[1][nb][0], acc[1][nb][1], a1, b);
    }
  }
}
// acc += sign * A * B over one KC chunk (A [row][k], B [k][col]).
template <bool NEG>
__device__ __forceinline__ void mma_nn(const double* stage, Acc& acc, const Lane& ln, bool do0, bool do1) {
  const double* sA = stage;
  const double* sB = stage + TILE * LDA_NT;
  const int r0 = (ln.rf[0] * 8 + ln.g) * LDA_NT + ln.q;
  const int r1 = (ln.rf[1] * 8 + ln.g) * LDA_NT + ln.q;
  const int rb = ln.q * LDB_NN + ln.g;
#pragma unroll
  for (int ks = 0; ks < KC / 4; ks++) {
    double a0 = sA[r0 + ks * 4], a1 = sA[r1 + ks * 4];
    if (NEG) { a0 = -a0; a1 = -a1; }
#pragma unroll
    for (int nb = 0; nb < 16; nb++) {
      double b = sB[rb + ks * 4 * LDB_NN + nb * 8];
      if (do0) dmma(acc[0][nb][0], acc[0][nb][1], a0, b);
      if (do1) dmma(acc[1][nb][0], acc[1][nb][1], a1, b);
    }
  }
}

// C-fragment (8x8, cols 2q,2q+1) -> A-fragment (8x4) of its column half `half`: A[g][q] = C[g][4*half+q]
__device__ __forceinline__ double c_to_a(double c0, double c1, int half, const Lane& ln) {
  int src = (ln.g << 2) + half * 2 + (ln.q >> 1);
  double v0 = __shfl_sync(0xffffffffu, c0, src);
  double v1 = __shfl_sync(0xffffffffu, c1, src);
  return (ln.q & 1) ? v1 : v0;
}

// In place  acc <- acc * W^T  with W lower triangular in smem (W[n][k], pitch LDW): phase 2 of a tile.
__device__ __forceinline__ void trimul_nt(Acc& acc, const double* sW, const Lane& ln, bool do0, bool do1) {
#pragma unroll
  for (int nb = 15; nb >= 0; nb--) {
    double o00 = 0, o01 = 0, o10 = 0, o11 = 0;
#pragma unroll
    for (int kb = 0; kb <= nb; kb++) {
#pragma unroll
      for (int half = 0; half < 2; half++) {
        double b = sW[(nb * 8 + ln.g) * LDW + kb * 8 + half * 4 + ln.q];
        double a0 = c_to_a(acc[0][kb][0], acc[0][kb][1], half, ln);
        double a1 = c_to_a(acc[1][kb][0], acc[1][kb][1], half, ln);
        if (do0) dmma(o00, o01, a0, b);
        if (do1) dmma(o10, o11, a1, b);
      }
    }
    acc[0][nb][0] = o00; acc[0][nb][1] = o01;
    acc[1][nb][0] = o10; acc[1][nb][1] = o11;
  }
}
// In place  acc <- acc * W  (W[k][n] lower triangular: k >= n): backward substitution step.
__device__ __forceinline__ void trimul_nn(Acc& acc, const double* sW, const Lane& ln, bool do0, bool do1) {
#pragma unroll
  for (int nb = 0; nb < 16; nb++) {
    double o00 = 0, o01 = 0, o10 = 0, o11 = 0;
#pragma unroll
    for (int kb = nb; kb < 16; kb++) {
#pragma unroll
      for (int half = 0; half < 2; half++) {
        double b = sW[(kb * 8 + half * 4 + ln.q) * LDW + nb * 8 + ln.g];
        double a0 = c_to_a(acc[0][kb][0], acc[0][kb][1], half, ln);
        double a1 = c_to_a(acc[1][kb][0], acc[1][kb][1], half, ln);
        if (do0) dmma(o00, o01, a0, b);
        if (do1) dmma(o10, o11, a1, b);
      }
    }
    acc[0][nb][0] = o00; acc[0][nb][1] = o01;
    acc[1][nb][0] = o10; acc[1][nb][1] = o11;
  }
}

// ----------------------------------------------------------------------------- tile I/O
__device__ __forceinline__ void tile_load_global(Acc& acc, const double* base, int ld, const Lane& ln) {
#pragma unroll
  for (int f = 0; f < 2; f++)
#pragma unroll
    for (int nb = 0; nb < 16; nb++) {
      double2 v = *reinterpret_cast<const double2*>(base + (size_t)(ln.rf[f] * 8 + ln.g) * ld + nb * 8 + 2 * ln.q);
      acc[f][nb][0] = v.x;
      acc[f][nb][1] = v.y;
    }
}
__device__ __forceinline__ void tile_store_global(const Acc& acc, double* base, int ld, const Lane& ln, bool do0,
                                                  bool do1) {
#pragma unroll
  for (int f = 0; f < 2; f++) {
    if (f == 0 ? !do0 : !do1) continue;
#pragma unroll
    for (int nb = 0; nb < 16; nb++)
      *reinterpret_cast<double2*>(base + (size_t)(ln.rf[f] * 8 + ln.g) * ld + nb * 8 + 2 * ln.q) =
          make_double2(acc[f][nb][0], acc[f][nb][1]);
  }
}
// Tile (ti, tj) of Ci = circ(c) + diag(Si), identity in the padding.
__device__ __forceinline__ void tile_generate(Acc& acc, const double* ct, const double* Si, int N, int nsys, int cplx,
                                              int ti, int tj, const Lane& ln) {
#pragma unroll
  for (int f = 0; f < 2; f++) {
    int r = ti * TILE + ln.rf[f] * 8 + ln.g;
#pragma unroll
    for (int nb = 0; nb < 16; nb++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        int c = tj * TILE + nb * 8 + 2 * ln.q + e;
        double v;
        if (r < nsys && c < nsys) {
          v = circ_entry(ct, N, cplx, r, c);
          if (r == c) v += Si[r];
        } else {
          v = (r == c) ? 1.0 : 0.0;
        }
        acc[f][nb][e] = v;
      }
  }
}

// ----------------------------------------------------------------------------- pipelined K loops
// acc -= A[0:128, k] * B[0:128, k]^T  for k in [0, nchunks*KC)
__device__ __forceinline__ void gemm_nt(double* pipe, Acc& acc, const Lane& ln, const double* A, const double* Bm,
                                        int ld, int nchunks, bool do0, bool do1, int nbmax0, int nbmax1) {
#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < nchunks) load_nt(pipe + s * STAGE_DOUBLES, A, Bm, ld, s * KC);
    cp_async_commit();
  }
  for (int c = 0; c < nchunks; c++) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    int nxt = c + STAGES - 1;
    if (nxt < nchunks) load_nt(pipe + (nxt % STAGES) * STAGE_DOUBLES, A, Bm, ld, nxt * KC);
    cp_async_commit();
    mma_nt(pipe + (c % STAGES) * STAGE_DOUBLES, acc, ln, do0, do1, nbmax0, nbmax1);
  }
  cp_async_wait<0>();
  __syncthreads();
}
// acc (+/-)= A[0:128, k0 + k] * B[k0 + k, 0:128]  for k in [0, nchunks*KC)
template <bool NEG>
__device__ __forceinline__ void gemm_nn(double* pipe, Acc& acc, const Lane& ln, const double* A, int lda,
                                        const double* Bm, int ldb, int kbase, int nchunks, bool do0, bool do1) {
#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < nchunks) load_nn(pipe + s * STAGE_DOUBLES, A, lda, Bm, ldb, kbase + s * KC);
    cp_async_commit();
  }
  for (int c = 0; c < nchunks; c++) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    int nxt = c + STAGES - 1;
    if (nxt < nchunks) load_nn(pipe + (nxt % STAGES) * STAGE_DOUBLES, A, lda, Bm, ldb, kbase + nxt * KC);
    cp_async_commit();
    mma_nn<NEG>(pipe + (c % STAGES) * STAGE_DOUBLES, acc, ln, do0, do1);
  }
  cp_async_wait<0>();
  __syncthreads();
}

// ----------------------------------------------------------------------------- diagonal tile
// In-place Cholesky of the 128x128 tile in smem (lower).  Returns false on a non-positive pivot
// (same test as LAPACK dpotrf: ajj <= 0 or NaN).  scol: 128 doubles of scratch.
__device__ bool potrf_tile(double* sD, double* scol) {
  const int t = threadIdx.x;
  for (int k = 0; k < TILE; k++) {
    __syncthreads();
    double d = sD[k * LDW + k];
    if (!(d > 0.0)) return false;  // uniform across the CTA
    double ajj = sqrt(d);
    double rinv = 1.0 / ajj;
    int i = k + 1 + (t >> 1);
    double lik = 0.0;
    if (i < TILE) {
      lik = sD[i * LDW + k] * rinv;
      if ((t & 1) == 0) scol[i] = lik;
    }
    __syncthreads();
    if (t == 0) sD[k * LDW + k] = ajj;
    if (i < TILE) {
      if ((t & 1) == 0) sD[i * LDW + k] = lik;
      for (int c = k + 1 + (t & 1); c <= i; c += 2) sD[i * LDW + c] -= lik * scol[c];
    }
  }
  __syncthreads();
  return true;
}
// In-place inverse of the lower-triangular tile (dtrti2, 'L', 'N').
__device__ void trtri_tile(double* sD) {
  const int t = threadIdx.x;
  for (int j = TILE - 1; j >= 0; j--) {
    __syncthreads();
    double ajj = 1.0 / sD[j * LDW + j];
    int i = j + 1 + (t >> 1);
    double sum = 0.0;
    if (i < TILE)
      for (int k = j + 1 + (t & 1); k <= i; k += 2) sum += sD[i * LDW + k] * sD[k * LDW + j];
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    __syncthreads();
    if (i < TILE && (t & 1) == 0) sD[i * LDW + j] = -ajj * sum;
    if (t == 0) sD[j * LDW + j] = ajj;
  }
  __syncthreads();
}

// ----------------------------------------------------------------------------- the sampler
// Shared memory:  [ctab: nsys (optional)] [pipe: PIPE_DOUBLES] [scol: TILE] [flag]
__global__ void __launch_bounds__(THREADS, 1) dense_sampler_kernel(DevBatch d, int it_begin, int it_end) {
  extern __shared__ __align__(16) double smem[];
  const Lane ln;
  const int slot = blockIdx.x;
  const int N = d.N, nsys = d.nsys, npad = d.npad, nblk = d.nblk, Kpad = d.Kpad, RT = d.RT;
  double* s_ct = smem;
  double* pipe = smem + (d.ctab_smem ? nsys : 0);
  double* scol = pipe + PIPE_DOUBLES;

  const size_t wrows = (size_t)npad + (size_t)RT * TILE;
  double* Wk = d.work + (size_t)slot * wrows * npad;   // L rows then RHS rows, pitch npad
  double* Rh = Wk + (size_t)npad * npad;               // RHS rows (sample-major)
  double* uT = d.uT + (size_t)slot * RT * TILE * Kpad;
  double* Si = d.Si + (size_t)slot * npad;

  long long tlast = clock64();
  for (int b = slot; b < d.B; b += gridDim.x) {
    if (d.failed[b]) continue;
    const int Tb = d.T[b], Mb = d.M[b];
    const double* ctg = d.ctab + (size_t)b * nsys;
    const double* ct = ctg;
    if (d.ctab_smem) {
      __syncthreads();
      for (int k = threadIdx.x; k < nsys; k += THREADS) s_ct[k] = ctg[k];
      ct = s_ct;
    }
    const double* h = d.h + (size_t)b * Kpad;
    const double* zT = d.zT + (size_t)b * d.Tmax * Kpad;
    const int* cidx = d.cidx + (size_t)b * d.nch;
    double* S = d.S + (size_t)b * N;
    const Philox ph(mix_key(d.seed, d.gid ? (unsigned long long)d.gid[b] : (unsigned long long)b));
    const int rhs_frags = (Tb + 7) >> 3;  // valid 8-row fragments of the RHS

    bool alive = true;
    for (int it = it_begin; it < it_end && alive; it++) {
      // ---------------- prepass
      __syncthreads();
      for (int k = threadIdx.x; k < npad; k += THREADS) {
        double si = 0.0;
        if (k < nsys) {
          double s = S[d.cplx ? (k >> 1) : k];
          si = (s != 0.0) ? 1.0 / s : 0.0;   // invert_no_zero, delay.py:1785
          if (d.cplx) si *= 2.0;             // delay.py:1794
        }
        Si[k] = si;
      }
      __syncthreads();
      {  // Y = sqrt(Si) * w1  (two adjacent delays per thread), zero in the padding
        const int half = npad >> 1;
        const long long tot = (long long)RT * TILE * half;
        for (long long e = threadIdx.x; e < tot; e += THREADS) {
          int t = (int)(e / half), k = (int)(e % half) * 2;
          double y0 = 0.0, y1 = 0.0;
          if (t < Tb && k < nsys && d.rng_mode != 2) {
            double n0, n1;
            if (d.rng_mode == 1) {
              unsigned long long blk = ((unsigned long long)t * nsys + k) >> 1;
              normal_pair(ph((unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_W1), n0, n1);
            } else {
              const double* w1 = d.w1 + d.w1_off[b] + (long long)it * nsys * Tb;
              n0 = w1[(long long)k * Tb + t];
              n1 = w1[(long long)(k + 1) * Tb + t];
            }
            y0 = sqrt(Si[k]) * n0;       // delay.py:1815
            y1 = sqrt(Si[k + 1]) * n1;
          }
          *reinterpret_cast<double2*>(Rh + (size_t)t * npad + k) = make_double2(y0, y1);
        }
      }
      {  // u = h * (z + w2)  on the full channel grid (delay.py:1814,1817), zero where masked/padded
        const int half = Kpad >> 1;
        const long long tot = (long long)RT * TILE * half;
        for (long long e = threadIdx.x; e < tot; e += THREADS) {
          int t = (int)(e / half), i = (int)(e % half);
          double u0 = 0.0, u1 = 0.0;
          if (t < Tb && i < d.nch) {
            int ci = cidx[i];
            if (ci >= 0) {
              double n0 = 0.0, n1 = 0.0;
              if (d.rng_mode == 1) {
                unsigned long long blk = (unsigned long long)t * half + i;
                normal_pair(ph((unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_W2), n0, n1);
              } else if (d.rng_mode == 0) {
                const double* w2 = d.w2 + d.w2_off[b] + (long long)it * 2 * Mb * Tb;
                n0 = w2[(long long)(2 * ci) * Tb + t];
                n1 = w2[(long long)(2 * ci + 1) * Tb + t];
              }
              double2 z = *reinterpret_cast<const double2*>(zT + (size_t)t * Kpad + 2 * i);
              u0 = h[2 * i] * (z.x + n0);
              u1 = h[2 * i + 1] * (z.y + n1);
            }
          }
          *reinterpret_cast<double2*>(uT + (size_t)t * Kpad + 2 * i) = make_double2(u0, u1);
        }
      }
      __syncthreads();
      PROF(0);

      Acc acc;
      // ---------------- phase 0: Y += u F
      for (int rt = 0; rt < RT; rt++) {
        const bool do0 = rt * 16 + ln.rf[0] < rhs_frags, do1 = rt * 16 + ln.rf[1] < rhs_frags;
        for (int cb = 0; cb < nblk; cb++) {
          double* ytile = Rh + (size_t)rt * TILE * npad + cb * TILE;
          tile_load_global(acc, ytile, npad, ln);
          gemm_nn<false>(pipe, acc, ln, uT + (size_t)rt * TILE * Kpad, Kpad, d.Ftab + cb * TILE, npad, 0, Kpad / KC,
                         do0, do1);
          tile_store_global(acc, ytile, npad, ln, do0, do1);
        }
      }
      __syncthreads();
      PROF(1);

      // ---------------- factor (+ forward solve on the RHS row tiles)
      for (int j = 0; j < nblk && alive; j++) {
        const double* Bp = Wk + (size_t)j * TILE * npad;
        // diagonal tile
        tile_generate(acc, ct, Si, N, nsys, d.cplx, j, j, ln);
        gemm_nt(pipe, acc, ln, Bp, Bp, npad, j * (TILE / KC), true, true, ln.rf[0], ln.rf[1]);
        PROF(2);
#pragma unroll
        for (int f = 0; f < 2; f++)
#pragma unroll
          for (int nb = 0; nb < 16; nb++)
            if (nb <= ln.rf[f]) {
              double* p = pipe + (ln.rf[f] * 8 + ln.g) * LDW + nb * 8 + 2 * ln.q;
              p[0] = acc[f][nb][0];
              p[1] = acc[f][nb][1];
            }
        alive = potrf_tile(pipe, scol);
        if (!alive) break;
        PROF(3);
        trtri_tile(pipe);
        {  // W_j = inv(L_jj) into the diagonal block, zeros above the diagonal
          double* dst = Wk + (size_t)j * TILE * npad + j * TILE;
          for (int idx = threadIdx.x; idx < TILE * TILE / 2; idx += THREADS) {
            int r = idx >> 6, c = (idx & 63) * 2;
            double v0 = (c <= r) ? pipe[r * LDW + c] : 0.0;
            double v1 = (c + 1 <= r) ? pipe[r * LDW + c + 1] : 0.0;
            *reinterpret_cast<double2*>(dst + (size_t)r * npad + c) = make_double2(v0, v1);
          }
        }
        __syncthreads();
        PROF(4);
        // tiles below the diagonal, then the RHS row tiles
        for (int i = j + 1; i < nblk + RT; i++) {
          bool do0 = true, do1 = true;
          double* out;
          if (i < nblk) {
            tile_generate(acc, ct, Si, N, nsys, d.cplx, i, j, ln);
            out = Wk + (size_t)i * TILE * npad + j * TILE;
          } else {
            int rt = i - nblk;
            do0 = rt * 16 + ln.rf[0] < rhs_frags;
            do1 = rt * 16 + ln.rf[1] < rhs_frags;
            out = Rh + (size_t)rt * TILE * npad + j * TILE;
            tile_load_global(acc, out, npad, ln);
          }
          PROF(5);
          gemm_nt(pipe, acc, ln, Wk + (size_t)i * TILE * npad, Bp, npad, j * (TILE / KC), do0, do1, 15, 15);
          PROF(6);
          load_tile(pipe, Wk + (size_t)j * TILE * npad + j * TILE, npad);
          PROF(7);
          trimul_nt(acc, pipe, ln, do0, do1);
          tile_store_global(acc, out, npad, ln, do0, do1);
          __syncthreads();
          PROF(8);
        }
      }
      if (!alive) {
        __syncthreads();
        if (threadIdx.x == 0) {
          d.failed[b] = 1;
          d.n_done[b] = it;
        }
        break;
      }

      // ---------------- backward solve  x^T = z^T L^-1
      for (int rt = 0; rt < RT; rt++) {
        const bool do0 = rt * 16 + ln.rf[0] < rhs_frags, do1 = rt * 16 + ln.rf[1] < rhs_frags;
        double* rrow = Rh + (size_t)rt * TILE * npad;
        for (int j = nblk - 1; j >= 0; j--) {
          double* xt = rrow + j * TILE;
          tile_load_global(acc, xt, npad, ln);
          gemm_nn<true>(pipe, acc, ln, rrow, npad, Wk + j * TILE, npad, (j + 1) * TILE, (nblk - 1 - j) * (TILE / KC),
                        do0, do1);
          PROF(9);
          load_tile(pipe, Wk + (size_t)j * TILE * npad + j * TILE, npad);
          trimul_nn(acc, pipe, ln, do0, do1);
          tile_store_global(acc, xt, npad, ln, do0, do1);
          __syncthreads();
          PROF(10);
        }
      }

      if (d.wiener) {
        double* xo = d.xout + (size_t)d.xoff[b] * nsys;
        for (long long e = threadIdx.x; e < (long long)Tb * nsys; e += THREADS) {
          int t = (int)(e / nsys), k = (int)(e % nsys);
          xo[e] = Rh[(size_t)t * npad + k];
        }
        if (threadIdx.x == 0) d.n_done[b] = 1;
        break;
      }

      // ---------------- power-spectrum draw (delay.py:1862-1880)
      for (int k = threadIdx.x; k < N; k += THREADS) {
        double var = 0.0;
        const int ncomp = d.cplx ? 2 : 1;
        for (int comp = 0; comp < ncomp; comp++) {
          const double* col = Rh + (d.cplx ? 2 * k + comp : k);
          double mean = 0.0;
          for (int t = 0; t < Tb; t++) mean += col[(size_t)t * npad];
          mean /= (double)Tb;
          double ss = 0.0;
          for (int t = 0; t < Tb; t++) {
            double dv = col[(size_t)t * npad] - mean;
            ss += dv * dv;
          }
          var += ss / (double)Tb;
        }
        double chi2;
        if (d.rng_mode == 1) {
          // chi^2_T = sum of T squared normals; pairs collapse to -2 ln(u)
          double acc2 = 0.0;
          const unsigned long long base = (unsigned long long)k * ((Tb + 3) / 4 + 1);
          for (int p = 0; p < Tb / 2; p += 2) {
            unsigned long long blk = base + (p >> 1);
            uint4 r = ph((unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_CHI2);
            acc2 += -2.0 * log(u01(r.x, r.y));
            if (p + 1 < Tb / 2) acc2 += -2.0 * log(u01(r.z, r.w));
          }
          if (Tb & 1) {
            unsigned long long blk = base + (Tb + 3) / 4;
            double n0, n1;
            normal_pair(ph((unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_CHI2), n0, n1);
            acc2 += n0 * n0;
          }
          chi2 = acc2;
        } else {
          chi2 = d.chi2[((size_t)b * d.niter + it) * N + k];
        }
        double snew = var * (double)Tb / chi2;
        S[k] = snew;
        d.samples[((size_t)it * d.B + b) * N + k] = snew;
      }
      if (threadIdx.x == 0) d.n_done[b] = it + 1;
      __syncthreads();
      PROF(11);
    }
  }
}

size_t dense_smem_bytes(const DevBatch& d) {
  return sizeof(double) * ((d.ctab_smem ? (size_t)d.nsys : 0) + PIPE_DOUBLES + TILE) + 16;
}

cudaError_t dense_configure(const DevBatch& d) {
  return cudaFuncSetAttribute(dense_sampler_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)dense_smem_bytes(d));
}

cudaError_t launch_dense_sampler(const DevBatch& d, int it_begin, int it_end, cudaStream_t st) {
  dense_sampler_kernel<<<d.nslots, THREADS, dense_smem_bytes(d), st>>>(d, it_begin, it_end);
  return cudaGetLastError();
}

}  // namespace dg
