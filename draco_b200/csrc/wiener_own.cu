// Own-kernel path of the Wiener delay-transform operator: ConstructWienerDelayTransform.process,
// draco/analysis/powerspec.py:288-316, per pixel
//     A = H G H^H + C o r r^T ,   cho_factor(A) ,   D = S F^H H^H A^-1 window
// on the fp64 tensor pipe (DMMA) through the tile machinery of the Gibbs sampler (dense_core.cuh) instead of cuBLAS
// ZGEMM / cuSOLVER ZPOTRF / cuBLAS ZTRSM (wiener_op.cu, solvers 1 and 2).
//
// Layout.  Every complex matrix is row-major with interleaved (re, im) doubles and a real row pitch LD = 2 F rounded
// up to 128; a host [F, F] array IS its transpose in the column-major convention of wiener_op.cu, so with
//   Pr[j][k] = P(k, j) ,  Gr[k][i] = G'(i, k)   (both = the arrays as they lie in memory)
// the three contractions are REAL tile GEMMs of exactly the complex flop count (8 M N K):
//   T1r = Pr  Gext            NN  [F x 2F] [2F x 2F]   Gext = the real 2x2 embedding of Gr
//   Ar  = T1r Pext^T          NT  [F x 2F] [2F x 2F]^T Pext rows 2i / 2i+1 = (re, im) / (-im, re) of Pr[i]: Ar = T1r Pr^H
//   Yr  = Fpr Pext^T          NT  [nd x 2F] ...        Yr = Fpr Pr^H = memory of R^T = F^H H^H
// (only the tiles of Ar on and below the diagonal are computed: A is Hermitian).  Ar[j][i] = A'(i, j), so the system
// solved for row d of the output is conj(Ar) x = Yr[d]: its real embedding E[2i+p][2j+q] (E symmetric positive
// definite, 2F x 2F, identity in the padding) goes through the left-looking DMMA Cholesky of the sampler with the nd
// right-hand sides riding along as row tiles, one persistent CTA per pixel; the solution rows ARE the interleaved
// complex rows D'[d][f] the container wants, and prior, window, mask and the complex64 rounding are fused into the
// copy-out.
#include "wiener_own.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "dense_core.cuh"
#include "ozaki_gemm.cuh"
#include "tma_host.h"
#include "tma_pipe.cuh"

namespace dgw {
using namespace dg;

static int round_up(int x, int m) { return (x + m - 1) / m * m; }

OwnGeom own_geom(int F, int nd, int sm_count) {
  OwnGeom g{};
  g.F = F;
  g.nd = nd;
  g.LD = round_up(2 * F, TILE);
  g.FR = round_up(F, TILE);
  g.NDR = round_up(nd, TILE);
  g.nblk = g.LD / TILE;
  g.RT = g.NDR / TILE;
  g.nslots = sm_count;
  g.work_stride = ((size_t)g.LD + (size_t)g.RT * TILE) * g.LD;
  return g;
}

OwnSizes own_sizes(const OwnGeom& g) {
  OwnSizes s{};
  s.Pr = (size_t)g.FR * g.LD;
  s.Pext = (size_t)g.LD * g.LD;
  s.T1r = (size_t)g.FR * g.LD;
  s.Ar = (size_t)g.FR * g.LD;
  s.Yr = (size_t)g.NDR * g.LD;
  s.Gext = (size_t)g.LD * g.LD;
  s.Fpr = (size_t)g.NDR * g.LD;
  s.work = (size_t)g.nslots * g.work_stride;
  s.Wp = (size_t)g.nslots * g.nblk * WP_DOUBLES;
  s.qP = oz::slice_bytes(g.FR, g.LD);
  s.qPe = oz::slice_bytes(g.LD, g.LD);
  s.eP = oz::exp_ints(g.FR, g.LD);
  s.ePe = oz::exp_ints(g.LD, g.LD);
  s.qG = oz::slice_bytes(g.LD, g.LD);
  s.qF = oz::slice_bytes(g.NDR, g.LD);
  s.eG = oz::exp_ints(g.LD, g.LD);
  s.eF = oz::exp_ints(g.NDR, g.LD);
  return s;
}

__device__ __forceinline__ double2 load_c(const void* p, size_t i, int is_complex) {
  if (is_complex) return reinterpret_cast<const double2*>(p)[i];
  return make_double2(reinterpret_cast<const double*>(p)[i], 0.0);
}

// Gext = real embedding of Gr[k][i] = FSFT'(i, k) b_i b_k (powerspec.py:266, 288): rows 2k / 2k+1 = (re, im) / (-im, re)
__global__ void own_embed_g_kernel(const double2* __restrict__ FSFT, const double* __restrict__ bp, double* __restrict__ Gext, int F,
                                   int LD) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.y;
  if (i >= F) return;
  const double s = sqrt(bp[i]) * sqrt(bp[k]);
  const double2 v = FSFT[(size_t)k * F + i];
  const double re = v.x * s, im = v.y * s;
  *reinterpret_cast<double2*>(Gext + (size_t)(2 * k) * LD + 2 * i) = make_double2(re, im);
  *reinterpret_cast<double2*>(Gext + (size_t)(2 * k + 1) * LD + 2 * i) = make_double2(-im, re);
}

// The transpose of Gext, written directly (solver 4: the int8 GEMM takes both operands with K contiguous, T1r = Pr (Gext^T)^T):
// GextT[c][r] = Gext[r][c]
__global__ void own_embed_gT_kernel(const double2* __restrict__ FSFT, const double* __restrict__ bp, double* __restrict__ GextT, int F,
                                    int LD) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (k >= F) return;
  const double s = sqrt(bp[i]) * sqrt(bp[k]);
  const double2 v = FSFT[(size_t)k * F + i];
  const double re = v.x * s, im = v.y * s;
  // Gext[2k][2i] = re, Gext[2k][2i+1] = im, Gext[2k+1][2i] = -im, Gext[2k+1][2i+1] = re
  *reinterpret_cast<double2*>(GextT + (size_t)(2 * i) * LD + 2 * k) = make_double2(re, -im);
  *reinterpret_cast<double2*>(GextT + (size_t)(2 * i + 1) * LD + 2 * k) = make_double2(im, re);
}

// Pr (interleaved) and Pext from the filter K of the chunk: P = memory of H = diag(M) K (powerspec.py:281-283)
__global__ void own_prep_p_kernel(const void* __restrict__ K, int k_cplx, const unsigned char* __restrict__ M, double* __restrict__ Pr,
                                  double* __restrict__ Pext, int F, int LD, size_t sPr, size_t sPext, size_t n) {
  const size_t FF = (size_t)F * F;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    const size_t r = e / FF, o = e - r * FF;
    const int j = (int)(o / F), k = (int)(o - (size_t)j * F);  // element (j, k) of the pixel's row-major array
    double2 v = make_double2(0.0, 0.0);
    if (M[r * F + j]) v = load_c(K, e, k_cplx);
    *reinterpret_cast<double2*>(Pr + r * sPr + (size_t)j * LD + 2 * k) = v;
    *reinterpret_cast<double2*>(Pext + r * sPext + (size_t)(2 * j) * LD + 2 * k) = v;
    *reinterpret_cast<double2*>(Pext + r * sPext + (size_t)(2 * j + 1) * LD + 2 * k) = make_double2(-v.y, v.x);
  }
}

// Fpr[d][k] = F'(k, d): the column-major [F, nd] array is the row-major [nd][F] one; only the pitch changes
__global__ void own_pad_fourier_kernel(const double2* __restrict__ Fp, double* __restrict__ Fpr, int F, int nd, int LD) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x, d = blockIdx.y;
  if (k >= F) return;
  *reinterpret_cast<double2*>(Fpr + (size_t)d * LD + 2 * k) = Fp[(size_t)d * F + k];
}

// Ar += C o r r^T (powerspec.py:274-277, 292) on and below the diagonal, unit diagonal on the masked channels.
// Ar[a][b] = A'(b, a); the noise term is symmetric in (a, b) up to the covariance's own symmetry: C is read at [a][b],
// as the column-major reading of wiener_op.cu's add_noise_kernel does.
__global__ void own_add_noise_kernel(double* __restrict__ Ar, const void* __restrict__ cov, int cov_cplx, const double2* __restrict__ rn,
                                     const unsigned char* __restrict__ M, int F, int LD, size_t sAr, size_t n) {
  const size_t FF = (size_t)F * F;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    const size_t r = e / FF, o = e - r * FF;
    const int a = (int)(o / F), b = (int)(o - (size_t)a * F);
    if (b > a) continue;
    const double2 ra = rn[r * F + a], rb = rn[r * F + b];
    const double2 rr = make_double2(ra.x * rb.x - ra.y * rb.y, ra.x * rb.y + ra.y * rb.x);
    const double2 c = load_c(cov, e, cov_cplx);
    double2* p = reinterpret_cast<double2*>(Ar + r * sAr + (size_t)a * LD + 2 * b);
    double2 v = *p;
    v.x += c.x * rr.x - c.y * rr.y;
    v.y += c.x * rr.y + c.y * rr.x;
    if (a == b && !M[r * F + a]) v = make_double2(1.0, 0.0);
    *p = v;
  }
}

// ------------------------------------------------------------------------------------------------ tile GEMM
// C[b] (+)= A[b] B[b]  over 128 x 128 output tiles, one tile per CTA at a time (persistent grid), fp64 DMMA with the
// cp.async double-buffered K loop of dense_core.cuh.  Every dimension is a multiple of 128 (buffers are zero-padded).
//   nn = 1:  C = A B      A [M][K] (lda), B [K][N] (ldb)
//   nn = 0:  C = A B^T    A [M][K], B [N][K], both with pitch lda
// lower_only: skip the tiles that hold no entry on or below the diagonal of a square complex (interleaved) result.
struct TileGemm {
  const double* A; long long sA; int lda;
  const double* B; long long sB; int ldb;
  double* C; long long sC; int ldc;
  int mt, nt, kchunks, nn, lower_only, batch;
};

__global__ void __launch_bounds__(THREADS, 1) own_tile_gemm_kernel(TileGemm g) {
  extern __shared__ __align__(16) double smem[];
  double* pipe = smem;
  const Lane ln;
  const long long per = (long long)g.mt * g.nt, total = per * g.batch;
  for (long long w = blockIdx.x; w < total; w += gridDim.x) {
    const int b = (int)(w / per), t = (int)(w - (long long)b * per);
    const int ti = t / g.nt, tj = t - ti * g.nt;
    // C holds complex entries interleaved along its rows: column tile tj starts at complex column 64 tj; the complex
    // lower triangle (column <= row) needs it only while 64 tj <= 128 ti + 127
    if (g.lower_only && tj > 2 * ti + 1) continue;  // CTA-uniform
    const double* A = g.A + b * g.sA + (size_t)ti * TILE * g.lda;
    Acc acc;
    acc_zero(acc);
    if (g.nn) {
      gemm_nn<false>(pipe, acc, ln, A, g.lda, g.B + b * g.sB + (size_t)tj * TILE, g.ldb, 0, g.kchunks, V_FULL);
    } else {
      gemm_nt(pipe, acc, ln, A, g.B + b * g.sB + (size_t)tj * TILE * g.lda, g.lda, g.kchunks, V_FULL);  // acc = -A B^T
#pragma unroll
      for (int f = 0; f < 2; f++)
#pragma unroll
        for (int nb = 0; nb < 16; nb++) {
          acc[f][nb][0] = -acc[f][nb][0];
          acc[f][nb][1] = -acc[f][nb][1];
        }
    }
    tile_store_global(acc, g.C + b * g.sC + (size_t)ti * TILE * g.ldc + (size_t)tj * TILE, g.ldc, ln, true, true);
  }
}

// ------------------------------------------------------------------------------------------------ factor + solve
struct SolveArgs {
  const double* Ar; size_t sAr;   // [R][FR][LD]
  const double* Yr; size_t sYr;   // [R][NDR][LD]
  double* work; size_t work_stride;
  double* Wp;
  const double* prior; const double* window; const unsigned char* M;
  float2* outd; int* info;
  int R, F, nd, LD, nblk, RT;
};

// adds tile (ti, tj) of the real embedding E of conj(Ar) into a staged tile (identity in the padding):
//   E[2i+p][2j+q] = Re a (p == q),  +Im a (p = 0, q = 1),  -Im a (p = 1, q = 0),   a = Ar[i][j],  i >= j
struct EmbedGen {
  const double* Ar;
  int F, LD;
  __device__ __forceinline__ void operator()(double* sT, int ti, int tj, bool lower_only) const {
    const int R0 = ti * TILE, C0 = tj * TILE, n2 = 2 * F;
#pragma unroll 2
    for (int e = threadIdx.x; e < TILE * TILE; e += THREADS) {
      const int r = e >> 7, c = e & 127;
      if (lower_only && c > r) continue;
      const int R = R0 + r, C = C0 + c;
      double v;
      if (R < n2 && C < n2) {
        const int i = R >> 1, p = R & 1, j = C >> 1, q = C & 1;
        if (i == j && p != q) {
          v = 0.0;  // the diagonal of a Hermitian matrix is real
        } else {
          const double2 a = *reinterpret_cast<const double2*>(Ar + (size_t)i * LD + 2 * j);
          v = (p == q) ? a.x : (p == 0 ? a.y : -a.y);
        }
      } else {
        v = (R == C) ? 1.0 : 0.0;
      }
      sT[r * LDW + c] += v;
    }
  }
};

__global__ void __launch_bounds__(THREADS, 1)
    own_solve_kernel(SolveArgs a, int use_tma, const __grid_constant__ CUtensorMap tmap) {
  extern __shared__ __align__(16) double smem[];
  const Lane ln;
  CholWS w;
  w.sWp = smem;
  w.pipe = smem + WP_DOUBLES;
  w.misc = w.pipe + PIPE_DOUBLES;
  w.npad = a.LD;
  w.nblk = a.nblk;
  w.RT = a.RT;
  w.rhs_frags = (a.nd + 7) >> 3;
  w.Wk = a.work + (size_t)blockIdx.x * a.work_stride;
  w.Rh = w.Wk + (size_t)a.LD * a.LD;
  w.Wpg = a.Wp + (size_t)blockIdx.x * a.nblk * WP_DOUBLES;
  // operand ring of the factorisation's NT updates (tma_pipe.cuh)
  TmaRing ring;
  if (use_tma) tma_ring_init(ring, reinterpret_cast<unsigned long long*>(w.misc + MISC_DOUBLES));
  NtTma nt_tma{&tmap, &ring, w.pipe, (int)blockIdx.x};
  const int half = a.LD / 2;
  // rows of the right-hand-side tiles that no pixel ever writes: keep them finite (they are staged, never used)
  for (long long e = threadIdx.x; e < (long long)a.RT * TILE * half; e += THREADS)
    reinterpret_cast<double2*>(w.Rh)[e] = make_double2(0.0, 0.0);
  for (int r = blockIdx.x; r < a.R; r += gridDim.x) {
    __syncthreads();
    // right-hand sides: the nd rows of Yr (interleaved complex = the embedded vectors); zero rows up to the fragment edge
    const double* Y = a.Yr + (size_t)r * a.sYr;
    const int rows = w.rhs_frags * 8;
    for (long long e = threadIdx.x; e < (long long)rows * half; e += THREADS) {
      const int d = (int)(e / half), k = (int)(e - (long long)d * half) * 2;
      double2 v = make_double2(0.0, 0.0);
      if (d < a.nd) v = *reinterpret_cast<const double2*>(Y + (size_t)d * a.LD + k);
      *reinterpret_cast<double2*>(w.Rh + (size_t)d * a.LD + k) = v;
    }
    __syncthreads();
    EmbedGen gen{a.Ar + (size_t)r * a.sAr, a.F, a.LD};
    const bool ok = use_tma ? chol_factor_forward(w, ln, gen, nt_tma) : chol_factor_forward(w, ln, gen);  // CTA-uniform
    if (!ok) {
      if (threadIdx.x == 0) a.info[r] = 1;
      continue;
    }
    chol_backward(w, ln);
    __syncthreads();
    // out[d][f] = complex64(prior_d window_f x[d][f]) (powerspec.py:314-316); masked channels are exact zeros
    float2* out = a.outd + (size_t)r * a.nd * a.F;
    const unsigned char* Mr = a.M + (size_t)r * a.F;
    for (long long e = threadIdx.x; e < (long long)a.nd * a.F; e += THREADS) {
      const int d = (int)(e / a.F), f = (int)(e - (long long)d * a.F);
      const double s = Mr[f] ? a.prior[d] * a.window[f] : 0.0;
      const double2 x = *reinterpret_cast<const double2*>(w.Rh + (size_t)d * a.LD + 2 * f);
      out[e] = make_float2((float)(s * x.x), (float)(s * x.y));
    }
    if (threadIdx.x == 0) a.info[r] = 0;
  }
}

// ------------------------------------------------------------------------------------------------ host
static size_t gemm_smem() { return sizeof(double) * PIPE_DOUBLES; }
static size_t solve_smem() { return sizeof(double) * (WP_DOUBLES + PIPE_DOUBLES + MISC_DOUBLES + 2 * TNST); }

cudaError_t own_configure() {
  cudaError_t e = cudaFuncSetAttribute(own_tile_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_smem());
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(own_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)solve_smem());
}

cudaError_t own_prepare_fourier(cudaStream_t st, const OwnGeom& g, const cuDoubleComplex* Fp, const OwnBufs& b, int ozaki) {
  own_pad_fourier_kernel<<<dim3((g.F + 127) / 128, g.nd), 128, 0, st>>>(reinterpret_cast<const double2*>(Fp), b.Fpr, g.F, g.nd, g.LD);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess || !ozaki) return e;
  return oz::split(st, b.Fpr, 0, g.LD, g.NDR, g.LD, oz::TM, b.qF, 0, b.eF, 0, 1);
}

static cudaError_t launch_gemm(cudaStream_t st, int sm_count, const TileGemm& t) {
  long long tiles = (long long)t.mt * t.nt * t.batch;
  int grid = (int)(tiles < sm_count ? tiles : sm_count);
  if (grid < 1) return cudaSuccess;
  own_tile_gemm_kernel<<<grid, THREADS, gemm_smem(), st>>>(t);
  return cudaGetLastError();
}

cudaError_t own_gemm(cudaStream_t st, int sm_count, const double* A, long long sA, int lda, const double* B, long long sB, int ldb,
                     double* C, long long sC, int ldc, int mt, int nt, int kchunks, int nn, int lower_only, int batch) {
  cudaError_t e = own_configure();
  if (e != cudaSuccess) return e;
  TileGemm t{A, sA, lda, B, sB, ldb, C, sC, ldc, mt, nt, kchunks, nn, lower_only, batch};
  return launch_gemm(st, sm_count, t);
}

// Solver 4: the three contractions on the INT8 tensor cores (ozaki_gemm.cuh: tcgen05.mma kind::i8, accumulators in TMEM)
// instead of the DMMA tile GEMM; everything else (masks, noise, factorisation, solves, epilogue) as solver 3.
//   T1r = Pr (Gext^T)^T      A operand: slices of Pr,  B operand: slices of Gext^T (once per elevation)
//   Ar  = T1r Pext^T         A operand: slices of T1r (in the buffer Pr's slices leave), B operand: slices of Pext; lower tiles
//   Yr  = Fpr Pext^T         A operand: slices of Fpr (once per call), B operand: the same slices of Pext
static cudaError_t own_chunk_ozaki(cudaStream_t st, const OwnGeom& g, const OwnBufs& b, const OwnSizes& z, int Rc, int sm_count,
                                   const cuDoubleComplex* FSFT, const double* beam_power_el, const void* K, int k_cplx, const void* C,
                                   int c_cplx, const unsigned char* M, const cuDoubleComplex* rn, const double* prior, const double* window,
                                   float2* outd, int* info, int64_t* launches) {
  const size_t nFF = (size_t)Rc * g.F * g.F;
  const int gridE = (int)std::min<size_t>((nFF + 255) / 256, (size_t)sm_count * 16);
  cudaError_t e;
  static const bool prof = getenv("DGWIENER_PROFILE") != nullptr;
  cudaEvent_t ev[12] = {};
  int nev = 0;
  auto mark = [&]() {
    if (prof && nev < 12) {
      cudaEventCreate(&ev[nev]);
      cudaEventRecord(ev[nev++], st);
    }
  };
  mark();
  // Gext is scratch here: it receives the transpose (its own layout is only needed by the DMMA NN GEMM)
  own_embed_gT_kernel<<<dim3((g.F + 127) / 128, g.F), 128, 0, st>>>(reinterpret_cast<const double2*>(FSFT), beam_power_el, b.Gext, g.F,
                                                                  g.LD);
  own_prep_p_kernel<<<gridE, 256, 0, st>>>(K, k_cplx, M, b.Pr, b.Pext, g.F, g.LD, z.Pr, z.Pext, nFF);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  mark();
  if ((e = oz::split(st, b.Gext, 0, g.LD, g.LD, g.LD, oz::TN, b.qG, 0, b.eG, 0, 1)) != cudaSuccess) return e;
  if ((e = oz::split(st, b.Pr, (long long)z.Pr, g.LD, g.FR, g.LD, oz::TM, b.qP, (long long)z.qP, b.eP, (long long)z.eP, Rc)) != cudaSuccess) return e;
  if ((e = oz::split(st, b.Pext, (long long)z.Pext, g.LD, g.LD, g.LD, oz::TN, b.qPe, (long long)z.qPe, b.ePe, (long long)z.ePe, Rc)) != cudaSuccess)
    return e;
  mark();
  oz::GemmParams t1{b.qP, (long long)z.qP, b.qG, 0, b.eP, (long long)z.eP, b.eG, 0, b.T1r, (long long)z.T1r, g.LD, g.FR, g.LD, g.LD, Rc, 0};
  if ((e = oz::gemm_nt(st, sm_count, t1)) != cudaSuccess) return e;
  mark();
  if ((e = oz::split(st, b.T1r, (long long)z.T1r, g.LD, g.FR, g.LD, oz::TM, b.qP, (long long)z.qP, b.eP, (long long)z.eP, Rc)) != cudaSuccess) return e;
  mark();
  oz::GemmParams t2{b.qP, (long long)z.qP, b.qPe, (long long)z.qPe, b.eP, (long long)z.eP, b.ePe, (long long)z.ePe, b.Ar, (long long)z.Ar,
                    g.LD, g.FR, g.LD, g.LD, Rc, 1};
  if ((e = oz::gemm_nt(st, sm_count, t2)) != cudaSuccess) return e;
  mark();
  own_add_noise_kernel<<<gridE, 256, 0, st>>>(b.Ar, C, c_cplx, reinterpret_cast<const double2*>(rn), M, g.F, g.LD, z.Ar, nFF);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  mark();
  oz::GemmParams t3{b.qF, 0, b.qPe, (long long)z.qPe, b.eF, 0, b.ePe, (long long)z.ePe, b.Yr, (long long)z.Yr, g.LD, g.NDR, g.LD, g.LD, Rc, 0};
  if ((e = oz::gemm_nt(st, sm_count, t3)) != cudaSuccess) return e;
  mark();
  SolveArgs a{b.Ar, z.Ar, b.Yr, z.Yr, b.work, g.work_stride, b.Wp, prior, window, M, outd, info, Rc, g.F, g.nd, g.LD, g.nblk, g.RT};
  const int grid = Rc < g.nslots ? Rc : g.nslots;
  CUtensorMap tm;
  const int use_tma = tma_work_map(&tm, b.work, (size_t)g.LD + (size_t)g.RT * TILE, g.LD, g.work_stride, g.nslots,
                                   sizeof(double) * WP_DOUBLES, TILE, TKC);
  own_solve_kernel<<<grid, THREADS, solve_smem(), st>>>(a, use_tma, tm);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  mark();
  if (prof) {
    cudaStreamSynchronize(st);
    static const char* names[9] = {"embed^T + prep", "split G^T, P, Pext", "int8 gemm T1 = P G", "split T1", "int8 gemm A = T1 P^H (lower)",
                                   "add noise", "int8 gemm Y = F' P^H", "factor + solve + epilogue", ""};
    for (int i = 0; i + 1 < nev; i++) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, ev[i], ev[i + 1]);
      fprintf(stderr, "[dgwiener ozaki] F=%d Rc=%d %-30s %9.3f ms\n", g.F, Rc, names[i], ms);
    }
    for (int i = 0; i < nev; i++) cudaEventDestroy(ev[i]);
  }
  if (launches) *launches += 12;
  return cudaSuccess;
}

cudaError_t own_chunk(cudaStream_t st, const OwnGeom& g, const OwnBufs& b, int Rc, int sm_count, const cuDoubleComplex* FSFT,
                      const double* beam_power_el, const void* K, int k_cplx, const void* C, int c_cplx, const unsigned char* M,
                      const cuDoubleComplex* rn, const double* prior, const double* window, float2* outd, int* info,
                      int64_t* launches, int ozaki) {
  const OwnSizes z = own_sizes(g);
  const size_t nFF = (size_t)Rc * g.F * g.F;
  const int gridE = (int)std::min<size_t>((nFF + 255) / 256, (size_t)sm_count * 16);
  cudaError_t e;
  // DGWIENER_PROFILE=1: CUDA-event time of every launch of the chunk on stderr (debug)
  static const bool prof = getenv("DGWIENER_PROFILE") != nullptr;
  cudaEvent_t ev[8] = {};
  int nev = 0;
  if (ozaki) return own_chunk_ozaki(st, g, b, z, Rc, sm_count, FSFT, beam_power_el, K, k_cplx, C, c_cplx, M, rn, prior, window, outd, info, launches);
  auto mark = [&]() {
    if (prof && nev < 8) {
      cudaEventCreate(&ev[nev]);
      cudaEventRecord(ev[nev++], st);
    }
  };
  mark();
  own_embed_g_kernel<<<dim3((g.F + 127) / 128, g.F), 128, 0, st>>>(reinterpret_cast<const double2*>(FSFT), beam_power_el, b.Gext, g.F,
                                                                 g.LD);
  own_prep_p_kernel<<<gridE, 256, 0, st>>>(K, k_cplx, M, b.Pr, b.Pext, g.F, g.LD, z.Pr, z.Pext, nFF);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  mark();
  const int kch = g.LD / KC;
  // T1r = Pr Gext
  TileGemm t1{b.Pr, (long long)z.Pr, g.LD, b.Gext, 0, g.LD, b.T1r, (long long)z.T1r, g.LD, g.FR / TILE, g.LD / TILE, kch, 1, 0, Rc};
  if ((e = launch_gemm(st, sm_count, t1)) != cudaSuccess) return e;
  mark();
  // Ar = T1r Pext^T (tiles on and below the complex diagonal)
  TileGemm t2{b.T1r, (long long)z.T1r, g.LD, b.Pext, (long long)z.Pext, g.LD, b.Ar, (long long)z.Ar, g.LD, g.FR / TILE, g.LD / TILE, kch, 0, 1, Rc};
  if ((e = launch_gemm(st, sm_count, t2)) != cudaSuccess) return e;
  mark();
  own_add_noise_kernel<<<gridE, 256, 0, st>>>(b.Ar, C, c_cplx, reinterpret_cast<const double2*>(rn), M, g.F, g.LD, z.Ar, nFF);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  mark();
  // Yr = Fpr Pext^T
  TileGemm t3{b.Fpr, 0, g.LD, b.Pext, (long long)z.Pext, g.LD, b.Yr, (long long)z.Yr, g.LD, g.NDR / TILE, g.LD / TILE, kch, 0, 0, Rc};
  if ((e = launch_gemm(st, sm_count, t3)) != cudaSuccess) return e;
  mark();
  SolveArgs a{b.Ar, z.Ar, b.Yr, z.Yr, b.work, g.work_stride, b.Wp, prior, window, M, outd, info, Rc, g.F, g.nd, g.LD, g.nblk, g.RT};
  const int grid = Rc < g.nslots ? Rc : g.nslots;
  CUtensorMap tm;
  const int use_tma = tma_work_map(&tm, b.work, (size_t)g.LD + (size_t)g.RT * TILE, g.LD, g.work_stride, g.nslots,
                                   sizeof(double) * WP_DOUBLES, TILE, TKC);
  own_solve_kernel<<<grid, THREADS, solve_smem(), st>>>(a, use_tma, tm);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  mark();
  if (prof) {
    cudaStreamSynchronize(st);
    static const char* names[6] = {"embed+prep", "gemm T1 = P G", "gemm A = T1 P^H (lower)", "add noise", "gemm Y = F' P^H", "factor + solve + epilogue"};
    for (int i = 0; i + 1 < nev; i++) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, ev[i], ev[i + 1]);
      fprintf(stderr, "[dgwiener own] F=%d Rc=%d %-28s %9.3f ms\n", g.F, Rc, names[i], ms);
    }
    for (int i = 0; i < nev; i++) cudaEventDestroy(ev[i]);
  }
  if (launches) *launches += 7;
  return cudaSuccess;
}

OzakiScratch own_gemm_ozaki_sizes(int M, int N, int K, int batch) {
  OzakiScratch z{};
  z.qa = (size_t)batch * oz::slice_bytes(M, K);
  z.qb = (size_t)batch * oz::slice_bytes(N, K);
  z.ea = (size_t)batch * oz::exp_ints(M, K);
  z.eb = (size_t)batch * oz::exp_ints(N, K);
  return z;
}

cudaError_t own_gemm_ozaki(cudaStream_t st, int sm_count, const double* A, const double* B, double* C, int M, int N, int K, int batch,
                           int8_t* qa, int8_t* qb, int* ea, int* eb) {
  const long long sQA = (long long)oz::slice_bytes(M, K), sQB = (long long)oz::slice_bytes(N, K);
  const long long sEA = (long long)oz::exp_ints(M, K), sEB = (long long)oz::exp_ints(N, K);
  cudaError_t e;
  if ((e = oz::split(st, A, (long long)M * K, K, M, K, oz::TM, qa, sQA, ea, sEA, batch)) != cudaSuccess) return e;
  if ((e = oz::split(st, B, (long long)N * K, K, N, K, oz::TN, qb, sQB, eb, sEB, batch)) != cudaSuccess) return e;
  oz::GemmParams p{qa, sQA, qb, sQB, ea, sEA, eb, sEB, C, (long long)M * N, N, M, N, K, batch, 0};
  return oz::gemm_nt(st, sm_count, p);
}

cudaError_t own_gemm_ozaki_pitched(cudaStream_t st, int sm_count, const double* A, int lda, const double* B, int ldb, double* C, int ldc,
                                   int M, int N, int K, int lower_only, int8_t* qa, int8_t* qb, int* ea, int* eb) {
  cudaError_t e;
  if ((e = oz::split(st, A, 0, lda, M, K, oz::TM, qa, 0, ea, 0, 1)) != cudaSuccess) return e;
  if ((e = oz::split(st, B, 0, ldb, N, K, oz::TN, qb, 0, eb, 0, 1)) != cudaSuccess) return e;
  oz::GemmParams p{qa, 0, qb, 0, ea, 0, eb, 0, C, 0, ldc, M, N, K, 1, lower_only};
  return oz::gemm_nt(st, sm_count, p);
}

}  // namespace dgw
