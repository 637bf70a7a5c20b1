// Own-kernel path of the NRML objective: LogLikePS._precompute / value / hessian, draco/analysis/delayopt.py:120-236,
//     C = U U^H + N ,  cho_factor(C) ,  Ut = C^-1 U ,  C^-1 X ,  W = (X - C) Ut ,  G = U^H Ut ,  K = Ut^H W
// on this repository's fp64 DMMA tile machinery instead of cuBLAS ZHERK / ZGEMM and cuSOLVER ZPOTRF / ZPOTRS (nrml.cu
// keeps those behind DGNRML_LIBRARY=1 for A/B).  One M x M system at a time fills the GPU, so unlike the samplers the
// factorisation is spread over the whole grid:
//
//   * every complex matrix is row-major with interleaved (re, im) doubles and a 128-padded real pitch (a column-major
//     complex [M, N] array IS the row-major [N][M] one); products are REAL GEMMs of exactly the complex flop count:
//     C = Um Um_ext^T (lower tiles), W = Ut XC_emb, G = Ut U_ext^T, K = W Ut_ext^T -- since r04 on the INT8 tensor cores
//     (Ozaki split, tcgen05.mma kind::i8, fp64-equivalent: ozaki_gemm.cuh through dgw::own_gemm_ozaki_pitched); DGNRML_NO_INT8=1
//     keeps the DMMA tile GEMM dgw::own_gemm (wiener_own.cu);
//   * the Hermitian system runs in its real embedding E (2M x 2M, identity in the padding) through a RIGHT-LOOKING
//     blocked Cholesky over 128-wide block columns: per block column one CTA factors the diagonal tile (factor_diag
//     of dense_core.cuh: 8-column panels, 8 x 8 inverses only), one CTA per tile below it applies inv(L_jj)^T, and a
//     2-D grid applies the rank-128 update to every trailing tile.  The right-hand sides -- the N columns of U and the
//     M columns of X, each an interleaved row -- ride along as extra row tiles (forward solve for free); the backward
//     solve is the mirror image (tile solve against inv(L_jj), rank-128 update of the columns to the left).
//     5 launches per block column; every tile operation is a DMMA GEMM on 128 x 128 x 128.
#include "nrml_own.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "dense_core.cuh"
#include "wiener_own.h"

namespace dgn {
using namespace dg;

static int round_up(int x, int m) { return (x + m - 1) / m * m; }

// ------------------------------------------------------------------------------------------------ elementwise
// Um[i][n] = MFm[i][n] sqrt(S_n) (delayopt.py:153) and its "conj" extension: rows 2i / 2i+1 = (re, im) / (-im, re)
__global__ void own_scale_um_kernel(const double* __restrict__ MFm, const double* __restrict__ sq, double* __restrict__ Um,
                                    double* __restrict__ UmExt, int M, int N, int LDN) {
  const size_t n_el = (size_t)M * N;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n_el; e += (size_t)gridDim.x * blockDim.x) {
    const int i = (int)(e / N), n = (int)(e - (size_t)i * N);
    const double2 m = *reinterpret_cast<const double2*>(MFm + (size_t)i * LDN + 2 * n);
    const double s = sq[n];
    const double2 v = make_double2(m.x * s, m.y * s);
    *reinterpret_cast<double2*>(Um + (size_t)i * LDN + 2 * n) = v;
    *reinterpret_cast<double2*>(UmExt + (size_t)(2 * i) * LDN + 2 * n) = v;
    *reinterpret_cast<double2*>(UmExt + (size_t)(2 * i + 1) * LDN + 2 * n) = make_double2(-v.y, v.x);
  }
}

// rows [n][.] of a column-major complex [M, ncols] array (ld M) -> interleaved rows with pitch LD (+ optional "conj"
// extension rows 2n / 2n+1 = (re, im) / (-im, re))
__global__ void own_rows_kernel(const double2* __restrict__ src, int M, int ncols, double* __restrict__ dst, int LD,
                                double* __restrict__ ext) {
  const size_t n_el = (size_t)M * ncols;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n_el; e += (size_t)gridDim.x * blockDim.x) {
    const int n = (int)(e / M), i = (int)(e - (size_t)n * M);
    const double2 v = src[e];
    if (dst) *reinterpret_cast<double2*>(dst + (size_t)n * LD + 2 * i) = v;
    if (ext) {
      *reinterpret_cast<double2*>(ext + (size_t)(2 * n) * LD + 2 * i) = v;
      *reinterpret_cast<double2*>(ext + (size_t)(2 * n + 1) * LD + 2 * i) = make_double2(-v.y, v.x);
    }
  }
}
// the same from interleaved rows (pitch LD) instead of a column-major array
__global__ void own_ext_rows_kernel(const double* __restrict__ rows, int nrows, int M, int LD, double* __restrict__ ext) {
  const size_t n_el = (size_t)M * nrows;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n_el; e += (size_t)gridDim.x * blockDim.x) {
    const int n = (int)(e / M), i = (int)(e - (size_t)n * M);
    const double2 v = *reinterpret_cast<const double2*>(rows + (size_t)n * LD + 2 * i);
    *reinterpret_cast<double2*>(ext + (size_t)(2 * n) * LD + 2 * i) = v;
    *reinterpret_cast<double2*>(ext + (size_t)(2 * n + 1) * LD + 2 * i) = make_double2(-v.y, v.x);
  }
}

// From Q = Um Um^H (lower triangle, Q[i][j] = C(i, j) for i >= j): add the noise variances (delayopt.py:143-144), write
// the lower triangle of the real embedding E of C into the factorisation workspace and the real embedding of
// XCrow[k][i] = (X - C)(i, k) (delayopt.py:151; X column-major, Hermitian) for the product W = (X - C) Ut.
//   E[2i+p][2j+q]      = Re c (p == q),  -Im c (p = 0, q = 1),  +Im c (p = 1, q = 0)
//   XCemb rows 2k / 2k+1 at columns (2i, 2i+1) = (re, im) / (-im, re) of XC(i, k)
__global__ void own_finish_kernel(const double* __restrict__ Q, const double* __restrict__ noise, const double2* __restrict__ X,
                                  double* __restrict__ E, double* __restrict__ XCemb, int M, int LDM) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
  if (i >= M || i < j) return;
  double2 c = *reinterpret_cast<const double2*>(Q + (size_t)i * LDM + 2 * j);
  if (i == j) c = make_double2(c.x + noise[i], 0.0);
  // E: block (i, j), i >= j
  E[(size_t)(2 * i) * LDM + 2 * j] = c.x;
  E[(size_t)(2 * i + 1) * LDM + 2 * j] = c.y;
  E[(size_t)(2 * i + 1) * LDM + 2 * j + 1] = c.x;
  if (i != j) E[(size_t)(2 * i) * LDM + 2 * j + 1] = -c.y;
  // XC(i, j) = X(i, j) - c  ->  XCrow[k = j][i]
  const double2 xl = X[(size_t)j * M + i];
  const double2 a = make_double2(xl.x - c.x, xl.y - c.y);
  *reinterpret_cast<double2*>(XCemb + (size_t)(2 * j) * LDM + 2 * i) = a;
  *reinterpret_cast<double2*>(XCemb + (size_t)(2 * j + 1) * LDM + 2 * i) = make_double2(-a.y, a.x);
  if (i != j) {  // XC(j, i) = X(j, i) - conj(c)  ->  XCrow[k = i][j]
    const double2 xu = X[(size_t)i * M + j];
    const double2 b = make_double2(xu.x - c.x, xu.y + c.y);
    *reinterpret_cast<double2*>(XCemb + (size_t)(2 * i) * LDM + 2 * j) = b;
    *reinterpret_cast<double2*>(XCemb + (size_t)(2 * i + 1) * LDM + 2 * j) = make_double2(-b.y, b.x);
  }
}

// t_a = -sum_i Re(conj(Ut[i, a]) W[i, a])  (delayopt.py:205, 232): one warp per row of the interleaved arrays
__global__ void own_row_dot_kernel(const double* __restrict__ Ut, const double* __restrict__ W, double* __restrict__ t, int M, int N,
                                   int LD) {
  const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (a >= N) return;
  const double* u = Ut + (size_t)a * LD;
  const double* w = W + (size_t)a * LD;
  double s = 0.0;
  for (int i = lane; i < 2 * M; i += 32) s += u[i] * w[i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) t[a] = -s;
}

// out[0] = log det C = sum over the embedded rows of log L_ii (each complex pivot appears twice: 2 log L_ii);
// out[1] = trace(C^-1 X) = sum_j Re (C^-1 X)(j, j), column j of C^-1 X being row j of `CiX`
__global__ void own_scalars_kernel(const double* __restrict__ logl, int nblk, const double* __restrict__ CiX, int M, int LD,
                                   double* __restrict__ out) {
  __shared__ double red[32];
  double s = 0.0;
  for (int j = threadIdx.x; j < M; j += blockDim.x) s += CiX[(size_t)j * LD + 2 * j];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += red[w];
    out[1] = t;
    double l = 0.0;
    for (int b = 0; b < nblk; b++) l += logl[b];
    out[0] = l;
  }
}

// ------------------------------------------------------------------------------------------------ grid Cholesky
// Workspace T: row tiles 0 .. nblk-1 = E (lower), then the right-hand-side row tiles; pitch LD.
// (1) diagonal tile of block column j: factor, keep the packed inverse blocks, report the pivots
__global__ void __launch_bounds__(THREADS, 1) rl_diag_kernel(double* __restrict__ T, int LD, int j, double* __restrict__ Wpg,
                                                            double* __restrict__ logl, int* __restrict__ flag) {
  extern __shared__ __align__(16) double smem[];
  double* sWp = smem;
  double* sT = smem + WP_DOUBLES;
  double* misc = sT + PIPE_DOUBLES;
  const Lane ln;
  if (*flag) return;  // an earlier block column met a non-positive pivot
  const double* src = T + (size_t)j * TILE * LD + (size_t)j * TILE;
  for (int e = threadIdx.x; e < TILE * (TILE / 2); e += THREADS) {
    const int r = e >> 6, c = (e & 63) * 2;
    *reinterpret_cast<double2*>(sT + r * LDW + c) = *reinterpret_cast<const double2*>(src + (size_t)r * LD + c);
  }
  __syncthreads();
  const bool ok = factor_diag(sT, sWp, misc, ln);
  if (!ok) {
    if (threadIdx.x == 0) *flag = j + 1;
    return;
  }
  for (int idx = threadIdx.x; idx < WP_DOUBLES / 2; idx += THREADS)
    reinterpret_cast<double2*>(Wpg + (size_t)j * WP_DOUBLES)[idx] = reinterpret_cast<const double2*>(sWp)[idx];
  // sum of log L_ii over the tile: the diagonal slots of the packed factor hold the 8 x 8 inverses, diag = 1 / L_ii
  if (threadIdx.x < TILE) {
    const int p = threadIdx.x >> 3, i = threadIdx.x & 7;
    double v = -log(sWp[wp_block(p, p) + i * 8 + (i ^ wp_swz(i))]);
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) misc[8 + (threadIdx.x >> 5)] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) logl[j] = misc[8] + misc[9] + misc[10] + misc[11];
}

// (2) tiles (i, j), i > j (matrix rows and right-hand sides):  T(i, j) <- T(i, j) inv(L_jj)^T
__global__ void __launch_bounds__(THREADS, 1) rl_panel_kernel(double* __restrict__ T, int LD, int j, const double* __restrict__ Wpg,
                                                             const int* __restrict__ flag) {
  extern __shared__ __align__(16) double smem[];
  double* sWp = smem;
  double* sT = smem + WP_DOUBLES;
  const Lane ln;
  if (*flag) return;
  const int i = j + 1 + blockIdx.x;
  double* tile = T + (size_t)i * TILE * LD + (size_t)j * TILE;
  for (int idx = threadIdx.x; idx < WP_DOUBLES / 2; idx += THREADS)
    reinterpret_cast<double2*>(sWp)[idx] = reinterpret_cast<const double2*>(Wpg + (size_t)j * WP_DOUBLES)[idx];
  for (int e = threadIdx.x; e < TILE * (TILE / 2); e += THREADS) {
    const int r = e >> 6, c = (e & 63) * 2;
    *reinterpret_cast<double2*>(sT + r * LDW + c) = *reinterpret_cast<const double2*>(tile + (size_t)r * LD + c);
  }
  __syncthreads();
  trisolve<false>(sT, sWp, ln, V_FULL);
  __syncthreads();
  copy_tile_out(sT, tile, LD, TILE);
}

// (3) trailing update: T(i, k) -= T(i, j) T(k, j)^T for j < k < nblk, k <= i < ntiles.  blockIdx = (i - k, k - j - 1)
__global__ void __launch_bounds__(THREADS, 1) rl_update_kernel(double* __restrict__ T, int LD, int j, int ntiles,
                                                              const int* __restrict__ flag) {
  extern __shared__ __align__(16) double smem[];
  const Lane ln;
  if (*flag) return;
  const int k = j + 1 + blockIdx.y, i = k + blockIdx.x;
  if (i >= ntiles) return;
  double* out = T + (size_t)i * TILE * LD + (size_t)k * TILE;
  Acc acc;
  tile_load_global(acc, out, LD, ln);
  gemm_nt(smem, acc, ln, T + (size_t)i * TILE * LD + (size_t)j * TILE, T + (size_t)k * TILE * LD + (size_t)j * TILE, LD, TILE / KC,
          V_FULL);
  tile_store_global(acc, out, LD, ln, true, true);
}

// (4) backward, block column j: right-hand-side tiles  Z(rt, j) <- Z(rt, j) inv(L_jj)
__global__ void __launch_bounds__(THREADS, 1) rl_back_diag_kernel(double* __restrict__ T, int LD, int j, int nblk,
                                                                 const double* __restrict__ Wpg) {
  extern __shared__ __align__(16) double smem[];
  double* sWp = smem;
  double* sT = smem + WP_DOUBLES;
  const Lane ln;
  double* tile = T + (size_t)(nblk + blockIdx.x) * TILE * LD + (size_t)j * TILE;
  for (int idx = threadIdx.x; idx < WP_DOUBLES / 2; idx += THREADS)
    reinterpret_cast<double2*>(sWp)[idx] = reinterpret_cast<const double2*>(Wpg + (size_t)j * WP_DOUBLES)[idx];
  for (int e = threadIdx.x; e < TILE * (TILE / 2); e += THREADS) {
    const int r = e >> 6, c = (e & 63) * 2;
    *reinterpret_cast<double2*>(sT + r * LDW + c) = *reinterpret_cast<const double2*>(tile + (size_t)r * LD + c);
  }
  __syncthreads();
  trisolve<true>(sT, sWp, ln, V_FULL);
  __syncthreads();
  copy_tile_out(sT, tile, LD, TILE);
}

// (5) backward update: Z(rt, k) -= X(rt, j) L(j, k) for k < j.  blockIdx = (rt, k)
__global__ void __launch_bounds__(THREADS, 1) rl_back_update_kernel(double* __restrict__ T, int LD, int j, int nblk) {
  extern __shared__ __align__(16) double smem[];
  const Lane ln;
  const int k = blockIdx.y;
  double* out = T + (size_t)(nblk + blockIdx.x) * TILE * LD + (size_t)k * TILE;
  Acc acc;
  tile_load_global(acc, out, LD, ln);
  gemm_nn<true>(smem, acc, ln, T + (size_t)(nblk + blockIdx.x) * TILE * LD + (size_t)j * TILE, LD,
                T + (size_t)j * TILE * LD + (size_t)k * TILE, LD, 0, TILE / KC, V_FULL);
  tile_store_global(acc, out, LD, ln, true, true);
}

// ------------------------------------------------------------------------------------------------ host
static size_t smem_factor() { return sizeof(double) * (WP_DOUBLES + PIPE_DOUBLES + MISC_DOUBLES); }
static size_t smem_panel() { return sizeof(double) * (WP_DOUBLES + PIPE_DOUBLES); }
static size_t smem_update() { return sizeof(double) * PIPE_DOUBLES; }

__global__ void own_pad_diag_kernel(double* __restrict__ E, int n_valid, int LD) {
  const int r = n_valid + blockIdx.x * blockDim.x + threadIdx.x;
  if (r < LD) E[(size_t)r * LD + r] = 1.0;
}

cudaError_t own_alloc(Own& o, int M, int N, int sm_count, cudaStream_t st) {
  o.M = M; o.N = N; o.sm_count = sm_count;
  o.LDM = round_up(2 * M, TILE); o.LDN = round_up(2 * N, TILE);
  o.MR = round_up(M, TILE); o.NR = round_up(N, TILE);
  o.nblk = o.LDM / TILE; o.rtU = o.NR / TILE; o.rtX = o.MR / TILE;
  const size_t rowsT = (size_t)o.LDM + o.NR + o.MR;
  struct { double** p; size_t n; } list[] = {
      {&o.MFm, (size_t)o.MR * o.LDN}, {&o.Um, (size_t)o.MR * o.LDN}, {&o.UmExt, (size_t)o.LDM * o.LDN}, {&o.Q, (size_t)o.MR * o.LDM},
      {&o.T, rowsT * o.LDM}, {&o.XCemb, (size_t)o.LDM * o.LDM}, {&o.UrowExt, (size_t)o.LDN * o.LDM}, {&o.UtrowExt, (size_t)o.LDN * o.LDM},
      {&o.Wrow, (size_t)o.NR * o.LDM}, {&o.G, (size_t)o.NR * o.LDN}, {&o.K, (size_t)o.NR * o.LDN}, {&o.Wpg, (size_t)o.nblk * WP_DOUBLES},
      {&o.logl, (size_t)o.nblk}};
  for (auto& it : list) {
    cudaError_t e = cudaMalloc((void**)it.p, it.n * sizeof(double));
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(*it.p, 0, it.n * sizeof(double), st);  // the padding of every matrix stays zero from here on
    if (e != cudaSuccess) return e;
  }
  cudaError_t e = cudaMalloc((void**)&o.flag, sizeof(int));
  if (e != cudaSuccess) return e;
  o.int8 = getenv("DGNRML_NO_INT8") == nullptr;
  if (o.int8) {  // scratch of the int8 contractions: the largest A operand (Um or Ut / W rows) and B operand (an extension)
    const dgw::OzakiScratch z1 = dgw::own_gemm_ozaki_sizes(o.MR, o.LDM, o.LDN, 1), z2 = dgw::own_gemm_ozaki_sizes(o.NR, o.LDN, o.LDM, 1);
    if ((e = cudaMalloc((void**)&o.qa, std::max(z1.qa, z2.qa))) != cudaSuccess) return e;
    if ((e = cudaMalloc((void**)&o.qb, std::max(z1.qb, z2.qb))) != cudaSuccess) return e;
    if ((e = cudaMalloc((void**)&o.ea, std::max(z1.ea, z2.ea) * sizeof(int))) != cudaSuccess) return e;
    if ((e = cudaMalloc((void**)&o.eb, std::max(z1.eb, z2.eb) * sizeof(int))) != cudaSuccess) return e;
  }
  if (o.LDM > 2 * M) own_pad_diag_kernel<<<(o.LDM - 2 * M + 127) / 128, 128, 0, st>>>(o.T, 2 * M, o.LDM);  // identity in the padding
  for (auto kern : {(const void*)rl_diag_kernel}) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_factor());
    if (e != cudaSuccess) return e;
  }
  for (auto kern : {(const void*)rl_panel_kernel, (const void*)rl_back_diag_kernel}) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_panel());
    if (e != cudaSuccess) return e;
  }
  for (auto kern : {(const void*)rl_update_kernel, (const void*)rl_back_update_kernel}) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_update());
    if (e != cudaSuccess) return e;
  }
  return cudaGetLastError();
}

void own_free(Own& o) {
  for (double* p : {o.MFm, o.Um, o.UmExt, o.Q, o.T, o.XCemb, o.UrowExt, o.UtrowExt, o.Wrow, o.G, o.K, o.Wpg, o.logl})
    if (p) cudaFree(p);
  if (o.flag) cudaFree(o.flag);
  for (void* p : {(void*)o.qa, (void*)o.qb, (void*)o.ea, (void*)o.eb})
    if (p) cudaFree(p);
  o = Own{};
}

__global__ void own_pitch_rows_kernel(const double2* __restrict__ src, int nrows, int ncols, double* __restrict__ dst, int LD) {
  const size_t n_el = (size_t)nrows * ncols;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n_el; e += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(e / ncols), c = (int)(e - (size_t)r * ncols);
    *reinterpret_cast<double2*>(dst + (size_t)r * LD + 2 * c) = src[e];
  }
}

cudaError_t own_set_mf(Own& o, const cuDoubleComplex* MF_rowmajor, cudaStream_t st) {
  own_pitch_rows_kernel<<<2048, 256, 0, st>>>(reinterpret_cast<const double2*>(MF_rowmajor), o.M, o.N, o.MFm, o.LDN);
  return cudaGetLastError();
}

// C = A B^T (both operands with K contiguous): the INT8 tensor-core GEMM (fp64-equivalent, ozaki_gemm.cuh) or the DMMA tile GEMM
static cudaError_t gemm_nt(Own& o, cudaStream_t st, const double* A, int lda, const double* B, int ldb, double* C, int ldc, int Mr, int Nc,
                           int K, int lower_only) {
  if (o.int8) return dgw::own_gemm_ozaki_pitched(st, o.sm_count, A, lda, B, ldb, C, ldc, Mr, Nc, K, lower_only, o.qa, o.qb, o.ea, o.eb);
  if (lda != ldb) return cudaErrorInvalidValue;  // the DMMA NT form shares one pitch
  return dgw::own_gemm(st, o.sm_count, A, 0, lda, B, 0, ldb, C, 0, ldc, Mr / TILE, Nc / TILE, K / KC, 0, lower_only, 1);
}

cudaError_t own_precompute(Own& o, cudaStream_t st, const double* sq, const cuDoubleComplex* U_colmajor,
                           const cuDoubleComplex* X_colmajor, const double* noise, double* t_dev, double* scal_dev,
                           int* info_host, int64_t* launches) {
  const int M = o.M, N = o.N, LDM = o.LDM, LDN = o.LDN, nblk = o.nblk;
  const int ntiles = nblk + o.rtU + o.rtX, nrhs = o.rtU + o.rtX;
  cudaError_t e;
  int64_t nl = 0;
  // U in both layouts: Um [M][N] (+ its extension) for C = U U^H; the rows of the column-major U are the right-hand sides
  own_scale_um_kernel<<<2048, 256, 0, st>>>(o.MFm, sq, o.Um, o.UmExt, M, N, LDN);
  own_rows_kernel<<<2048, 256, 0, st>>>(reinterpret_cast<const double2*>(U_colmajor), M, N, o.Ut(), LDM, o.UrowExt);
  own_rows_kernel<<<2048, 256, 0, st>>>(reinterpret_cast<const double2*>(X_colmajor), M, M, o.CiX(), LDM, nullptr);
  nl += 3;
  // Q = Um Um^H on and below the diagonal: Q[i][j] = C(i, j) - noise
  if ((e = gemm_nt(o, st, o.Um, LDN, o.UmExt, LDN, o.Q, LDM, o.MR, LDM, LDN, 1)) != cudaSuccess) return e;
  own_finish_kernel<<<dim3((M + 127) / 128, M), 128, 0, st>>>(o.Q, noise, reinterpret_cast<const double2*>(X_colmajor), o.T, o.XCemb, M,
                                                            LDM);
  nl += 2;
  if ((e = cudaMemsetAsync(o.flag, 0, sizeof(int), st)) != cudaSuccess) return e;
  // right-looking factorisation, the right-hand sides riding along
  for (int j = 0; j < nblk; j++) {
    rl_diag_kernel<<<1, THREADS, smem_factor(), st>>>(o.T, LDM, j, o.Wpg, o.logl, o.flag);
    rl_panel_kernel<<<ntiles - j - 1, THREADS, smem_panel(), st>>>(o.T, LDM, j, o.Wpg, o.flag);
    nl += 2;
    if (j + 1 < nblk) {
      rl_update_kernel<<<dim3(ntiles - j - 1, nblk - j - 1), THREADS, smem_update(), st>>>(o.T, LDM, j, ntiles, o.flag);
      nl += 1;
    }
  }
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  int flag = 0;
  if ((e = cudaMemcpyAsync(&flag, o.flag, sizeof(int), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
  if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
  *info_host = flag;
  if (launches) *launches += nl;
  if (flag) return cudaSuccess;
  nl = 0;
  // backward solve of every right-hand side
  for (int j = nblk - 1; j >= 0; j--) {
    rl_back_diag_kernel<<<nrhs, THREADS, smem_panel(), st>>>(o.T, LDM, j, nblk, o.Wpg);
    nl += 1;
    if (j > 0) {
      rl_back_update_kernel<<<dim3(nrhs, j), THREADS, smem_update(), st>>>(o.T, LDM, j, nblk);
      nl += 1;
    }
  }
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  // W = (X - C) Ut :  Wrow = Utrow XCemb ;  t ;  log det, trace
  // (the real embedding of the Hermitian X - C is symmetric, so the int8 path takes it as the [N][K] operand of an NT product;
  // the two readings differ by the rounding asymmetry of the caller's X(i, j) against conj X(j, i))
  if (o.int8) {
    if ((e = gemm_nt(o, st, o.Ut(), LDM, o.XCemb, LDM, o.Wrow, LDM, o.NR, LDM, LDM, 0)) != cudaSuccess) return e;
  } else if ((e = dgw::own_gemm(st, o.sm_count, o.Ut(), 0, LDM, o.XCemb, 0, LDM, o.Wrow, 0, LDM, o.NR / TILE, LDM / TILE, LDM / KC, 1, 0,
                                1)) != cudaSuccess) {
    return e;
  }
  own_row_dot_kernel<<<(N + 7) / 8, 256, 0, st>>>(o.Ut(), o.Wrow, t_dev, M, N, LDM);
  own_scalars_kernel<<<1, 256, 0, st>>>(o.logl, nblk, o.CiX(), M, LDM, scal_dev);
  nl += 3;
  if (launches) *launches += nl;
  return cudaGetLastError();
}

cudaError_t own_hessian_terms(Own& o, cudaStream_t st, int exact, int64_t* launches) {
  cudaError_t e;
  // G[b][a] = sum_i Utrow[b][i] conj(Urow[a][i]) = (U^H Ut)(a, b)
  if ((e = gemm_nt(o, st, o.Ut(), o.LDM, o.UrowExt, o.LDM, o.G, o.LDN, o.NR, o.LDN, o.LDM, 0)) != cudaSuccess) return e;
  if (launches) *launches += 1;
  if (exact) {  // K[b][a] = sum_i Wrow[b][i] conj(Utrow[a][i]) = (Ut^H W)(a, b)
    own_ext_rows_kernel<<<2048, 256, 0, st>>>(o.Ut(), o.N, o.M, o.LDM, o.UtrowExt);
    if ((e = gemm_nt(o, st, o.Wrow, o.LDM, o.UtrowExt, o.LDM, o.K, o.LDN, o.NR, o.LDN, o.LDM, 0)) != cudaSuccess) return e;
    if (launches) *launches += 2;
  }
  return cudaGetLastError();
}

}  // namespace dgn
