// Own-kernel path of the Wiener delay-transform operator (dgwiener_construct, solver = 3): every contraction and the
// factorisation / solves of draco/analysis/powerspec.py:288-316 on hand-written fp64 DMMA kernels (wiener_own.cu)
// instead of cuBLAS / cuSOLVER.  Internal interface between wiener_op.cu (orchestration, mask / noise kernels) and
// wiener_own.cu.
#pragma once
#include <cuComplex.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace dgw {

struct OwnGeom {
  int F, nd;
  int LD;     // real row pitch of every matrix: 2 F rounded up to 128 (interleaved re / im)
  int FR;     // F rounded up to 128 (row tiles of the F-row matrices)
  int NDR;    // nd rounded up to 128
  int nblk;   // LD / 128: block columns of the embedded system
  int RT;     // NDR / 128: right-hand-side row tiles
  int nslots; // concurrent factorisations (one persistent CTA each)
  size_t work_stride;  // doubles per slot: (LD + RT * 128) * LD
};
OwnGeom own_geom(int F, int nd, int sm_count);

// doubles of each per-pixel buffer (allocate R times) and of the shared ones
struct OwnSizes {
  size_t Pr, Pext, T1r, Ar, Yr;      // per pixel
  size_t Gext, Fpr, work, Wp;        // once
  // int8 contractions (ozaki_gemm.cuh; solver 4): slice bytes / exponent ints of the operands
  size_t qP, qPe, eP, ePe;           // per pixel: Pr (then T1r) as A operand, Pext as B operand
  size_t qG, qF, eG, eF;             // once: Gext^T as B operand, Fpr as A operand
};
OwnSizes own_sizes(const OwnGeom& g);

struct OwnBufs {
  double *Pr, *Pext, *T1r, *Ar, *Yr;  // [R][...]; must be zero-filled once after allocation (the padding stays zero)
  double *Gext, *Fpr, *work, *Wp;
  // solver 4 only (null otherwise)
  int8_t *qP, *qPe, *qG, *qF;
  int *eP, *ePe, *eG, *eF;
};

cudaError_t own_configure();
// Fpr (zero-padded row-major copy of F'^T: [nd][F] complex) from the column-major F' [F, nd]
// (ozaki: also its int8 slices, for the contraction Y = F' P^H on the int8 tensor cores)
cudaError_t own_prepare_fourier(cudaStream_t st, const OwnGeom& g, const cuDoubleComplex* Fp, const OwnBufs& b, int ozaki);
// One chunk of Rc pixels of one elevation: out[r][d][f] = complex64(prior_d window_f (A_r^-1 R_r^T)[d][f]) with
// A_r = H_r G H_r^H + N_r (powerspec.py:288-316); info[r] = 1 where A_r is not positive definite.
cudaError_t own_chunk(cudaStream_t st, const OwnGeom& g, const OwnBufs& b, int Rc, int sm_count, const cuDoubleComplex* FSFT,
                      const double* beam_power_el, const void* K, int k_cplx, const void* C, int c_cplx, const unsigned char* M,
                      const cuDoubleComplex* rn, const double* prior, const double* window, float2* outd, int* info,
                      int64_t* launches, int ozaki);

// The tile GEMM by itself (used by the NRML objective, nrml_own.cu): C[b] = A[b] B[b] (nn = 1; A [M][K] pitch lda,
// B [K][N] pitch ldb) or A[b] B[b]^T (nn = 0; B [N][K], both with pitch lda) over mt x nt output tiles of 128 x 128, K =
// 32 kchunks; every dimension a multiple of 128 / 32 (zero-padded buffers).  lower_only: C holds complex entries
// interleaved along its rows and only the tiles that touch its lower triangle are formed.
cudaError_t own_gemm(cudaStream_t st, int sm_count, const double* A, long long sA, int lda, const double* B, long long sB, int ldb,
                     double* C, long long sC, int ldc, int mt, int nt, int kchunks, int nn, int lower_only, int batch);

// The int8 tensor-core GEMM by itself (ozaki_gemm.cuh) on device-resident fp64 operands: C[b] = A[b] B[b]^T, A [M][K],
// B [N][K], C [M][N] row-major and dense; M % 128 == N % 64 == K % 128 == 0.  Scratch: slice bytes / exponent ints from
// own_gemm_ozaki_sizes.
struct OzakiScratch {
  size_t qa, qb;  // bytes
  size_t ea, eb;  // ints
};
OzakiScratch own_gemm_ozaki_sizes(int M, int N, int K, int batch);
cudaError_t own_gemm_ozaki(cudaStream_t st, int sm_count, const double* A, const double* B, double* C, int M, int N, int K, int batch,
                           int8_t* qa, int8_t* qb, int* ea, int* eb);
// One product with row pitches (the drop-in for own_gemm(..., nn = 0, ..., batch = 1) of the NRML objective): C = A B^T, A [M][K]
// pitch lda, B [N][K] pitch ldb, C [M][N] pitch ldc; lower_only as in own_gemm.  Scratch sized by own_gemm_ozaki_sizes(M, N, K, 1).
cudaError_t own_gemm_ozaki_pitched(cudaStream_t st, int sm_count, const double* A, int lda, const double* B, int ldb, double* C, int ldc,
                                   int M, int N, int K, int lower_only, int8_t* qa, int8_t* qb, int* ea, int* eb);

}  // namespace dgw
