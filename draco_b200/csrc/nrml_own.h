// Own-kernel path of the NRML objective (nrml_own.cu); internal interface used by nrml.cu.
#pragma once
#include <cuComplex.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace dgn {

struct Own {
  int M = 0, N = 0;
  int LDM = 0, LDN = 0;  // real pitches: 2 M / 2 N rounded up to 128
  int MR = 0, NR = 0;    // M / N rounded up to 128
  int nblk = 0;          // LDM / 128 block columns of the embedded system
  int rtU = 0, rtX = 0;  // right-hand-side row tiles: columns of U, columns of X
  int sm_count = 0;
  double *MFm = nullptr, *Um = nullptr, *UmExt = nullptr, *Q = nullptr, *T = nullptr, *XCemb = nullptr, *UrowExt = nullptr,
         *UtrowExt = nullptr, *Wrow = nullptr, *G = nullptr, *K = nullptr, *Wpg = nullptr, *logl = nullptr;
  int* flag = nullptr;
  // the products C = U U^H, W = (X - C) Ut, G = U^H Ut, K = Ut^H W on the INT8 tensor cores (ozaki_gemm.cuh) unless
  // DGNRML_NO_INT8 is set when the objective is created: slices / exponents of one A and one B operand
  int int8 = 0;
  int8_t *qa = nullptr, *qb = nullptr;
  int *ea = nullptr, *eb = nullptr;
  // views into T
  double* Ut() const { return T + (size_t)nblk * 128 * LDM; }                 // [NR][LDM] rows = columns of C^-1 U
  double* CiX() const { return T + (size_t)(nblk + rtU) * 128 * LDM; }        // [MR][LDM] rows = columns of C^-1 X
  int ldG() const { return LDN / 2; }                                         // complex pitch of G / K ([b][a] = G(a, b))
};

cudaError_t own_alloc(Own& o, int M, int N, int sm_count, cudaStream_t st);
void own_free(Own& o);
// MF as the host holds it (row-major [M, N] complex128), already on the device
cudaError_t own_set_mf(Own& o, const cuDoubleComplex* MF_rowmajor, cudaStream_t st);
// _precompute (delayopt.py:120-164) at sq = sqrt(S): factor C = U U^H + N, Ut = C^-1 U, C^-1 X, W = (X - C) Ut,
// t_a = -Re(Ut_a . W_a).  U_colmajor: the column-major [M, N] U (its memory = the rows to solve for);
// X_colmajor: [M, M].  *info_host: 0, or 1 + the block column whose pivot failed.  scal_dev[0] = log det C,
// scal_dev[1] = trace(C^-1 X).  Synchronises the stream once (for the flag).
cudaError_t own_precompute(Own& o, cudaStream_t st, const double* sq, const cuDoubleComplex* U_colmajor,
                           const cuDoubleComplex* X_colmajor, const double* noise, double* t_dev, double* scal_dev,
                           int* info_host, int64_t* launches);
// G = U^H Ut and (exact) K = Ut^H W as [b][a] complex arrays with pitch ldG() (delayopt.py:221-236)
cudaError_t own_hessian_terms(Own& o, cudaStream_t st, int exact, int64_t* launches);

}  // namespace dgn
