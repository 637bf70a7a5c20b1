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
};
OwnSizes own_sizes(const OwnGeom& g);

struct OwnBufs {
  double *Pr, *Pext, *T1r, *Ar, *Yr;  // [R][...]; must be zero-filled once after allocation (the padding stays zero)
  double *Gext, *Fpr, *work, *Wp;
};

cudaError_t own_configure();
// Fpr (zero-padded row-major copy of F'^T: [nd][F] complex) from the column-major F' [F, nd]
cudaError_t own_prepare_fourier(cudaStream_t st, const OwnGeom& g, const cuDoubleComplex* Fp, const OwnBufs& b);
// One chunk of Rc pixels of one elevation: out[r][d][f] = complex64(prior_d window_f (A_r^-1 R_r^T)[d][f]) with
// A_r = H_r G H_r^H + N_r (powerspec.py:288-316); info[r] = 1 where A_r is not positive definite.
cudaError_t own_chunk(cudaStream_t st, const OwnGeom& g, const OwnBufs& b, int Rc, int sm_count, const cuDoubleComplex* FSFT,
                      const double* beam_power_el, const void* K, int k_cplx, const void* C, int c_cplx, const unsigned char* M,
                      const cuDoubleComplex* rn, const double* prior, const double* window, float2* outd, int* info,
                      int64_t* launches);

}  // namespace dgw
