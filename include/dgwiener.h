/* dgwiener.h -- C ABI of the Wiener delay-transform operator on the GPU (part of libdgnrml.so).
 *
 * SURVEY 8 row f4.  Replaces the loops over polarisation / elevation / right ascension of
 * ConstructWienerDelayTransform.process, draco/analysis/powerspec.py:229-319: for every pixel the signal + noise
 * covariance in the masked, filtered frequency domain
 *     A = H (F S F^H o b b^T) H^H + C o r r^T,   H = diag(M) K,
 * its Cholesky factorisation (scipy.linalg.cho_factor / cho_solve there) and the projection
 *     D = S (F^H H^H A^-1) window   ->  [delay, freq]
 * that takes a filtered, masked map spectrum to delay space.  One call builds the operators of ONE polarisation
 * (the unit over which the reference gathers the filter and the covariance, powerspec.py:251-252).  The task class
 * (draco_b200/powerspec.py) keeps the container handling, the window, the delay axis and the prior on the host,
 * exactly as the reference does.
 *
 * All pointers are HOST pointers to C-contiguous (numpy row-major) arrays owned by the caller; nothing is
 * retained after the call returns.
 *
 * Return codes: 0 ok; 1 = a covariance is not positive definite (scipy raises LinAlgError there,
 * powerspec.py:296; a failing pixel is reported through bad_el / bad_ra);
 * <0 bad argument (its index); >1 CUDA / library error (dgwiener_last_error).
 */
#ifndef DGWIENER_H
#define DGWIENER_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int32_t device_id;
  int32_t nfreq;            /* F: channels of the map                                        */
  int32_t ndelay;           /* delays kept (tau >= 0), powerspec.py:214                      */
  int32_t nra, nel;         /* pixels: right ascension, local elevations                     */
  int32_t filter_is_complex; /* 0: filter is float64, 1: complex128 (RingMap "complex_filter") */
  int32_t cov_is_complex;    /* 0: freq_cov is float64, 1: complex128                          */
  int32_t solver;           /* 0 auto (= 4 from F = 256 when the budget holds a full wave of pixels, else 3), 3 this
                               library's own DMMA kernels for every contraction,
                               the factorisation and the solves; 4 as 3 with the three contractions on the INT8
                               tensor cores (Ozaki split of the fp64 operands into 8 signed 7-bit slices,
                               tcgen05.mma kind::i8 with int32 accumulators in TMEM: fp64-equivalent results);
                               1 / 2 cuBLAS + cuSOLVER (batched factorisation / per matrix on several streams) */
  const void* fourier;      /* [F, ndelay] complex128: exp(2 pi i freq tau) / sqrt(ntau), powerspec.py:229 */
  const double* prior;      /* [ndelay] diagonal signal prior, powerspec.py:233              */
  const double* window;     /* [F] spectral window (zero = channel not used), powerspec.py:207 */
  const void* filter;       /* [nra, F, F]  K, powerspec.py:252 after its transpose          */
  const void* freq_cov;     /* [nra, F, F]  C, powerspec.py:251                              */
  const double* weight;     /* [nel, nra, F] map weights, powerspec.py:265                   */
  const double* beam_power; /* [nel, F] dirty-beam power (its square root scales the prior), powerspec.py:266 */
  void* out;                /* complex64; operator of pixel (ra, el) = [ndelay, F] block at out + 8 * (ra * out_stride_ra + el * out_stride_el) */
  int64_t out_stride_ra, out_stride_el; /* in complex64 elements                               */
  int64_t workspace_bytes;  /* device memory budget for one chunk of pixels (0: 32 GiB, at most half of the free memory) */
  int32_t out_on_device;    /* 1: out is a device pointer (same strides): the operator never visits the host */
  int32_t reserved;
  /* results */
  int32_t bad_el, bad_ra;   /* a pixel whose factorisation failed (return code 1)            */
  double ms_device;         /* CUDA-event time of the whole call on the device, copies included */
  double ms_compute;        /* sum of the CUDA-event times of the kernels + library calls alone */
  double flop;              /* algorithmic fp64 flop of the call (complex MAC = 8)           */
  int64_t launches;         /* kernels + library calls launched                              */
} dgwiener_problem;

int dgwiener_construct(dgwiener_problem* p);

/* Apply stored operators to a ring map: ApplyWienerDelayTransform.process, powerspec.py:430-458, for ONE
 * polarisation: for every pixel (ra, el)
 *     spectrum[el, ra, :] = op[ra, el] @ map[:, ra, el]                                  (powerspec.py:446)
 *     weight[el, ra, :]   = invert_no_zero(|op[ra, el]|^2 @ invert_no_zero(w[:, ra, el])) (powerspec.py:449-452)
 * one hand-written kernel, bound by the single read of the complex64 operator from HBM.  map / weight are the
 * container's own [freq, ra, el] slices; the outputs are the [el, ra, delay] block of the DelayTransform
 * container's [baseline = pol * nel + el, sample = ra, delay] datasets. */
typedef struct {
  int32_t device_id;
  int32_t nfreq, ndelay, nra, nel;
  int32_t op_on_device;     /* 1: op is a device pointer (e.g. written by dgwiener_construct with out_on_device) */
  const void* op;           /* complex64; [ndelay, F] block of pixel (ra, el) at op + 8 * (ra * op_stride_ra + el * op_stride_el) */
  int64_t op_stride_ra, op_stride_el;
  const double* map;        /* [F, nra, nel] host */
  const double* weight;     /* [F, nra, nel] host */
  void* spectrum;           /* [nel, nra, ndelay] complex128 host */
  float* spectrum_weight;   /* [nel, nra, ndelay] float32 host */
  int64_t workspace_bytes;  /* device budget for one chunk of host-resident operators (0: 2 GiB) */
  /* results */
  double ms_kernel;         /* CUDA-event time of the apply kernel launches alone */
  double bytes;             /* algorithmic bytes: 8 ndelay F + 16 F + 20 ndelay per pixel */
  int64_t launches;
} dgwiener_apply_problem;

int dgwiener_apply(dgwiener_apply_problem* p);

/* The contraction kernel of solver 4 by itself:  C[b] = A[b] B[b]^T  on the INT8 tensor cores (fp64 operands split into 8
 * signed 7-bit slices per row and block of 256 columns, tcgen05.mma kind::i8, int32 accumulators in TMEM; fp64-equivalent
 * results: csrc/ozaki_gemm.cuh).  It stands where the reference calls numpy's matmul on complex128 operands
 * (draco/analysis/powerspec.py:288-304: H (FSF^H o bb^T) H^H and F^H H^H); exported so that tests hold the kernel BIT FOR BIT to
 * the host emulation of its integer arithmetic (oracle/ozaki_emulation.py, tests/test_gpu_ozaki_gemm.py).
 * A [batch][M][K], B [batch][N][K], C [batch][M][N]: float64, C-contiguous HOST pointers; M % 128 == N % 64 == K % 128 == 0
 * (-5 otherwise).  ms_device (may be NULL): device time of split + GEMM. */
int dgwiener_gemm_ozaki(int32_t device_id, const double* A, const double* B, double* C, int32_t M, int32_t N, int32_t K, int32_t batch,
                        double* ms_device);
int dgwiener_sizeof(void);
int dgwiener_apply_sizeof(void);
const char* dgwiener_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* DGWIENER_H */
