/* dgnrml.h -- C ABI of the NRML / maximum-posterior likelihood object on the GPU (libdgnrml.so).
 *
 * SURVEY 8 row f4.  Replaces the linear algebra of draco/analysis/delayopt.py:61-236 (class LogLikePS:
 * _precompute, value, gradient, hessian), which delay_power_spectrum_maxpost (delayopt.py:424-561) evaluates
 * inside scipy.optimize.minimize(method="Newton-CG").  The optimiser's control flow stays on the host exactly as
 * in the reference (scipy is the reference's dependency); every evaluation of the objective -- the M x M complex
 * Hermitian covariance C = MF diag(S) MF^H + N, its Cholesky factor, C^-1 U, the trace term, the gradient and the
 * N x N Hessian -- runs on the device.
 *
 * The constructor takes exactly what LogLikePS.__init__ takes after its frequency selection (delayopt.py:93-118):
 * the data covariance X, the masked Fourier matrix MF and the noise variances, all host arrays in numpy's
 * row-major layout.  libdgnrml.so is separate from libdgibbs.so because it links cuBLAS / cuSOLVER (plain
 * library ZHERK / ZGEMM / ZPOTRF / ZPOTRS; everything else is hand-written CUDA); the Gibbs path has no such
 * dependency.
 *
 * Return codes: 0 ok; 1 = the covariance is not positive definite (scipy raises LinAlgError there,
 * delayopt.py:530-548); <0 bad argument; >1 CUDA / library error (dgnrml_last_error).
 */
#ifndef DGNRML_H
#define DGNRML_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dgnrml_s* dgnrml;

enum { DGNRML_VALUE = 1, DGNRML_GRADIENT = 2, DGNRML_HESSIAN = 4 };

/* M retained channels, N delays (= parameters), nsamp = samples behind X (delayopt.py:114).
 * X [M, M] complex128, MF [M, N] complex128, noise [M] float64 (variances), row-major host arrays.
 * log_lo / log_hi: bounds applied to the parameters x = log S (delayopt.py:118, 137). */
int dgnrml_create(int32_t device_id, int32_t M, int32_t N, int32_t nsamp, int32_t exact_hessian, double log_lo,
                  double log_hi, const void* X, const void* MF, const double* noise, dgnrml* out);

/* Evaluate at x [N] (host).  what = OR of DGNRML_*; value [1], gradient [N], hessian [N, N] row-major (host;
 * may be NULL when not requested).  Repeated calls with the same x reuse the factorisation (delayopt.py:134). */
int dgnrml_eval(dgnrml s, const double* x, int32_t what, double* value, double* gradient, double* hessian);

/* Starting point of the optimiser, delayopt.py:505-506: S0 [N] = (data @ pinv(MF^T, rtol)).var(axis=0) with data [T, M]
 * the windowed samples (host, row-major complex128, the channels MF keeps).  Returns 1 if the eigendecomposition
 * behind the pseudo-inverse fails. */
int dgnrml_initial_guess(dgnrml s, const void* data, int32_t T, double rtol, double* S0);

int dgnrml_destroy(dgnrml s);
const char* dgnrml_last_error(dgnrml s);
/* milliseconds spent on the device by the last dgnrml_eval (CUDA events) and its algorithmic fp64 flop */
int dgnrml_last_timing(dgnrml s, double* ms, double* flop);

#ifdef __cplusplus
}
#endif
#endif /* DGNRML_H */
