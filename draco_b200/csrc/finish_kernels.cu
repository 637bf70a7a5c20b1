// End of a batch: median of the trailing samples + fftshift, front-padding of short chains.
// Reference: DelayPowerSpectrumGibbs._estimator, draco/analysis/delay.py:1256-1265 and
// DelayPowerSpectrumBase._evaluate, delay.py:1175-1178.
#include "common.cuh"
#include "internal.h"

namespace dg {

// one thread per (baseline, delay)
__global__ void finish_kernel(DevBatch d, const double* __restrict__ S0, int ntail, double* __restrict__ spectrum,
                              int shift_samples, int* __restrict__ n_done_out, unsigned char* __restrict__ success_out,
                              double* __restrict__ final_S) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)d.B * d.N) return;
  const int b = (int)(idx / d.N), k = (int)(idx % d.N);
  const int N = d.N, niter = d.niter;
  const int nd = d.n_done[b];
  const size_t stride = (size_t)d.B * N;
  const double* col = d.samples + (size_t)b * N + k;

  if (k == 0) {
    if (n_done_out) n_done_out[b] = nd;
    // an empty chain is a failure even when nothing failed numerically (nsamp = 0): `if samples: ... else: success = False`,
    // delay.py:1261-1265
    if (success_out) success_out[b] = (nd == niter && nd > 0 && !d.failed[b]) ? 1 : 0;
  }
  if (final_S) final_S[idx] = d.S[idx];

  if (nd == 0) {
    spectrum[idx] = S0[idx];  // estimator failed at the first sample: the prior, unshifted (delay.py:1261-1265)
  } else {
    // samples[-ntail:] of a list of nd entries; ntail == 0 selects the whole list (delay.py:1259)
    int m = (ntail == 0 || ntail > nd) ? nd : ntail;
    int start = nd - m;
    // order statistics (m-1)/2 and m/2 by rank counting
    const int k_lo = (m - 1) >> 1, k_hi = m >> 1;
    double v_lo = 0.0, v_hi = 0.0;
    for (int i = 0; i < m; i++) {
      double v = col[(size_t)(start + i) * stride];
      int less = 0, eq = 0;
      for (int j = 0; j < m; j++) {
        double u = col[(size_t)(start + j) * stride];
        less += (u < v);
        eq += (u == v);
      }
      if (less <= k_lo && k_lo < less + eq) v_lo = v;
      if (less <= k_hi && k_hi < less + eq) v_hi = v;
    }
    double med = (k_lo == k_hi) ? v_lo : 0.5 * (v_lo + v_hi);
    int ks = k + N / 2;  // np.fft.fftshift
    if (ks >= N) ks -= N;
    spectrum[(size_t)b * N + ks] = med;
  }

  if (shift_samples && nd < niter) {
    // spectrum_samples[-nd:, b] = samples, zeros in front (delay.py:1177-1178); in place, back to front
    double* w = d.samples + (size_t)b * N + k;
    const int pad = niter - nd;
    for (int r = niter - 1; r >= 0; r--) w[(size_t)r * stride] = (r >= pad) ? w[(size_t)(r - pad) * stride] : 0.0;
  }
}

cudaError_t launch_finish(const DevBatch& d, const double* S0, double median_frac, double* spectrum,
                          double* samples_out, int* n_done_out, unsigned char* success_out, double* final_S,
                          cudaStream_t st) {
  long long total = (long long)d.B * d.N;
  int ntail = (int)((double)d.niter * median_frac);  // int(nsamp * median_frac), delay.py:1256
  int blocks = (int)((total + 255) / 256);
  finish_kernel<<<blocks, 256, 0, st>>>(d, S0, ntail, spectrum, samples_out != nullptr, n_done_out, success_out,
                                         final_S);
  return cudaGetLastError();
}

}  // namespace dg
