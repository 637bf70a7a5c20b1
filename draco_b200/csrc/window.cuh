// Generalised cosine-sum windows (draco/util/tools.py:547-601) and exact twiddles, shared by the set-up kernels and
// the plain FFT delay transform.  The coefficient table is per translation unit (static __constant__).
#pragma once

namespace dg {

// cosine-sum coefficients, draco/util/tools.py:584-595 (digit for digit)
static __constant__ double c_window[8][4] = {
    {1, 0, 0, 0},
    {0.5, -0.5, 0, 0},
    {0.5, -0.5, 0, 0},
    {0.53836, -0.46164, 0, 0},
    {0.42, -0.5, 0.08, 0},
    {0.355768, -0.487396, 0.144232, -0.012604},
    {0.3635819, -0.4891775, 0.1365995, -0.0106411},
    {0.35875, -0.48829, 0.14128, -0.01168},
};

__device__ __forceinline__ double window_value(int window_id, double x) {
  if (window_id < 0) return 1.0;
  double w = 0.0;
#pragma unroll
  for (int j = 0; j < 4; j++) w += c_window[window_id][j] * cos(2.0 * 3.141592653589793 * (double)j * x);
  return (x >= 0.0 && x <= 1.0) ? w : 0.0;
}

// exact twiddle: cos/sin(2 pi m / N) with the integer phase m already reduced mod N
__device__ __forceinline__ void twiddle(long long m, int N, double& c, double& s) {
  sincospi(2.0 * (double)m / (double)N, &s, &c);
}

}  // namespace dg
