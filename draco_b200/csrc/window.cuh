// Generalised cosine-sum windows (draco/util/tools.py:547-601) and exact twiddles, shared by the set-up kernels and
// the plain FFT delay transform.  The coefficient table is per translation unit (static __constant__).
#pragma once
#include "common.cuh"

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

// window_id: -1 none, 0..7 the table above, 8 "triangular", DG_WINDOW_TUKEY_BASE + r * 1e6 "tukey-r" (tools.py:566-581;
// include/dgibbs.h)
__device__ __forceinline__ double window_value(int window_id, double x) {
  if (window_id < 0) return 1.0;
  double w;
  if (window_id == DG_WINDOW_TRIANGULAR) {
    w = 1.0 - 2.0 * fabs(x - 0.5);
  } else if (window_id >= DG_WINDOW_TUKEY_BASE) {
    const double r = (double)(window_id - DG_WINDOW_TUKEY_BASE) / 1e6;  // the correctly rounded double of "0.X"
    const double alpha = 0.5 * r;
    w = 1.0;
    if (x < alpha) w = 0.5 * (1.0 + cos(3.141592653589793 * (x - alpha) / alpha));
    if (x >= 1.0 - alpha) w = 0.5 * (1.0 + cos(3.141592653589793 * (x - (1.0 - alpha)) / alpha));
  } else {
    w = 0.0;
#pragma unroll
    for (int j = 0; j < 4; j++) w += c_window[window_id][j] * cos(2.0 * 3.141592653589793 * (double)j * x);
  }
  return (x >= 0.0 && x <= 1.0) ? w : 0.0;
}

// exact twiddle: cos/sin(2 pi m / N) with the integer phase m already reduced mod N
__device__ __forceinline__ void twiddle(long long m, int N, double& c, double& s) {
  sincospi(2.0 * (double)m / (double)N, &s, &c);
}

}  // namespace dg
