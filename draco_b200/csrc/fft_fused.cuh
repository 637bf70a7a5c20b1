// Stockham inverse FFT with fused first / last stages and per-stage twiddle tables, shared by the plain FFT delay
// transform (delay_fft.cu) and the projection of the diagonal-branch baselines (project_fft.cu).
#pragma once
#include "dense_core.cuh"

namespace dg {

// One stage's twiddle tables: t[i * cnt + k] = e^{2 pi i (m_i k s) / N}, k < cnt, i < ntab (exact integer phase)
static __global__ void fft_stage_tw_kernel(double2* __restrict__ t, int N, int cnt, int s, int m0, int m1, int m2, int ntab) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= cnt) return;
  const int mult[3] = {m0, m1, m2};
  for (int i = 0; i < ntab; i++) {
    double c, sn;
    sincospi(2.0 * (double)(((long long)mult[i] * k * s) % N) / (double)N, &sn, &c);
    t[(size_t)i * cnt + k] = make_double2(c, sn);
  }
}

// Tables of every stage after the first, in stage order (the layout fft_inverse_fused walks through).
static cudaError_t build_stage_twiddles(double2* t, int N, cudaStream_t st) {
  int Ns = 1;
  size_t off = 0;
  while (Ns * 8 <= N) {
    if (Ns > 1) {
      fft_stage_tw_kernel<<<(Ns + 255) / 256, 256, 0, st>>>(t + off, N, Ns, N / (8 * Ns), 1, 2, 4, 3);
      off += 3 * (size_t)Ns;
    }
    Ns <<= 3;
  }
  if (Ns * 4 == N)
    fft_stage_tw_kernel<<<(N / 4 + 255) / 256, 256, 0, st>>>(t + off, N, N / 4, 1, 1, 2, 3, 3);
  else if (Ns * 2 == N)
    fft_stage_tw_kernel<<<(N / 2 + 255) / 256, 256, 0, st>>>(t + off, N, N / 2, 1, 1, 0, 0, 1);
  return cudaGetLastError();
}

// Inverse (sign +) Stockham FFT of N = 2^m >= 8 points whose first stage reads its inputs through ld(i) (global
// memory: the windowed row) and whose last stage hands its outputs to st(k, value) (global memory: scaled, shifted)
// -- two passes over shared memory and two barriers fewer than staging the row first.  Same butterflies and swizzle
// as fft_inverse_smem (dense_core.cuh); the twiddles come from per-stage tables laid out contiguously in the butterfly
// index (build_stage_twiddles): a table look-up costs 4 L1 wavefronts per warp, where the strided look-ups into one
// shared e^(2 pi i m / N) table cost up to 32 each and made the LSU data pipe the bound (ncu: 79 % busy, two thirds
// of it twiddle loads); af() runs once the thread's own first-stage inputs are consumed
// (the hook for prefetching the next transform's inputs).  The caller synchronises the CTA between two transforms.
template <int NT, class Load, class AfterFirst, class Store>
__device__ __forceinline__ void fft_inverse_fused(double2* x, double2* y, const double2* stw, int N, Load ld, AfterFirst af,
                                                  Store st) {
  int Ns = 1;
  bool swz = false;
  const int e8 = N >> 3;
  while (Ns * 8 <= N) {
    const bool first = Ns == 1, last = Ns * 8 == N;
    const bool swz_out = first && !last;
    for (int j = threadIdx.x; j < e8; j += NT) {
      const int k = j & (Ns - 1);
      double2 v[8];
      if (first) {
#pragma unroll
        for (int r = 0; r < 8; r++) v[r] = ld(j + r * e8);
      } else {
#pragma unroll
        for (int r = 0; r < 8; r++) v[r] = x[fft_swz(j + r * e8, swz)];
        // w^(k s), w^(2 k s), w^(4 k s) from this stage's table (contiguous in k: coalesced), the rest by products
        const double2 w1 = stw[k], w2 = stw[Ns + k], w4 = stw[2 * Ns + k];
        const double2 w3 = cmul(w1, w2), w5 = cmul(w1, w4), w6 = cmul(w2, w4);
        const double2 w7 = cmul(w3, w4);
        v[1] = cmul(v[1], w1); v[2] = cmul(v[2], w2); v[3] = cmul(v[3], w3); v[4] = cmul(v[4], w4);
        v[5] = cmul(v[5], w5); v[6] = cmul(v[6], w6); v[7] = cmul(v[7], w7);
      }
      dft4_inv(v[0], v[2], v[4], v[6]);
      dft4_inv(v[1], v[3], v[5], v[7]);
      const double h = 0.70710678118654752440;
      const double2 o1 = make_double2((v[3].x - v[3].y) * h, (v[3].x + v[3].y) * h);
      const double2 o2 = make_double2(-v[5].y, v[5].x);
      const double2 o3 = make_double2(-(v[7].x + v[7].y) * h, (v[7].x - v[7].y) * h);
      const int j0 = ((j - k) << 3) + k;
      const double2 r0 = cadd(v[0], v[1]), r1 = cadd(v[2], o1), r2 = cadd(v[4], o2), r3 = cadd(v[6], o3);
      const double2 r4 = csub(v[0], v[1]), r5 = csub(v[2], o1), r6 = csub(v[4], o2), r7 = csub(v[6], o3);
      if (last) {
        st(j0, r0); st(j0 + Ns, r1); st(j0 + 2 * Ns, r2); st(j0 + 3 * Ns, r3);
        st(j0 + 4 * Ns, r4); st(j0 + 5 * Ns, r5); st(j0 + 6 * Ns, r6); st(j0 + 7 * Ns, r7);
      } else {
        y[fft_swz(j0, swz_out)] = r0;
        y[fft_swz(j0 + Ns, swz_out)] = r1;
        y[fft_swz(j0 + 2 * Ns, swz_out)] = r2;
        y[fft_swz(j0 + 3 * Ns, swz_out)] = r3;
        y[fft_swz(j0 + 4 * Ns, swz_out)] = r4;
        y[fft_swz(j0 + 5 * Ns, swz_out)] = r5;
        y[fft_swz(j0 + 6 * Ns, swz_out)] = r6;
        y[fft_swz(j0 + 7 * Ns, swz_out)] = r7;
      }
    }
    if (first) af();  // the first-stage inputs of this thread are consumed
    if (last) return;
    __syncthreads();
    double2* tmp = x; x = y; y = tmp;
    if (!first) stw += 3 * Ns;
    Ns <<= 3;
    swz = swz_out;
  }
  if (Ns * 4 == N) {  // last stage, radix 4
    const int quarter = N >> 2;
    for (int j = threadIdx.x; j < quarter; j += NT) {
      double2 v0 = x[fft_swz(j, swz)], v1 = x[fft_swz(j + quarter, swz)], v2 = x[fft_swz(j + 2 * quarter, swz)],
              v3 = x[fft_swz(j + 3 * quarter, swz)];
      v1 = cmul(v1, stw[j]);
      v2 = cmul(v2, stw[quarter + j]);
      v3 = cmul(v3, stw[2 * quarter + j]);
      dft4_inv(v0, v1, v2, v3);
      st(j, v0); st(j + quarter, v1); st(j + 2 * quarter, v2); st(j + 3 * quarter, v3);
    }
  } else {  // last stage, radix 2
    const int half = N >> 1;
    for (int j = threadIdx.x; j < half; j += NT) {
      const double2 v0 = x[fft_swz(j, swz)], v1 = cmul(x[fft_swz(j + half, swz)], stw[j]);
      st(j, cadd(v0, v1));
      st(j + half, csub(v0, v1));
    }
  }
}

}  // namespace dg
