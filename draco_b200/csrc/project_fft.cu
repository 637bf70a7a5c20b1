// Projection of the diagonal-branch baselines:  yhat[t, :] = F^T (h z[t])  (delay.py:1800-1818 with the draws of
// the diagonal shortcut taken out: they are added analytically by the diagonal sampler).
//
// One masked inverse DFT per sample, as the in-smem Stockham FFT of dense_core.cuh (two real samples per complex
// FFT in real mode).  Unlike the dense sampler's fused prepass this kernel has no factorisation to share the SM with,
// so a CTA keeps only the two FFT buffers (2 N double2) and several CTAs are resident per SM: the FFT stages of one
// overlap the global loads / stores of the others.  HBM-bound work: reads T*M*16 B, writes T*npad*8 B per baseline.
// The arithmetic is the same as project_rows_fft's in DGIBBS_RNG_ZERO mode, so both routes give identical bits.
#include "dense_core.cuh"

namespace dg {

constexpr int PTHREADS = 256;  // 3 CTAs of 8 warps per SM at N = 2048 (512 threads measured slower: 3.3 vs 2.6 ms)

__global__ void __launch_bounds__(PTHREADS) diag_project_fft_kernel(DevBatch d, int ppb, long long nwork) {
  extern __shared__ __align__(16) double2 fsm[];
  double2* bufA = fsm;
  double2* bufB = fsm + d.N;
  const int N = d.N, nsys = d.nsys, ypitch = d.ypitch, nch = d.nch, Kpad = d.Kpad;
  const int step = d.cplx ? 1 : 2;
  for (long long w = blockIdx.x; w < nwork; w += gridDim.x) {
    const int li = (int)(w / ppb), t0 = (int)(w % ppb) * step;
    const int b = d.blist[li];
    const int Tb = d.T[b];
    if (t0 >= Tb || d.failed[b]) continue;  // CTA-uniform
    const double* h = d.h + (size_t)b * Kpad;
    const double* zT = d.zT + (size_t)b * d.Tmax * Kpad + (size_t)t0 * Kpad;
    double* yo = d.yhat + ((size_t)d.ycompact[b] * d.Tmax + t0) * ypitch;
    const bool two = !d.cplx && t0 + 1 < Tb;
    __syncthreads();  // the previous epilogue is done with the buffers
    // masked / padded channels carry h = 0 and z = 0 (setup kernel), so no channel-index lookup is needed; the global
    // loads of PU channels are issued before the first use
    constexpr int PU = 4;
    if (!d.cplx) {
      for (int i0 = threadIdx.x; i0 < nch; i0 += PU * PTHREADS) {
        double2 hh[PU], za[PU], zb[PU];
#pragma unroll
        for (int j = 0; j < PU; j++) {
          const int i = i0 + j * PTHREADS;
          hh[j] = za[j] = zb[j] = make_double2(0.0, 0.0);
          if (i < nch) {
            hh[j] = *reinterpret_cast<const double2*>(h + 2 * i);
            za[j] = *reinterpret_cast<const double2*>(zT + 2 * i);
            if (two) zb[j] = *reinterpret_cast<const double2*>(zT + Kpad + 2 * i);
          }
        }
#pragma unroll
        for (int j = 0; j < PU; j++) {
          const int i = i0 + j * PTHREADS;
          if (i >= nch) break;
          const double2 U0 = make_double2(hh[j].x * za[j].x, hh[j].y * za[j].y);
          const double2 U1 = make_double2(hh[j].x * zb[j].x, hh[j].y * zb[j].y);
          if (i == 0 || i == N / 2) {
            bufA[i] = make_double2(U0.x, U1.x);
          } else {
            bufA[i] = make_double2(0.5 * (U0.x - U1.y), 0.5 * (U0.y + U1.x));      // (U0 + i U1) / 2
            bufA[N - i] = make_double2(0.5 * (U0.x + U1.y), 0.5 * (U1.x - U0.y));  // (conj U0 + i conj U1) / 2
          }
        }
      }
    } else {
      for (int i0 = threadIdx.x; i0 < N; i0 += PU * PTHREADS) {
        double2 hh[PU], za[PU];
#pragma unroll
        for (int j = 0; j < PU; j++) {
          const int i = i0 + j * PTHREADS;
          hh[j] = za[j] = make_double2(0.0, 0.0);
          if (i < N) {
            hh[j] = *reinterpret_cast<const double2*>(h + 2 * i);
            za[j] = *reinterpret_cast<const double2*>(zT + 2 * i);
          }
        }
#pragma unroll
        for (int j = 0; j < PU; j++) {
          const int i = i0 + j * PTHREADS;
          if (i < N) bufA[i] = make_double2(hh[j].x * za[j].x, hh[j].y * za[j].y);
        }
      }
    }
    __syncthreads();
    const double2* res = fft_inverse_smem<PTHREADS>(bufA, bufB, d.twtab, N);
    if (!d.cplx) {
      for (int p = threadIdx.x; p < ypitch / 2; p += PTHREADS) {
        const int k = 2 * p;
        double2 ya = make_double2(0.0, 0.0), yb = ya;
        if (k < nsys) {
          const double2 r0 = res[k], r1 = res[k + 1];
          ya = make_double2(r0.x, r1.x);
          yb = make_double2(r0.y, r1.y);
        }
        *reinterpret_cast<double2*>(yo + k) = ya;
        if (two) *reinterpret_cast<double2*>(yo + ypitch + k) = yb;
      }
    } else {
      for (int p = threadIdx.x; p < ypitch / 2; p += PTHREADS) {
        double2 ya = make_double2(0.0, 0.0);
        if (2 * p < nsys) ya = res[p];
        *reinterpret_cast<double2*>(yo + 2 * p) = ya;
      }
    }
  }
}

// false: this geometry is not served (the caller falls back to the dense kernel's projection phase)
bool diag_project_fft_fits(const DevBatch& d) { return d.use_fft && d.twtab != nullptr && (size_t)d.N * 32 <= 200 * 1024; }

cudaError_t launch_diag_project_fft(const DevBatch& d, int sm_count, cudaStream_t st) {
  if (d.nlist <= 0) return cudaSuccess;
  const size_t smem = (size_t)d.N * 2 * sizeof(double2);
  cudaError_t e = cudaFuncSetAttribute(diag_project_fft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int occ = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, diag_project_fft_kernel, PTHREADS, smem);
  if (e != cudaSuccess) return e;
  if (occ < 1) return cudaErrorInvalidConfiguration;
  const int ppb = d.cplx ? d.Tmax : (d.Tmax + 1) / 2;
  const long long nwork = (long long)d.nlist * ppb;
  long long grid = (long long)occ * sm_count;
  if (grid > nwork) grid = nwork;
  diag_project_fft_kernel<<<(unsigned)grid, PTHREADS, smem, st>>>(d, ppb, nwork);
  return cudaGetLastError();
}

}  // namespace dg
