// Projection of the diagonal-branch baselines:  yhat[t, :] = F^T (h z[t])  (delay.py:1800-1818 with the draws of
// the diagonal shortcut taken out: they are added analytically by the diagonal sampler).
//
// One masked inverse DFT per sample as the Stockham FFT of fft_fused.cuh (two real samples per complex FFT in real
// mode): the Hermitian extension of h z is staged once per channel (complex mode: the first radix-8 stage takes h z
// straight from global memory), the last stage writes y straight to yhat, and the twiddles come from per-stage tables
// with coalesced loads.  Unlike the dense
// sampler's fused prepass this kernel has no factorisation to share the SM with, so a CTA keeps only the two FFT
// buffers (2 N double2) and several CTAs are resident per SM: the FFT stages of one overlap the global loads / stores
// of the others.  HBM-bound work: reads T*M*16 B, writes T*npad*8 B per baseline.
#include "dense_core.cuh"
#include "fft_fused.cuh"

namespace dg {

constexpr int PTHREADS = 256;  // 8 warps per CTA (512 threads measured slower: 3.3 vs 2.6 ms)
// Resident CTAs per SM the register allocation must allow.  Without a bound ptxas takes 98 registers: TWO CTAs per SM
// at N = 2048 (ncu r02p: occupancy limited by registers and shared memory alike, 16 warps, long-scoreboard stalls
// 4.8 per issue) although three 64 KB buffers fit; bounded to 3 it takes 79, no spills: 1.83 -> 1.47 ms for 2048
// cfg3 baselines (profiles/r04c_proj_variants.log).  0 = no bound.
#ifndef DGIBBS_PROJ_MINB
#define DGIBBS_PROJ_MINB 3
#endif
#if DGIBBS_PROJ_MINB > 0
#define DGIBBS_PROJ_BOUNDS __launch_bounds__(PTHREADS, DGIBBS_PROJ_MINB)
#else
#define DGIBBS_PROJ_BOUNDS __launch_bounds__(PTHREADS)
#endif

__global__ void DGIBBS_PROJ_BOUNDS diag_project_fft_kernel(DevBatch d, int ppb, long long nwork) {
  extern __shared__ __align__(16) double2 fsm[];
  double2* bufA = fsm;
  double2* bufB = fsm + d.N;
  const int N = d.N, nsys = d.nsys, ypitch = d.ypitch, Kpad = d.Kpad;
  const int step = d.cplx ? 1 : 2;
  for (long long w = blockIdx.x; w < nwork; w += gridDim.x) {
    const int li = (int)(w / ppb), t0 = (int)(w % ppb) * step;
    const int b = d.blist[li];
    const int Tb = d.T[b];
    if (t0 >= Tb || d.failed[b]) continue;  // CTA-uniform
    // The whitened samples are formed on the fly from the raw data with the setup kernel's per-channel factors, in
    // the setup kernel's order of operations (z = ((v w) sqrt(Ni_r)), u = h z): same bits as going through zT, which
    // is never written for these baselines (DevBatch::skip_zT_diag).
    const double2* h2 = reinterpret_cast<const double2*>(d.h + (size_t)b * Kpad);
    const double2* wf = d.wsre + (size_t)b * (Kpad / 2);
    const int* cidx = d.cidx + (size_t)b * d.nch;
    const int Mb = d.M[b];
    const bool ident = d.cident[b] != 0;  // kept channels are 0 .. Mb-1: no dependent index load (masked ones have h = 0)
    const long long da = d.data_off[b] + (long long)t0 * Mb, db = da + Mb;  // element offsets of the two samples
    const int c64 = d.data_c64;
    double* yo = d.yhat + ((size_t)d.ycompact[b] * d.Tmax + t0) * ypitch;
    const bool two = !d.cplx && t0 + 1 < Tb;
    __syncthreads();  // the previous transform's last stage is done with the buffers
    if (!d.cplx) {
      // Stage the Hermitian extension A + iB of the two samples: a thread forms u of channel i once and writes both
      // element i and element N - i (half the global loads of taking them per transform input); PU channels in flight.
      const int half = N >> 1;
      constexpr int PU = 4;
      for (int i0 = threadIdx.x; i0 <= half; i0 += PU * PTHREADS) {
        double2 U0[PU], U1[PU];
#pragma unroll
        for (int j = 0; j < PU; j++) {
          const int i = i0 + j * PTHREADS;
          U0[j] = U1[j] = make_double2(0.0, 0.0);
          int ci = -1;
          if (i <= half) ci = ident ? (i < Mb ? i : Mb - 1) : cidx[i];
          if (ci >= 0) {
            const double2 hh = h2[i], f = wf[i], va = ld_vis(d.data, c64, da + ci);
            const double fim = (i == 0 || i == half) ? 0.0 : f.y;  // delay.py:1692: DC and Nyquist are real
            double2 vb = make_double2(0.0, 0.0);
            if (two) vb = ld_vis(d.data, c64, db + ci);
            U0[j] = make_double2(hh.x * ((va.x * f.x) * f.y), hh.y * ((va.y * f.x) * fim));
            U1[j] = make_double2(hh.x * ((vb.x * f.x) * f.y), hh.y * ((vb.y * f.x) * fim));
          }
        }
#pragma unroll
        for (int j = 0; j < PU; j++) {
          const int i = i0 + j * PTHREADS;
          if (i > half) break;
          if (i == 0 || i == half) {
            bufA[i] = make_double2(U0[j].x, U1[j].x);
          } else {
            bufA[i] = make_double2(0.5 * (U0[j].x - U1[j].y), 0.5 * (U0[j].y + U1[j].x));      // (U0 + i U1) / 2
            bufA[N - i] = make_double2(0.5 * (U0[j].x + U1[j].y), 0.5 * (U1[j].x - U0[j].y));  // (conj U0 + i conj U1) / 2
          }
        }
      }
      __syncthreads();
      fft_inverse_fused<PTHREADS>(
          bufA, bufB, d.stwtab, N, [&](int p) { return bufA[p]; }, []() {},
          [&](int k, double2 v) {
            yo[k] = v.x;
            if (two) yo[ypitch + k] = v.y;
          });
      for (int k = nsys + threadIdx.x; k < ypitch; k += PTHREADS) {  // rows are zero beyond nsys
        yo[k] = 0.0;
        if (two) yo[ypitch + k] = 0.0;
      }
    } else {
      fft_inverse_fused<PTHREADS>(
          bufA, bufB, d.stwtab, N,
          [&](int p) {
            const int ci = ident ? (p < Mb ? p : Mb - 1) : cidx[p];
            if (ci < 0) return make_double2(0.0, 0.0);
            const double2 hh = h2[p], f = wf[p], va = ld_vis(d.data, c64, da + ci);
            return make_double2(hh.x * ((va.x * f.x) * f.y), hh.y * ((va.y * f.x) * f.y));
          },
          []() {}, [&](int k, double2 v) { *reinterpret_cast<double2*>(yo + 2 * k) = v; });
      for (int k = nsys + threadIdx.x; k < ypitch; k += PTHREADS) yo[k] = 0.0;
    }
  }
}

cudaError_t launch_build_stage_twiddles(int N, double2* tw, cudaStream_t st) { return build_stage_twiddles(tw, N, st); }

// false: this geometry is not served (the caller falls back to the dense kernel's projection phase)
bool diag_project_fft_fits(const DevBatch& d) { return d.use_fft && d.stwtab != nullptr && (size_t)d.N * 32 <= 200 * 1024; }

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
