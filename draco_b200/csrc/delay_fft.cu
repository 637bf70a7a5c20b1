// Plain (unmasked) delay transform by inverse FFT: delay_spectrum_fft, draco/analysis/delay.py:2102-2130, as used by
// the task DelaySpectrumFFT (delay.py:960-980): every row (one time sample of one baseline, N complex channels) is
// multiplied by the generalised cosine window at x = i / N (delay.py:2124-2127), transformed with numpy's ifft
// convention (sign +, scale 1/N) and, for the task, fftshifted (delay.py:980).
//
// HBM-bound: 16 N bytes read and 16 N bytes written per row.  N = 2^m <= 4096 runs as the fused-stage Stockham FFT of
// fft_fused.cuh (first stage fed from a cp.async-prefetched row, last stage stored straight to HBM, per-stage twiddle
// tables; several CTAs per SM); any other N <= 6400 is a direct O(N^2) DFT against an exact twiddle table in shared
// memory.
#include <cstdio>
#include <mutex>
#include <string>

#include "dense_core.cuh"
#include "fft_fused.cuh"
#include "window.cuh"

#include "../../include/dgibbs.h"

namespace dg {

constexpr int FTHREADS = 256;

__global__ void delay_fft_window_kernel(double* __restrict__ wtab, int N, int window_id) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) wtab[i] = window_value(window_id, (double)i / (double)N);
}

// tw[m] = e^{2 pi i m / N}, all m < N (the direct DFT needs the full circle)
__global__ void delay_fft_fulltw_kernel(double2* __restrict__ tw, int N) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= N) return;
  double c, s;
  twiddle(m, N, c, s);
  tw[m] = make_double2(c, s);
}

// One CTA of NT = clamp(N / 8, 32, 256) threads per row at a time (every thread owns a radix-8 butterfly), as many
// CTAs per SM as three N-point buffers allow.  The row after the current one is prefetched with cp.async into the
// third buffer while the current one is transformed: a thread copies exactly the elements its own first-stage
// butterflies will read (j + r N/8), so its own cp.async.wait_group is all the synchronisation the inputs need, and
// it refills a slot only after it has consumed it.  HBM reads, the shared-memory passes and the stores overlap.
template <int NT>
__global__ void __launch_bounds__(NT) delay_fft_pow2_kernel(const double2* __restrict__ x, double2* __restrict__ y,
                                                            const double* __restrict__ wtab,
                                                            const double2* __restrict__ tw, long long nrows, int N,
                                                            int shift) {
  extern __shared__ __align__(16) double2 fsm[];
  double2* bufA = fsm;
  double2* bufB = fsm + N;
  double2* bufC = fsm + 2 * N;  // raw inputs of the next row
  const double invN = 1.0 / (double)N;
  const int off = shift ? N / 2 : 0;
  const int e8 = N >> 3;
  auto prefetch = [&](long long r) {
    if (r < nrows) {
      const double2* xr = x + r * N;
      for (int j = threadIdx.x; j < e8; j += NT)
#pragma unroll
        for (int q = 0; q < 8; q++) cp_async16(bufC + j + q * e8, xr + j + q * e8);
    }
    cp_async_commit();
  };
  prefetch(blockIdx.x);
  for (long long r = blockIdx.x; r < nrows; r += gridDim.x) {
    double2* yr = y + r * N;
    cp_async_wait<0>();
    __syncthreads();  // the previous row's last stage is done reading bufA / bufB
    fft_inverse_fused<NT>(
        bufA, bufB, tw, N,
        [&](int i) {
          const double2 v = bufC[i];
          const double w = wtab[i];
          return make_double2(v.x * w, v.y * w);
        },
        [&]() { prefetch(r + gridDim.x); },
        [&](int k, double2 v) { yr[(k + off) & (N - 1)] = make_double2(v.x * invN, v.y * invN); });
  }
  cp_async_wait<0>();
}

__global__ void __launch_bounds__(FTHREADS) delay_dft_direct_kernel(const double2* __restrict__ x, double2* __restrict__ y,
                                                                    const double* __restrict__ wtab,
                                                                    const double2* __restrict__ twg, long long nrows, int N,
                                                                    int shift) {
  extern __shared__ __align__(16) double2 fsm[];
  double2* tw = fsm;      // [N]
  double2* xw = fsm + N;  // [N] windowed row
  for (int m = threadIdx.x; m < N; m += FTHREADS) tw[m] = twg[m];
  const double invN = 1.0 / (double)N;
  const int off = shift ? N / 2 : 0;
  for (long long r = blockIdx.x; r < nrows; r += gridDim.x) {
    const double2* xr = x + r * N;
    __syncthreads();
    for (int i = threadIdx.x; i < N; i += FTHREADS) {
      const double2 v = xr[i];
      const double w = wtab[i];
      xw[i] = make_double2(v.x * w, v.y * w);
    }
    __syncthreads();
    double2* yr = y + r * N;
    for (int k = threadIdx.x; k < N; k += FTHREADS) {
      double ar = 0.0, ai = 0.0;
      int m = 0;
      for (int f = 0; f < N; f++) {
        const double2 a = xw[f], t = tw[m];
        ar += a.x * t.x - a.y * t.y;
        ai += a.x * t.y + a.y * t.x;
        m += k;
        if (m >= N) m -= N;
      }
      int ko = k + off;
      if (ko >= N) ko -= N;
      yr[ko] = make_double2(ar * invN, ai * invN);
    }
  }
}

// DelaySpectrumToPowerSpectrum.process (delay.py:1061-1106): ps[b, n] = var_t spec[b, t, n] over the unmasked samples
// (numpy's _var: mean = sum * (1 / count) per component, sum of |x - mean|^2 in sample order, divided by the count);
// with a mask, a baseline without any unmasked sample gives NaN -> 0 and sets the baseline's flag.
__global__ void __launch_bounds__(256) sample_variance_kernel(const double2* __restrict__ spec,
                                                             const unsigned char* __restrict__ mask, double* __restrict__ var,
                                                             unsigned char* __restrict__ bad, int T, int N) {
  const int b = blockIdx.y;
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const double2* x = spec + (size_t)b * T * N + n;
  const unsigned char* mk = mask ? mask + (size_t)b * T : nullptr;
  double sr = 0.0, si = 0.0;
  int cnt = 0;
  for (int t = 0; t < T; t++) {
    if (mk && mk[t]) continue;
    const double2 v = x[(size_t)t * N];
    sr += v.x;
    si += v.y;
    cnt++;
  }
  const double scl = 1.0 / (double)cnt;  // inf for an empty selection: 0 * inf = NaN, as numpy's 0 / 0
  const double mr = sr * scl, mi = si * scl;
  double acc = 0.0;
  for (int t = 0; t < T; t++) {
    if (mk && mk[t]) continue;
    const double2 v = x[(size_t)t * N];
    const double dx = v.x - mr, dy = v.y - mi;
    acc += __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
  }
  double out = cnt > 0 ? acc / (double)cnt : __longlong_as_double(0x7ff8000000000000LL);
  if (mask && out != out) {  // delay.py:1099-1104
    out = 0.0;
    if (bad) bad[b] = 1;  // every thread that sees a NaN writes the same value
  }
  var[(size_t)b * N + n] = out;
}

}  // namespace dg

using namespace dg;

#define FCU(call)                                                                                             \
  do {                                                                                                        \
    cudaError_t e_ = (call);                                                                                  \
    if (e_ != cudaSuccess) {                                                                                  \
      char msg_[512];                                                                                         \
      snprintf(msg_, sizeof msg_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      dg_set_error(h, msg_);                                                                                  \
      return (int)e_;                                                                                         \
    }                                                                                                         \
  } while (0)

// delay_spectrum_fft (delay.py:2102-2130) for nrows rows at once; shift = 1 adds the fftshift of
// DelaySpectrumFFT._estimator (delay.py:980).
extern "C" int dgibbs_delay_fft(dgibbs_handle h, int32_t on_device, int64_t nrows, int32_t N, int32_t window_id,
                                int32_t shift, const void* data, void* out, double* ms_kernel) {
  if (!h) return -1;
  auto bad = [&](const char* what) {
    dg_set_error(h, (std::string("bad argument: ") + what).c_str());
    return -2;
  };
  if (nrows < 0 || N <= 0) return bad("nrows >= 0 and N > 0 are required");
  if (!window_id_valid(window_id)) return bad("window_id");
  if (ms_kernel) *ms_kernel = 0.0;
  if (nrows == 0) return 0;
  if (!data || !out) return bad("data and out are required");
  const bool pow2 = N >= 8 && (N & (N - 1)) == 0 && N <= 4096;
  if (!pow2 && (size_t)N * 32 > 200 * 1024) return bad("N must be a power of two <= 4096 or any N <= 6400");
  FCU(cudaSetDevice(dg_device(h)));
  cudaStream_t st = dg_stream(h);
  const size_t bytes = (size_t)nrows * (size_t)N * 16;
  const double2* x = (const double2*)data;
  double2* y = (double2*)out;
  void* q;
  if (!on_device) {
    FCU(dg_pool_get(h, "dfft_in", bytes, &q));
    FCU(cudaMemcpyAsync(q, data, bytes, cudaMemcpyHostToDevice, st));
    x = (const double2*)q;
    FCU(dg_pool_get(h, "dfft_out", bytes, &q));
    y = (double2*)q;
  }
  // window and twiddle tables: rebuilt only when (N, window) or the pooled buffer changed since the last call
  FCU(dg_pool_get(h, "dfft_window", sizeof(double) * (size_t)N, &q));
  double* wtab = (double*)q;
  FCU(dg_pool_get(h, "dfft_tw", sizeof(double2) * 2 * (size_t)N, &q));  // stage tables: <= 3 (8 + 64 + 512) + 3 N / 4 entries
  double2* tw = (double2*)q;
  long long *tag_w = dg_tag(h, "dfft_window"), *tag_wp = dg_tag(h, "dfft_window_ptr");
  long long *tag_t = dg_tag(h, "dfft_tw"), *tag_tp = dg_tag(h, "dfft_tw_ptr");
  const long long key_w = ((long long)N << 24) | (long long)(window_id + 2), key_t = ((long long)N << 1) | (pow2 ? 1 : 0);
  if (*tag_w != key_w || *tag_wp != (long long)(uintptr_t)wtab) {
    delay_fft_window_kernel<<<(N + 255) / 256, 256, 0, st>>>(wtab, N, window_id);
    FCU(cudaGetLastError());
    *tag_w = key_w;
    *tag_wp = (long long)(uintptr_t)wtab;
  }
  if (*tag_t != key_t || *tag_tp != (long long)(uintptr_t)tw) {
    if (pow2)
      FCU(build_stage_twiddles(tw, N, st));
    else
      delay_fft_fulltw_kernel<<<(N + 255) / 256, 256, 0, st>>>(tw, N);
    FCU(cudaGetLastError());
    *tag_t = key_t;
    *tag_tp = (long long)(uintptr_t)tw;
  }

  const size_t smem = (size_t)N * (pow2 ? 3 : 2) * sizeof(double2);
  void (*kern)(const double2*, double2*, const double*, const double2*, long long, int, int) = delay_dft_direct_kernel;
  int nthreads = FTHREADS;
  if (pow2) {
    if (N >= 4096) { kern = delay_fft_pow2_kernel<512>; nthreads = 512; }  // one CTA per SM: every butterfly its own thread
    else if (N >= 2048) { kern = delay_fft_pow2_kernel<256>; nthreads = 256; }
    else if (N == 1024) { kern = delay_fft_pow2_kernel<128>; nthreads = 128; }
    else if (N == 512) { kern = delay_fft_pow2_kernel<64>; nthreads = 64; }
    else { kern = delay_fft_pow2_kernel<32>; nthreads = 32; }
  }
  FCU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  FCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, nthreads, smem));
  if (occ < 1) return bad("the transform does not fit in shared memory");
  long long grid = (long long)occ * dg_sm_count(h);
  if (grid > nrows) grid = nrows;
  const cudaEvent_t e0 = dg_event(h, 0), e1 = dg_event(h, 1);  // the handle's own events: nothing to leak on an early return
  FCU(cudaEventRecord(e0, st));
  kern<<<(unsigned)grid, nthreads, smem, st>>>(x, y, wtab, tw, (long long)nrows, N, shift ? 1 : 0);
  FCU(cudaEventRecord(e1, st));
  FCU(cudaGetLastError());
  if (!on_device) FCU(cudaMemcpyAsync(out, y, bytes, cudaMemcpyDeviceToHost, st));
  FCU(cudaStreamSynchronize(st));
  float ms = 0.f;
  FCU(cudaEventElapsedTime(&ms, e0, e1));
  if (ms_kernel) *ms_kernel = (double)ms;
  return 0;
}

// DelaySpectrumToPowerSpectrum (delay.py:1061-1106): variance over the sample axis of a [B, T, N] complex128 delay
// spectrum; mask [B, T] (1 = masked sample, may be NULL); bad [B] receives 1 where the variance was NaN (may be NULL,
// only written with a mask, as in the reference).
extern "C" int dgibbs_sample_variance(dgibbs_handle h, int32_t on_device, int32_t B, int32_t T, int32_t N, const void* spec,
                                      const uint8_t* mask, double* var, uint8_t* bad, double* ms_kernel) {
  if (!h) return -1;
  auto badarg = [&](const char* what) {
    dg_set_error(h, (std::string("bad argument: ") + what).c_str());
    return -2;
  };
  if (B < 0 || T <= 0 || N <= 0) return badarg("B >= 0, T > 0 and N > 0 are required");
  if (B > 65535) return badarg("at most 65535 baselines per call");
  if (ms_kernel) *ms_kernel = 0.0;
  if (B == 0) return 0;
  if (!spec || !var) return badarg("spec and var are required");
  FCU(cudaSetDevice(dg_device(h)));
  cudaStream_t st = dg_stream(h);
  const size_t nb = (size_t)B * T * N * 16;
  const double2* x = (const double2*)spec;
  const unsigned char* mk = mask;
  double* v = var;
  unsigned char* bd = bad;
  void* q;
  if (!on_device) {
    FCU(dg_pool_get(h, "svar_in", nb, &q));
    FCU(cudaMemcpyAsync(q, spec, nb, cudaMemcpyHostToDevice, st));
    x = (const double2*)q;
    if (mask) {
      FCU(dg_pool_get(h, "svar_mask", (size_t)B * T, &q));
      FCU(cudaMemcpyAsync(q, mask, (size_t)B * T, cudaMemcpyHostToDevice, st));
      mk = (const unsigned char*)q;
    }
    FCU(dg_pool_get(h, "svar_out", (size_t)B * N * 8, &q));
    v = (double*)q;
    if (bad) {
      FCU(dg_pool_get(h, "svar_bad", (size_t)B, &q));
      bd = (unsigned char*)q;
    }
  }
  if (bd) FCU(cudaMemsetAsync(bd, 0, (size_t)B, st));
  const cudaEvent_t e0 = dg_event(h, 0), e1 = dg_event(h, 1);  // the handle's own events: nothing to leak on an early return
  FCU(cudaEventRecord(e0, st));
  sample_variance_kernel<<<dim3((N + 255) / 256, B), 256, 0, st>>>(x, mk, v, bd, T, N);
  FCU(cudaEventRecord(e1, st));
  FCU(cudaGetLastError());
  if (!on_device) {
    FCU(cudaMemcpyAsync(var, v, (size_t)B * N * 8, cudaMemcpyDeviceToHost, st));
    if (bad) FCU(cudaMemcpyAsync(bad, bd, (size_t)B, cudaMemcpyDeviceToHost, st));
  }
  FCU(cudaStreamSynchronize(st));
  float ms = 0.f;
  FCU(cudaEventElapsedTime(&ms, e0, e1));
  if (ms_kernel) *ms_kernel = (double)ms;
  return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// dgibbs_delay_fft_noalloc: the same transform for callers that must not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The scratch of dgibbs_delay_fft -- the window table and the per-stage twiddle tables, which that entry
// point takes from the handle's grow-only pool (cudaMalloc on first use) -- lives here in a file-scope __device__ array
// of fixed size, so the transform is limited to N = 2^m, 8 <= N <= DGIBBS_NOALLOC_MAX_N, device-resident rows.
//
// The array is part of the library image: every GPU that loads libdgibbs.so holds its 160 KB until the process exits.
// In HOST code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address; its device address is
// looked up once per device with cudaGetSymbolAddress (with that device current) and handed to the kernels.
// Every call on a device shares the one array and finds the previous call's tables in it: both tables are rebuilt in
// full for (N, window_id) at the head of every call, on the handle's stream, so nothing has to start at zero and
// nothing stale can be read -- but two calls that overlap on the GPU (different streams / handles on one device)
// race on it.  Calls on one stream are ordered and safe.
namespace dg {
constexpr int NOALLOC_MAX_N = DGIBBS_NOALLOC_MAX_N;
constexpr int NOALLOC_WINDOW_DOUBLES = NOALLOC_MAX_N;      // window table
constexpr int NOALLOC_TW_DOUBLES = 4 * NOALLOC_MAX_N;      // 2 N double2 of stage twiddles (as dgibbs_delay_fft sizes them)
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_WINDOW_DOUBLES + NOALLOC_TW_DOUBLES];
static_assert(sizeof(double) * NOALLOC_WINDOW_DOUBLES % 256 == 0, "the twiddle tables start on a 256-byte boundary");

constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};  // device address of g_noalloc_scratch, per device

// Set-up, once per device, with that device current.
static cudaError_t noalloc_scratch_address(int device, double** out) {
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    double* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_noalloc_ptr[device] = p;
  }
  *out = g_noalloc_ptr[device];
  return cudaSuccess;
}
}  // namespace dg

extern "C" int dgibbs_delay_fft_noalloc(dgibbs_handle h, int64_t nrows, int32_t N, int32_t window_id, int32_t shift,
                                        const void* data, void* out, int32_t sync) {
  if (!h) return -1;
  auto bad = [&](const char* what) {
    dg_set_error(h, (std::string("bad argument: ") + what).c_str());
    return -2;
  };
  if (nrows < 0) return bad("nrows >= 0 is required");
  if (!window_id_valid(window_id)) return bad("window_id");
  if (N < 8 || (N & (N - 1)) != 0 || N > NOALLOC_MAX_N) {
    char msg[160];
    snprintf(msg, sizeof msg, "dgibbs_delay_fft_noalloc serves N = 2^m with 8 <= N <= %d (static scratch); got N = %d: "
                              "use dgibbs_delay_fft", NOALLOC_MAX_N, (int)N);
    return bad(msg);
  }
  const int dev = dg_device(h);
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return bad("device index beyond the per-device table of the static scratch");
  FCU(cudaSetDevice(dev));
  double* scratch = nullptr;
  FCU(noalloc_scratch_address(dev, &scratch));
  if (nrows == 0) return 0;  // set-up only (e.g. before a graph capture)
  if (!data || !out) return bad("data and out are required (device pointers)");
  if (data == out) return bad("out may not alias data");
  cudaStream_t st = dg_stream(h);
  double* wtab = scratch;
  double2* tw = reinterpret_cast<double2*>(scratch + NOALLOC_WINDOW_DOUBLES);
  // both tables are rebuilt in full for this (N, window): the array keeps the last call's contents
  delay_fft_window_kernel<<<(N + 255) / 256, 256, 0, st>>>(wtab, N, window_id);
  FCU(build_stage_twiddles(tw, N, st));

  const size_t smem = (size_t)N * 3 * sizeof(double2);
  void (*kern)(const double2*, double2*, const double*, const double2*, long long, int, int);
  int nthreads;
  if (N >= 4096) { kern = delay_fft_pow2_kernel<512>; nthreads = 512; }
  else if (N >= 2048) { kern = delay_fft_pow2_kernel<256>; nthreads = 256; }
  else if (N == 1024) { kern = delay_fft_pow2_kernel<128>; nthreads = 128; }
  else if (N == 512) { kern = delay_fft_pow2_kernel<64>; nthreads = 64; }
  else { kern = delay_fft_pow2_kernel<32>; nthreads = 32; }
  FCU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  FCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, nthreads, smem));
  if (occ < 1) return bad("the transform does not fit in shared memory");
  long long grid = (long long)occ * dg_sm_count(h);
  if (grid > nrows) grid = nrows;
  kern<<<(unsigned)grid, nthreads, smem, st>>>((const double2*)data, (double2*)out, wtab, tw, (long long)nrows, N, shift ? 1 : 0);
  FCU(cudaGetLastError());
  if (sync) FCU(cudaStreamSynchronize(st));
  return 0;
}
