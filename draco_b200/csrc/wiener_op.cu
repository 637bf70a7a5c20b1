// Wiener delay-transform operator on the GPU: the per-pixel linear algebra of ConstructWienerDelayTransform.process,
// draco/analysis/powerspec.py:229-319 (C ABI in include/dgwiener.h; linked into libdgnrml.so with cuBLAS / cuSOLVER).
//
// The reference loops over elevation and right ascension and, per pixel, forms H G H^H + N (two F^3 complex GEMMs),
// inverts it on the valid channels through cho_factor / cho_solve against the identity (F right-hand sides), forms
// R^T = F^H H^H and multiplies it into the inverse.  Here a chunk of pixels is resident in HBM and
//   * every host array is used in the memory layout it has: a row-major [F, F] matrix IS its transpose in cuBLAS's
//     column-major convention, so no transposition kernel runs; with P = H^T, G' = G^T the transposed covariance is
//     A' = P^H G' P + N' and the transposed operator D' = A'^-1 (P^H F'^T), i.e. D' is written in the layout the
//     output container wants ([delay, freq] row-major);
//   * G' P is ONE GEMM for the whole chunk ([F, F] x [F, R F]); P^H (G' P) and P^H F' are strided-batched GEMMs;
//   * the masked channels are decoupled by a unit diagonal (their rows of H and of N are exactly zero), so the
//     "valid" sub-matrix never has to be gathered: the factor of the padded matrix is the factor of the sub-matrix;
//   * the solve runs on the ndelay = F/2 columns of R^T instead of the F columns of an explicit inverse
//     (25 F^3 instead of 33 F^3 flop per pixel);
//   * mask / noise scaling / window / prior / complex64 rounding are the hand-written kernels below.
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <cusolverDn.h>

#include <algorithm>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/dgwiener.h"
#include "wiener_own.h"

namespace {

typedef cuDoubleComplex zc;
constexpr int NSTREAM = 4;

thread_local std::string g_err;

struct WLibs {
  bool ready = false;
  cublasHandle_t blas = nullptr;
  cusolverDnHandle_t solver = nullptr;  // main stream: batched factorisation
  cusolverDnHandle_t ssolver[NSTREAM] = {};
  cublasHandle_t sblas[NSTREAM] = {};
  cudaStream_t stream = nullptr, sstream[NSTREAM] = {};
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, evA = nullptr, evB = nullptr, evS[NSTREAM] = {};
};
WLibs g_wlibs[64];

#define WCU(call)                                                                  \
  do {                                                                             \
    cudaError_t e_ = (call);                                                       \
    if (e_ != cudaSuccess) {                                                       \
      g_err = std::string(#call) + " failed: " + cudaGetErrorString(e_);           \
      return 100 + (int)e_;                                                        \
    }                                                                              \
  } while (0)
#define WBL(call)                                                                  \
  do {                                                                             \
    int e_ = (int)(call);                                                          \
    if (e_ != 0) {                                                                 \
      g_err = std::string(#call) + " failed with library status " + std::to_string(e_); \
      return 1000 + e_;                                                            \
    }                                                                              \
  } while (0)

// ------------------------------------------------------------------------------------------------ kernels
// F' (column-major [F, ndelay]) = conj(fourier) transposed in memory: F'(k, d) = conj(fourier[k, d])
// FS(k, d) = F'(k, d) prior_d
__global__ void fourier_prep_kernel(const zc* __restrict__ fourier, const double* __restrict__ prior, zc* __restrict__ Fp,
                                    zc* __restrict__ FS, int F, int nd) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x, d = blockIdx.y;
  if (k >= F) return;
  const zc v = fourier[(size_t)k * nd + d];
  Fp[(size_t)d * F + k] = make_cuDoubleComplex(v.x, -v.y);
  FS[(size_t)d * F + k] = make_cuDoubleComplex(v.x * prior[d], -v.y * prior[d]);
}

// FSFT'(i, j) = sum_d FS(i, d) conj(F'(j, d))  (powerspec.py:234; column-major [F, F], once per call): 16 x 16 output
// tile per block, the two operand slabs of 16 delays staged through shared memory (own-kernel path; the library paths
// call ZGEMM)
__global__ void fsft_kernel(const zc* __restrict__ FS, const zc* __restrict__ Fp, zc* __restrict__ FSFT, int F, int nd) {
  __shared__ double2 sa[16][17], sb[16][17];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int i = blockIdx.x * 16 + tx, j = blockIdx.y * 16 + ty;
  double re = 0.0, im = 0.0;
  for (int d0 = 0; d0 < nd; d0 += 16) {
    const int d = d0 + ty, ia = blockIdx.x * 16 + tx, jb = blockIdx.y * 16 + tx;
    sa[ty][tx] = (d < nd && ia < F) ? FS[(size_t)d * F + ia] : make_double2(0.0, 0.0);
    sb[ty][tx] = (d < nd && jb < F) ? Fp[(size_t)d * F + jb] : make_double2(0.0, 0.0);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; k++) {
      const double2 a = sa[k][tx], b = sb[k][ty];  // a conj(b)
      re += a.x * b.x + a.y * b.y;
      im += a.y * b.x - a.x * b.y;
    }
    __syncthreads();
  }
  if (i < F && j < F) FSFT[(size_t)j * F + i] = make_cuDoubleComplex(re, im);
}

// G'(i, j) = FSFT'(i, j) b_i b_j, b = sqrt(beam power)  (powerspec.py:266, 288)
__global__ void beam_scale_kernel(const zc* __restrict__ FSFT, const double* __restrict__ bp, zc* __restrict__ G, int F) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
  if (i >= F) return;
  const double s = sqrt(bp[i]) * sqrt(bp[j]);
  const zc v = FSFT[(size_t)j * F + i];
  G[(size_t)j * F + i] = make_cuDoubleComplex(v.x * s, v.y * s);
}

__device__ __forceinline__ zc load_zc(const void* p, size_t i, int is_complex) {
  if (is_complex) return ((const zc*)p)[i];
  return make_cuDoubleComplex(((const double*)p)[i], 0.0);
}

// Per pixel and channel: M = (window > 0) && (w > 0);  r = sqrt(invert_no_zero(w C_ff)) (window > 0)
// (powerspec.py:273, 281-282).  A real covariance keeps numpy's real arithmetic (negative -> NaN), a complex one
// the principal complex square root.
__global__ void mask_noise_kernel(const double* __restrict__ w, const void* __restrict__ cov, int cov_cplx,
                                  const double* __restrict__ window, unsigned char* __restrict__ M, zc* __restrict__ rn,
                                  int R, int F) {
  const size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= (size_t)R * F) return;
  const int f = (int)(e % F);
  const size_t r = e / F;
  const bool win = window[f] > 0.0;
  const double wt = w[e];
  M[e] = (win && wt > 0.0) ? 1 : 0;
  const zc c = load_zc(cov, r * (size_t)F * F + (size_t)f * F + f, cov_cplx);
  zc out = make_cuDoubleComplex(0.0, 0.0);
  const double xr = wt * c.x, xi = wt * c.y;
  if (win && (xr != 0.0 || xi != 0.0)) {
    if (!cov_cplx) {
      out.x = sqrt(1.0 / xr);
    } else {
      const double den = xr * xr + xi * xi;
      const double ir = xr / den, ii = -xi / den;  // 1 / x
      const double mod = hypot(ir, ii);
      double sr = sqrt(0.5 * (mod + ir)), si = sqrt(0.5 * (mod - ir));  // principal square root
      if (!(sr == sr)) sr = 0.0;
      if (!(si == si)) si = 0.0;
      out.x = sr;
      out.y = ii < 0.0 ? -si : si;
    }
  }
  rn[e] = out;
}

// P = memory of H = diag(M) K (row f of K kept or zeroed), as complex128  (powerspec.py:281-283)
__global__ void mask_filter_kernel(const void* __restrict__ K, int k_cplx, const unsigned char* __restrict__ M,
                                   zc* __restrict__ P, size_t n, int F) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    const size_t row = e / F;  // = pixel * F + f
    P[e] = M[row] ? load_zc(K, e, k_cplx) : make_cuDoubleComplex(0.0, 0.0);
  }
}

// A += C o r r^T (powerspec.py:274-277, 292), unit diagonal on the masked channels
__global__ void add_noise_kernel(zc* __restrict__ A, const void* __restrict__ cov, int cov_cplx, const zc* __restrict__ rn,
                                 const unsigned char* __restrict__ M, size_t n, int F) {
  const size_t FF = (size_t)F * F;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    const size_t r = e / FF, o = e - r * FF;
    const int a = (int)(o / F), b = (int)(o - (size_t)a * F);
    const zc ra = rn[r * F + a], rb = rn[r * F + b];
    const zc rr = make_cuDoubleComplex(ra.x * rb.x - ra.y * rb.y, ra.x * rb.y + ra.y * rb.x);
    const zc c = load_zc(cov, e, cov_cplx);
    zc v = A[e];
    v.x += c.x * rr.x - c.y * rr.y;
    v.y += c.x * rr.y + c.y * rr.x;
    if (a == b && !M[r * F + a]) v = make_cuDoubleComplex(1.0, 0.0);
    A[e] = v;
  }
}

// out[d, f] = complex64( prior_d Z[d, f] window_f )  (powerspec.py:314-316); masked channels are exact zeros
__global__ void epilogue_kernel(const zc* __restrict__ Z, const double* __restrict__ prior, const double* __restrict__ window,
                                const unsigned char* __restrict__ M, float2* __restrict__ out, size_t n, int F, int nd) {
  const size_t per = (size_t)F * nd;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    const size_t r = e / per, o = e - r * per;
    const int d = (int)(o / F), f = (int)(o - (size_t)d * F);
    const double s = M[r * F + f] ? prior[d] * window[f] : 0.0;
    const zc z = Z[e];
    out[e] = make_float2((float)(s * z.x), (float)(s * z.y));
  }
}

__global__ void fill_ptrs_kernel(zc** Aptr, zc** Yptr, zc* A, zc* Y, int R, size_t sa, size_t sy) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  Aptr[r] = A + (size_t)r * sa;
  Yptr[r] = Y + (size_t)r * sy;
}

// ApplyWienerDelayTransform, powerspec.py:430-452: one CTA per pixel (ra, el).  The map and the inverse weights of the
// pixel are gathered into shared memory once (they are strided by nra * nel in the container's [freq, ra, el] layout);
// every warp then streams rows of the complex64 operator (16 bytes per lane and load, four loads in flight) and keeps
// the three sums in fp64.  |op|^2 is formed in fp32 as numpy does for a complex64 operand (np.abs(op) ** 2).
template <int NT>
__global__ void __launch_bounds__(NT) apply_kernel(const float2* __restrict__ op, long long stride_ra, long long stride_el,
                                                   int ra0, const double* __restrict__ map, const double* __restrict__ wgt,
                                                   double2* __restrict__ spec, float* __restrict__ sw, int F, int nd, int nra,
                                                   int nel) {
  extern __shared__ double sm[];
  double* m = sm;
  double* v = sm + F;
  const int ee = blockIdx.x % nel, rl = blockIdx.x / nel, rr = ra0 + rl;
  for (int f = threadIdx.x; f < F; f += NT) {
    const size_t o = ((size_t)f * nra + rr) * nel + ee;
    m[f] = map[o];
    const double w = wgt[o];
    v[f] = w != 0.0 ? 1.0 / w : 0.0;
  }
  __syncthreads();
  const float2* base = op + (size_t)rl * stride_ra + (size_t)ee * stride_el;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool vec = (F % 2 == 0) && ((((size_t)base) & 15) == 0);
  for (int d = warp; d < nd; d += NT / 32) {
    const float2* row = base + (size_t)d * F;
    double sr = 0.0, si = 0.0, sv = 0.0;
    if (vec) {
      const float4* row4 = reinterpret_cast<const float4*>(row);
      const int n4 = F / 2;
#pragma unroll 4
      for (int q = lane; q < n4; q += 32) {
        const float4 c = __ldg(row4 + q);
        const double m0 = m[2 * q], m1 = m[2 * q + 1];
        const float a0 = hypotf(c.x, c.y), a1 = hypotf(c.z, c.w);
        sr += (double)c.x * m0 + (double)c.z * m1;
        si += (double)c.y * m0 + (double)c.w * m1;
        sv += (double)(a0 * a0) * v[2 * q] + (double)(a1 * a1) * v[2 * q + 1];
      }
    } else {
      for (int f = lane; f < F; f += 32) {
        const float2 c = row[f];
        const float a = hypotf(c.x, c.y);
        sr += (double)c.x * m[f];
        si += (double)c.y * m[f];
        sv += (double)(a * a) * v[f];
      }
    }
    for (int o = 16; o > 0; o >>= 1) {
      sr += __shfl_xor_sync(0xffffffffu, sr, o);
      si += __shfl_xor_sync(0xffffffffu, si, o);
      sv += __shfl_xor_sync(0xffffffffu, sv, o);
    }
    if (lane == 0) {
      const size_t o = ((size_t)ee * nra + rr) * nd + d;
      spec[o] = make_double2(sr, si);
      sw[o] = sv != 0.0 ? (float)(1.0 / sv) : 0.f;
    }
  }
}

int init_libs(int dev) {
  WLibs& L = g_wlibs[dev];
  if (L.ready) return 0;
  WCU(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking));
  WBL(cublasCreate(&L.blas));
  WBL(cublasSetStream(L.blas, L.stream));
  WBL(cusolverDnCreate(&L.solver));
  WBL(cusolverDnSetStream(L.solver, L.stream));
  for (cudaEvent_t* e : {&L.ev0, &L.ev1, &L.evA, &L.evB}) WCU(cudaEventCreate(e));
  for (int i = 0; i < NSTREAM; i++) {
    WCU(cudaStreamCreateWithFlags(&L.sstream[i], cudaStreamNonBlocking));
    WBL(cusolverDnCreate(&L.ssolver[i]));
    WBL(cusolverDnSetStream(L.ssolver[i], L.sstream[i]));
    WBL(cublasCreate(&L.sblas[i]));
    WBL(cublasSetStream(L.sblas[i], L.sstream[i]));
    WCU(cudaEventCreateWithFlags(&L.evS[i], cudaEventDisableTiming));
  }
  L.ready = true;
  return 0;
}

struct DevBufs {
  std::vector<void*> ptrs;
  ~DevBufs() {
    for (void* p : ptrs) cudaFree(p);
  }
  template <class T>
  cudaError_t alloc(T** p, size_t bytes) {
    cudaError_t e = cudaMalloc((void**)p, bytes ? bytes : 16);
    if (e == cudaSuccess) ptrs.push_back((void*)*p);
    return e;
  }
};

}  // namespace

extern "C" int dgwiener_sizeof(void) { return (int)sizeof(dgwiener_problem); }
extern "C" int dgwiener_apply_sizeof(void) { return (int)sizeof(dgwiener_apply_problem); }
extern "C" const char* dgwiener_last_error(void) { return g_err.c_str(); }

extern "C" int dgwiener_construct(dgwiener_problem* p) {
  g_err.clear();
  if (!p) return -1;
  const int F = p->nfreq, nd = p->ndelay, nra = p->nra, nel = p->nel;
  if (p->device_id < 0 || p->device_id >= 64) return -2;
  if (F <= 0 || nd <= 0 || nra < 0 || nel < 0) return -3;
  if (!p->fourier || !p->prior || !p->window) return -4;
  if (nra > 0 && nel > 0 && (!p->filter || !p->freq_cov || !p->weight || !p->beam_power || !p->out)) return -5;
  p->bad_el = p->bad_ra = -1;
  p->ms_device = p->ms_compute = p->flop = 0.0;
  p->launches = 0;
  {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || p->device_id >= count) {
      g_err = "no CUDA device " + std::to_string(p->device_id) + " (there is no CPU fallback)";
      cudaGetLastError();
      return 100;
    }
  }
  if (nra == 0 || nel == 0) return 0;
  WCU(cudaSetDevice(p->device_id));
  if (int rc = init_libs(p->device_id)) return rc;
  WLibs& L = g_wlibs[p->device_id];
  cudaStream_t st = L.stream;

  const size_t FF = (size_t)F * F, FD = (size_t)F * nd;
  const size_t kbytes = p->filter_is_complex ? 16 : 8, cbytes = p->cov_is_complex ? 16 : 8;
  // solver 3 (and 0 = auto): every contraction, the factorisation and the solves on this repository's own DMMA kernels
  // (wiener_own.cu); 1 / 2: cuBLAS ZGEMM + cuSOLVER ZPOTRF (batched / per matrix) + cuBLAS ZTRSM
  // 4: as 3 with the three contractions on the INT8 tensor cores (Ozaki split, ozaki_gemm.cuh: tcgen05.mma kind::i8 + TMEM)
  const bool own = p->solver == 3 || p->solver == 0 || p->solver == 4;
  int sm_count = 148;
  WCU(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, p->device_id));
  const dgw::OwnGeom og = dgw::own_geom(F, nd, sm_count);
  const dgw::OwnSizes oz = dgw::own_sizes(og);
  const size_t pp_own = FF * (kbytes + cbytes) + 8 * (oz.Pr + oz.Pext + oz.T1r + oz.Ar + oz.Yr) + FD * 8 + (size_t)F * (8 + 16 + 1) + 64;
  const size_t pp_int8 = pp_own + oz.qP + oz.qPe + 4 * (oz.eP + oz.ePe);
  // default budget of a chunk: 32 GiB of the 180 GB, at most half of what is free (a full wave of F = 1024 pixels needs 15 / 23 GB)
  size_t budget = (size_t)32 << 30;
  if (p->workspace_bytes > 0) {
    budget = (size_t)p->workspace_bytes;
  } else {
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) budget = std::max<size_t>((size_t)1 << 30, std::min(budget, free_b / 2));
  }
  // 0 (auto): the int8 contractions from F = 256 (measured faster there: profiles/r04i_perf_wiener_op.json.log) provided the
  // budget still holds a full wave of pixels with their slices; the DMMA contractions otherwise
  const size_t wave = std::min<size_t>((size_t)nra, (size_t)og.nslots);
  const int ozaki = (p->solver == 4 || (p->solver == 0 && F >= 256 && budget / pp_int8 >= wave)) ? 1 : 0;
  const size_t per_pixel = own ? (ozaki ? pp_int8 : pp_own)
                               : FF * (kbytes + cbytes + 3 * 16) + FD * (16 + 8) + (size_t)F * (8 + 16 + 1) + 64;
  int R = (int)std::max<size_t>(1, std::min<size_t>(std::min<size_t>((size_t)nra, 32768), budget / per_pixel));
  // the factorisation runs one pixel per SM at a time: a chunk that is not the last one should fill whole waves
  // (r04g: 296 pixels in chunks of 209 + 87 ran the solve kernel for 3 waves instead of 2)
  if (own && R < nra && R > og.nslots) R -= R % og.nslots;
  // measured (profiles/r01j_perf_wiener_op.json.log): the batched factorisation wins up to F = 512 at least
  const bool batched = p->solver == 1;

  DevBufs bufs;
  zc *Fp, *FS, *FSFT, *G, *P, *T1, *A, *Y, *rn;
  double *prior, *window, *bp, *w;
  void *K, *C;
  unsigned char* M;
  float2* outd;
  int* info;
  zc **Aptr, **Yptr;
  WCU(bufs.alloc(&Fp, FD * 16)); WCU(bufs.alloc(&FS, FD * 16)); WCU(bufs.alloc(&FSFT, FF * 16)); WCU(bufs.alloc(&G, FF * 16));
  WCU(bufs.alloc(&prior, (size_t)nd * 8)); WCU(bufs.alloc(&window, (size_t)F * 8)); WCU(bufs.alloc(&bp, (size_t)nel * F * 8));
  WCU(bufs.alloc(&K, (size_t)R * FF * kbytes)); WCU(bufs.alloc(&C, (size_t)R * FF * cbytes));
  dgw::OwnBufs ob{};
  if (own) {
    // T1 only carries the host Fourier operator to the device; P / A / Y are replaced by the interleaved-real buffers
    WCU(bufs.alloc(&T1, std::max(FD, FF) * 16));
    P = A = Y = nullptr;
    struct { double** p; size_t n; } own_list[] = {{&ob.Pr, R * oz.Pr}, {&ob.Pext, R * oz.Pext}, {&ob.T1r, R * oz.T1r}, {&ob.Ar, R * oz.Ar},
                                                   {&ob.Yr, R * oz.Yr}, {&ob.Gext, oz.Gext}, {&ob.Fpr, oz.Fpr}, {&ob.work, oz.work}, {&ob.Wp, oz.Wp}};
    for (auto& it : own_list) {
      WCU(bufs.alloc(it.p, it.n * 8));
      WCU(cudaMemsetAsync(*it.p, 0, it.n * 8, L.stream));  // the padding of every matrix stays zero from here on
    }
    if (ozaki) {
      WCU(bufs.alloc(&ob.qP, R * oz.qP)); WCU(bufs.alloc(&ob.qPe, R * oz.qPe)); WCU(bufs.alloc(&ob.qG, oz.qG)); WCU(bufs.alloc(&ob.qF, oz.qF));
      WCU(bufs.alloc(&ob.eP, R * oz.eP * 4)); WCU(bufs.alloc(&ob.ePe, R * oz.ePe * 4)); WCU(bufs.alloc(&ob.eG, oz.eG * 4));
      WCU(bufs.alloc(&ob.eF, oz.eF * 4));
    }
    WCU(dgw::own_configure());
  } else {
    WCU(bufs.alloc(&P, (size_t)R * FF * 16)); WCU(bufs.alloc(&T1, (size_t)R * FF * 16)); WCU(bufs.alloc(&A, (size_t)R * FF * 16));
    WCU(bufs.alloc(&Y, (size_t)R * FD * 16));
  }
  WCU(bufs.alloc(&outd, (size_t)R * FD * 8));
  WCU(bufs.alloc(&w, (size_t)R * F * 8)); WCU(bufs.alloc(&rn, (size_t)R * F * 16)); WCU(bufs.alloc(&M, (size_t)R * F));
  WCU(bufs.alloc(&info, (size_t)R * sizeof(int)));
  WCU(bufs.alloc(&Aptr, (size_t)R * sizeof(zc*))); WCU(bufs.alloc(&Yptr, (size_t)R * sizeof(zc*)));
  // per-stream cuSOLVER workspaces (per-matrix mode)
  zc* swork[NSTREAM] = {};
  int lwork = 0;
  if (!batched && !own) {
    WBL(cusolverDnZpotrf_bufferSize(L.ssolver[0], CUBLAS_FILL_MODE_LOWER, F, A, F, &lwork));
    for (int i = 0; i < NSTREAM; i++) WCU(bufs.alloc(&swork[i], (size_t)std::max(lwork, 1) * 16));
  }

  const zc one = make_cuDoubleComplex(1.0, 0.0), zero = make_cuDoubleComplex(0.0, 0.0);
  std::vector<int> hinfo((size_t)R);
  int64_t launches = 0;
  double ms_compute = 0.0, flop = 0.0;

  WCU(cudaEventRecord(L.ev0, st));
  WCU(cudaMemcpyAsync(prior, p->prior, (size_t)nd * 8, cudaMemcpyHostToDevice, st));
  WCU(cudaMemcpyAsync(window, p->window, (size_t)F * 8, cudaMemcpyHostToDevice, st));
  WCU(cudaMemcpyAsync(bp, p->beam_power, (size_t)nel * F * 8, cudaMemcpyHostToDevice, st));
  // the host Fourier operator passes through T1 (free at this point) on its way to F' / F' S
  WCU(cudaMemcpyAsync(T1, p->fourier, FD * 16, cudaMemcpyHostToDevice, st));
  fourier_prep_kernel<<<dim3((F + 127) / 128, nd), 128, 0, st>>>(T1, prior, Fp, FS, F, nd);
  // FSFT' = (F' S) F'^H  (powerspec.py:234)
  if (own) {
    fsft_kernel<<<dim3((F + 15) / 16, (F + 15) / 16), dim3(16, 16), 0, st>>>(FS, Fp, FSFT, F, nd);
    WCU(cudaGetLastError());
  } else {
    WBL(cublasZgemm(L.blas, CUBLAS_OP_N, CUBLAS_OP_C, F, F, nd, &one, FS, F, Fp, F, &zero, FSFT, F));
  }
  launches += 2;
  flop += 8.0 * F * (double)F * nd;
  if (own) {
    WCU(dgw::own_prepare_fourier(st, og, Fp, ob, ozaki));
    launches += 1 + ozaki;
  }

  for (int r0 = 0; r0 < nra; r0 += R) {
    const int Rc = std::min(R, nra - r0);
    const size_t nFF = (size_t)Rc * FF;
    WCU(cudaMemcpyAsync(K, (const char*)p->filter + (size_t)r0 * FF * kbytes, nFF * kbytes, cudaMemcpyHostToDevice, st));
    WCU(cudaMemcpyAsync(C, (const char*)p->freq_cov + (size_t)r0 * FF * cbytes, nFF * cbytes, cudaMemcpyHostToDevice, st));
    for (int ee = 0; ee < nel; ee++) {
      WCU(cudaMemcpyAsync(w, p->weight + ((size_t)ee * nra + r0) * F, (size_t)Rc * F * 8, cudaMemcpyHostToDevice, st));
      WCU(cudaEventRecord(L.evA, st));
      const int gridE = (int)std::min<size_t>((nFF + 255) / 256, 148 * 16);
      if (own) {
        mask_noise_kernel<<<(unsigned)(((size_t)Rc * F + 255) / 256), 256, 0, st>>>(w, C, p->cov_is_complex, window, M, rn, Rc, F);
        launches += 1;
        WCU(cudaMemsetAsync(info, 0, (size_t)Rc * sizeof(int), st));
        WCU(dgw::own_chunk(st, og, ob, Rc, sm_count, FSFT, bp + (size_t)ee * F, K, p->filter_is_complex, C, p->cov_is_complex, M, rn,
                           prior, window, outd, info, &launches, ozaki));
      } else {
      beam_scale_kernel<<<dim3((F + 127) / 128, F), 128, 0, st>>>(FSFT, bp + (size_t)ee * F, G, F);
      mask_noise_kernel<<<(unsigned)(((size_t)Rc * F + 255) / 256), 256, 0, st>>>(w, C, p->cov_is_complex, window, M, rn, Rc, F);
      mask_filter_kernel<<<gridE, 256, 0, st>>>(K, p->filter_is_complex, M, P, nFF, F);
      // T1 = G' [P_1 ... P_R]: one GEMM for the chunk  (the H (FSFT o b b^T) half of powerspec.py:288)
      WBL(cublasZgemm(L.blas, CUBLAS_OP_N, CUBLAS_OP_N, F, Rc * F, F, &one, G, F, P, F, &zero, T1, F));
      // A'_r = P_r^H T1_r  (the ... H^H half)
      WBL(cublasZgemmStridedBatched(L.blas, CUBLAS_OP_C, CUBLAS_OP_N, F, F, F, &one, P, F, (long long)FF, T1, F, (long long)FF,
                                    &zero, A, F, (long long)FF, Rc));
      add_noise_kernel<<<gridE, 256, 0, st>>>(A, C, p->cov_is_complex, rn, M, nFF, F);
      // Y_r = P_r^H F' = memory of R^T = F^H H^H  (powerspec.py:311)
      WBL(cublasZgemmStridedBatched(L.blas, CUBLAS_OP_C, CUBLAS_OP_N, F, nd, F, &one, P, F, (long long)FF, Fp, F, 0LL, &zero, Y,
                                    F, (long long)FD, Rc));
      launches += 7;
      // Cholesky factorisation + solve on the ndelay columns of R^T  (powerspec.py:296-304, 314-316)
      if (batched) {
        fill_ptrs_kernel<<<(Rc + 127) / 128, 128, 0, st>>>(Aptr, Yptr, A, Y, Rc, FF, FD);
        WBL(cusolverDnZpotrfBatched(L.solver, CUBLAS_FILL_MODE_LOWER, F, Aptr, F, info, Rc));
        WBL(cublasZtrsmBatched(L.blas, CUBLAS_SIDE_LEFT, CUBLAS_FILL_MODE_LOWER, CUBLAS_OP_N, CUBLAS_DIAG_NON_UNIT, F, nd, &one,
                               (const zc* const*)Aptr, F, Yptr, F, Rc));
        WBL(cublasZtrsmBatched(L.blas, CUBLAS_SIDE_LEFT, CUBLAS_FILL_MODE_LOWER, CUBLAS_OP_C, CUBLAS_DIAG_NON_UNIT, F, nd, &one,
                               (const zc* const*)Aptr, F, Yptr, F, Rc));
        launches += 4;
      } else {
        WCU(cudaEventRecord(L.evB, st));
        for (int i = 0; i < NSTREAM; i++) WCU(cudaStreamWaitEvent(L.sstream[i], L.evB, 0));
        for (int r = 0; r < Rc; r++) {
          const int i = r % NSTREAM;
          WBL(cusolverDnZpotrf(L.ssolver[i], CUBLAS_FILL_MODE_LOWER, F, A + (size_t)r * FF, F, swork[i], lwork, info + r));
          WBL(cublasZtrsm(L.sblas[i], CUBLAS_SIDE_LEFT, CUBLAS_FILL_MODE_LOWER, CUBLAS_OP_N, CUBLAS_DIAG_NON_UNIT, F, nd, &one,
                          A + (size_t)r * FF, F, Y + (size_t)r * FD, F));
          WBL(cublasZtrsm(L.sblas[i], CUBLAS_SIDE_LEFT, CUBLAS_FILL_MODE_LOWER, CUBLAS_OP_C, CUBLAS_DIAG_NON_UNIT, F, nd, &one,
                          A + (size_t)r * FF, F, Y + (size_t)r * FD, F));
        }
        for (int i = 0; i < NSTREAM; i++) {
          WCU(cudaEventRecord(L.evS[i], L.sstream[i]));
          WCU(cudaStreamWaitEvent(st, L.evS[i], 0));
        }
        launches += 3 * (int64_t)Rc;
      }
      const size_t nFD = (size_t)Rc * FD;
      epilogue_kernel<<<(int)std::min<size_t>((nFD + 255) / 256, 148 * 16), 256, 0, st>>>(Y, prior, window, M, outd, nFD, F, nd);
      launches += 1;
      }  // library path
      WCU(cudaEventRecord(L.ev1, st));
      WCU(cudaMemcpyAsync(hinfo.data(), info, (size_t)Rc * sizeof(int), cudaMemcpyDeviceToHost, st));
      char* dst = (char*)p->out + 8 * ((size_t)r0 * p->out_stride_ra + (size_t)ee * p->out_stride_el);
      WCU(cudaMemcpy2DAsync(dst, (size_t)p->out_stride_ra * 8, outd, FD * 8, FD * 8, (size_t)Rc,
                            p->out_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
      WCU(cudaGetLastError());
      WCU(cudaStreamSynchronize(st));
      float ms = 0.f;
      cudaEventElapsedTime(&ms, L.evA, L.ev1);
      ms_compute += ms;
      flop += (double)Rc * (16.0 * F * (double)F * F + 4.0 / 3.0 * F * (double)F * F + 16.0 * F * (double)F * nd);
      for (int r = 0; r < Rc; r++)
        if (hinfo[r] != 0) {
          p->bad_el = ee;
          p->bad_ra = r0 + r;
          g_err = "covariance of pixel (ra " + std::to_string(r0 + r) + ", el " + std::to_string(ee) +
                  ") is not positive definite" + (own ? std::string() : " (leading minor " + std::to_string(hinfo[r]) + ")");
          return 1;
        }
    }
  }
  WCU(cudaEventRecord(L.ev1, st));
  WCU(cudaStreamSynchronize(st));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, L.ev0, L.ev1);
  p->ms_device = ms;
  p->ms_compute = ms_compute;
  p->flop = flop;
  p->launches = launches;
  return 0;
}

extern "C" int dgwiener_gemm_ozaki(int32_t device_id, const double* A, const double* B, double* C, int32_t M, int32_t N, int32_t K,
                                   int32_t batch, double* ms_device) {
  g_err.clear();
  if (device_id < 0 || device_id >= 64) return -1;
  if (M < 0 || N < 0 || K < 0 || batch < 0 || M % 128 || N % 64 || K % 128) {
    g_err = "dgwiener_gemm_ozaki: M % 128, N % 64 and K % 128 must be 0";
    return -5;
  }
  if (ms_device) *ms_device = 0.0;
  if ((size_t)M * N * batch == 0) return 0;
  if (!A || !B || !C) return -2;
  {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device_id >= count) {
      g_err = "no CUDA device " + std::to_string(device_id) + " (there is no CPU fallback)";
      cudaGetLastError();
      return 100;
    }
  }
  WCU(cudaSetDevice(device_id));
  if (int rc = init_libs(device_id)) return rc;
  WLibs& L = g_wlibs[device_id];
  cudaStream_t st = L.stream;
  int sm_count = 148;
  WCU(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device_id));
  const dgw::OzakiScratch z = dgw::own_gemm_ozaki_sizes(M, N, K, batch);
  DevBufs bufs;
  double *dA, *dB, *dC;
  int8_t *qa, *qb;
  int *ea, *eb;
  const size_t nA = (size_t)batch * M * K, nB = (size_t)batch * N * K, nC = (size_t)batch * M * N;
  WCU(bufs.alloc(&dA, nA * 8)); WCU(bufs.alloc(&dB, nB * 8)); WCU(bufs.alloc(&dC, nC * 8));
  WCU(bufs.alloc(&qa, z.qa)); WCU(bufs.alloc(&qb, z.qb)); WCU(bufs.alloc(&ea, z.ea * 4)); WCU(bufs.alloc(&eb, z.eb * 4));
  if (K > 0) {
    WCU(cudaMemcpyAsync(dA, A, nA * 8, cudaMemcpyHostToDevice, st));
    WCU(cudaMemcpyAsync(dB, B, nB * 8, cudaMemcpyHostToDevice, st));
  }
  WCU(cudaEventRecord(L.evA, st));
  if (K > 0) {
    WCU(dgw::own_gemm_ozaki(st, sm_count, dA, dB, dC, M, N, K, batch, qa, qb, ea, eb));
  } else {
    WCU(cudaMemsetAsync(dC, 0, nC * 8, st));
  }
  WCU(cudaEventRecord(L.ev1, st));
  WCU(cudaMemcpyAsync(C, dC, nC * 8, cudaMemcpyDeviceToHost, st));
  WCU(cudaStreamSynchronize(st));
  if (ms_device) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, L.evA, L.ev1);
    *ms_device = ms;
  }
  return 0;
}

extern "C" int dgwiener_apply(dgwiener_apply_problem* p) {
  g_err.clear();
  if (!p) return -1;
  const int F = p->nfreq, nd = p->ndelay, nra = p->nra, nel = p->nel;
  if (p->device_id < 0 || p->device_id >= 64) return -2;
  if (F <= 0 || nd <= 0 || nra < 0 || nel < 0) return -3;
  p->ms_kernel = p->bytes = 0.0;
  p->launches = 0;
  {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || p->device_id >= count) {
      g_err = "no CUDA device " + std::to_string(p->device_id) + " (there is no CPU fallback)";
      cudaGetLastError();
      return 100;
    }
  }
  if (nra == 0 || nel == 0) return 0;
  if (!p->op || !p->map || !p->weight || !p->spectrum || !p->spectrum_weight) return -4;
  if ((size_t)F * 16 > 200 * 1024) {
    g_err = "nfreq too large for the shared-memory staging of the apply kernel";
    return -3;
  }
  WCU(cudaSetDevice(p->device_id));
  if (int rc = init_libs(p->device_id)) return rc;
  WLibs& L = g_wlibs[p->device_id];
  cudaStream_t st = L.stream;
  constexpr int NT = 256;
  const size_t smem = (size_t)F * 16;
  WCU(cudaFuncSetAttribute(apply_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));

  const size_t nmap = (size_t)F * nra * nel, nout = (size_t)nel * nra * nd, FD = (size_t)F * nd;
  DevBufs bufs;
  double *map, *wgt;
  double2* spec;
  float* sw;
  float2* opd = nullptr;
  WCU(bufs.alloc(&map, nmap * 8)); WCU(bufs.alloc(&wgt, nmap * 8));
  WCU(bufs.alloc(&spec, nout * 16)); WCU(bufs.alloc(&sw, nout * 4));
  WCU(cudaMemcpyAsync(map, p->map, nmap * 8, cudaMemcpyHostToDevice, st));
  WCU(cudaMemcpyAsync(wgt, p->weight, nmap * 8, cudaMemcpyHostToDevice, st));
  int R = nra;
  if (!p->op_on_device) {
    const size_t budget = p->workspace_bytes > 0 ? (size_t)p->workspace_bytes : ((size_t)2 << 30);
    R = (int)std::max<size_t>(1, std::min<size_t>((size_t)nra, budget / ((size_t)nel * FD * 8)));
    WCU(bufs.alloc(&opd, (size_t)R * nel * FD * 8));
  }
  double ms_kernel = 0.0;
  for (int r0 = 0; r0 < nra; r0 += R) {
    const int Rc = std::min(R, nra - r0);
    const float2* opc;
    long long sra, sel;
    if (p->op_on_device) {
      opc = (const float2*)p->op;
      sra = p->op_stride_ra;
      sel = p->op_stride_el;
    } else {  // pack the chunk's [ndelay, F] blocks as [Rc, nel, ndelay, F]
      const char* src = (const char*)p->op + 8 * (size_t)r0 * p->op_stride_ra;
      if (p->op_stride_el == (long long)FD && p->op_stride_ra == (long long)(nel * FD)) {
        WCU(cudaMemcpyAsync(opd, src, (size_t)Rc * nel * FD * 8, cudaMemcpyHostToDevice, st));
      } else {
        for (int r = 0; r < Rc; r++)
          WCU(cudaMemcpy2DAsync(opd + (size_t)r * nel * FD, FD * 8, src + 8 * (size_t)r * p->op_stride_ra,
                                (size_t)p->op_stride_el * 8, FD * 8, (size_t)nel, cudaMemcpyHostToDevice, st));
      }
      opc = opd;
      sra = (long long)(nel * FD);
      sel = (long long)FD;
    }
    WCU(cudaEventRecord(L.evA, st));
    apply_kernel<NT><<<(unsigned)((size_t)Rc * nel), NT, smem, st>>>(opc, sra, sel, r0, map, wgt, spec, sw, F, nd, nra, nel);
    WCU(cudaEventRecord(L.ev1, st));
    WCU(cudaGetLastError());
    WCU(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, L.evA, L.ev1);
    ms_kernel += ms;
    p->launches += 1;
  }
  WCU(cudaMemcpyAsync(p->spectrum, spec, nout * 16, cudaMemcpyDeviceToHost, st));
  WCU(cudaMemcpyAsync(p->spectrum_weight, sw, nout * 4, cudaMemcpyDeviceToHost, st));
  WCU(cudaStreamSynchronize(st));
  p->ms_kernel = ms_kernel;
  p->bytes = (double)nra * nel * (8.0 * FD + 16.0 * F + 20.0 * nd);
  return 0;
}
