// Device-side time / frequency cuts of a raw [B, T, nfreq] slab: DelayTransformBase._cut_data,
// draco/analysis/delay.py:516-597, for every baseline of the slab at once (the reference loops over baselines in
// Python, delay.py:1146-1167).  One CTA per baseline; a thread owns whole frequency columns and walks the time axis,
// which is coalesced across the CTA.  The doubly fancy-indexed arrays of delay.py:567-568 come out of numpy with
// the TIME axis contiguous, so numpy reduces every column with its pairwise summation (8 interleaved accumulators,
// blocks of 128 scalars); pw_sum below is that algorithm, and the cut data, the averaged weights and both masks come
// out bit-identical to the reference's (scale_freq excepted: its all-element std is one long pairwise sum in numpy
// and a tree reduction here, equal to rounding).
//
// Output layout: baseline b's kept block (T_b, M_b) row-major starts at data_out + b*T*nfreq; its M_b averaged
// weights and channel indices start at Ni_out / fsel_out + b*nfreq -- the ragged batch of dgibbs_problem with
// data_off = b*T*nfreq and chan_off = b*nfreq, so dgibbs_run consumes it in place (no host round trip).
#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/dgibbs.h"
#include "internal.h"

namespace dg {

constexpr int CUT_THREADS = 256;
constexpr int CUT_WARPS = CUT_THREADS / 32;

struct CutArgs {
  int B, T, F;
  int remove_mean, scale_freq;
  double time_frac, freq_frac, boost;
  const void* data;  // complex128 or complex64 (template parameter D)
  const void* weight;
  const long long* cind;
  double2* dout;
  double* Ni;
  long long* fsel;
  int* Tout;
  int* Mout;
  unsigned char* alive;
  unsigned char* kt;
  unsigned char* kf;
};

// mean over the kept samples and weight_boost in the arithmetic numpy uses for the weight dtype (delay.py:589-593):
// float32 weights are summed, divided and boosted in float32
__device__ __forceinline__ double boosted_mean(float s, int n, double boost) {
  return (double)(((float)((double)s / (double)n)) * (float)boost);
}
__device__ __forceinline__ double boosted_mean(double s, int n, double boost) { return (s / (double)n) * boost; }

template <typename R>
struct cplx_t {
  R x, y;
};
template <typename R>
__device__ __forceinline__ cplx_t<R> operator+(cplx_t<R> a, cplx_t<R> b) {
  return cplx_t<R>{a.x + b.x, a.y + b.y};
}
using cplx = cplx_t<double>;
template <typename R>
struct vec2_of;
template <>
struct vec2_of<double> {
  using type = double2;
};
template <>
struct vec2_of<float> {
  using type = float2;
};

// numpy's pairwise summation (numpy/_core/src/umath/loops_utils.h, pairwise_sum_@TYPE@) of g(s) .. g(s + n - 1).
// L accumulators interleaved over consecutive elements: 8 for real scalars, 4 for complex (= 8 scalars).
template <typename V, int L, class Get>
__device__ __forceinline__ V pw_leaf(Get g, int s, int n) {
  if (n < L) {
    V r{};
    for (int i = 0; i < n; i++) r = r + g(s + i);
    return r;
  }
  V r[L];
#pragma unroll
  for (int j = 0; j < L; j++) r[j] = g(s + j);
  int i = L;
  for (; i < n - n % L; i += L) {
#pragma unroll
    for (int j = 0; j < L; j++) r[j] = r[j] + g(s + i + j);
  }
  V res;
  if constexpr (L == 8)
    res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
  else
    res = (r[0] + r[1]) + (r[2] + r[3]);
  for (; i < n; i++) res = res + g(s + i);
  return res;
}
template <typename V, int L, class Get>
__device__ V pw_sum(Get g, int n) {
  constexpr int BLK = 16 * L;  // 128 scalars
  if (n <= BLK) return pw_leaf<V, L>(g, 0, n);
  // recursion of the reference algorithm (halves rounded down to a multiple of 8 scalars) with an explicit stack
  struct Frame {
    int s, n, stage;
    V left;
  };
  Frame st[24];
  int sp = 0;
  st[0] = Frame{0, n, 0, V{}};
  V ret{};
  while (sp >= 0) {
    Frame& f = st[sp];
    const int half = L == 8 ? (f.n / 2 - (f.n / 2) % 8) : (f.n - f.n % 8) / 2;
    if (f.stage == 0) {
      if (f.n <= BLK) {
        ret = pw_leaf<V, L>(g, f.s, f.n);
        sp--;
      } else {
        f.stage = 1;
        st[sp + 1] = Frame{f.s, half, 0, V{}};
        sp++;
      }
    } else if (f.stage == 1) {
      f.left = ret;
      f.stage = 2;
      st[sp + 1] = Frame{f.s + half, f.n - half, 0, V{}};
      sp++;
    } else {
      ret = f.left + ret;
      sp--;
    }
  }
  return ret;
}

// sum over the CTA, broadcast to every thread (deterministic order)
__device__ double cta_sum(double v, double* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double s = 0.0;
  for (int w = 0; w < CUT_WARPS; w++) s += red[w];
  return s;
}

// W: weight scalar (float / double); D: scalar of the complex visibilities (double = complex128, float = complex64).
// A complex64 slab is cut in the arithmetic numpy uses for it (delay.py:574 on a complex64 array): the column sums are
// float32 pairwise sums, the division by the count runs in complex128 (the count is an int64 scalar, so numpy picks the
// complex128 loop and casts the quotient back), the subtraction is float32; the result is widened to complex128 exactly
// as _complex_to_alternating_real does afterwards (delay.py:1631).
template <typename W, typename D>
__global__ void __launch_bounds__(CUT_THREADS) cut_data_kernel(CutArgs a) {
  using D2 = typename vec2_of<D>::type;
  extern __shared__ __align__(16) unsigned char cut_smem[];
  const int T = a.T, F = a.F, b = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* wavg = reinterpret_cast<double*>(cut_smem);  // [F] boosted mean weight of a frequency
  double* red = wavg + F;                              // [16]
  int* rowcnt = reinterpret_cast<int*>(red + 16);      // [T] positive weights of a time sample
  int* tlist = rowcnt + T;                             // [T] kept time samples, ascending
  int* fflag = tlist + T;                              // [F]
  int* flist = fflag + F;                              // [F] kept frequencies, ascending
  int* sc = flist + F;                                 // [16]
  const W* w = reinterpret_cast<const W*>(a.weight) + (size_t)b * T * F;
  const D2* dat = reinterpret_cast<const D2*>(a.data) + (size_t)b * T * F;

  // ---- non_zero_time: fraction of positive weights of each sample > time_frac (delay.py:548-554)
  for (int t = warp; t < T; t += CUT_WARPS) {
    int c = 0;
    for (int f = lane; f < F; f += 32) c += w[(size_t)t * F + f] > (W)0 ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) rowcnt[t] = c;
  }
  __syncthreads();
  if (warp == 0) {
    int n = 0;
    bool any = false;
    for (int t0 = 0; t0 < T; t0 += 32) {
      const int t = t0 + lane;
      const int c = t < T ? rowcnt[t] : 0;
      const bool k = t < T && ((double)c / (double)F > a.time_frac);
      any = any || c > 0;
      const unsigned m = __ballot_sync(0xffffffffu, k);
      if (k) tlist[n + __popc(m & ((1u << lane) - 1u))] = t;
      n += __popc(m);
      if (t < T && a.kt) a.kt[(size_t)b * T + t] = k ? 1 : 0;
    }
    any = __any_sync(0xffffffffu, any);
    if (lane == 0) {
      sc[0] = n;
      sc[1] = any ? 1 : 0;
    }
  }
  __syncthreads();
  const int nkt = sc[0];
  const bool any_weight = sc[1] != 0;

  // ---- non_zero_freq over the kept samples (delay.py:556-562) and the time-averaged weights (delay.py:589-593)
  for (int f = tid; f < F; f += CUT_THREADS) {
    int c = 0;
    for (int i = 0; i < nkt; i++) c += w[(size_t)tlist[i] * F + f] > (W)0 ? 1 : 0;
    const W s = pw_sum<W, 8>([&](int i) { return w[(size_t)tlist[i] * F + f]; }, nkt);
    const bool k = (double)c / (double)nkt > a.freq_frac;  // no kept sample: 0/0 = NaN compares false, as in numpy
    fflag[f] = k ? 1 : 0;
    wavg[f] = nkt > 0 ? boosted_mean(s, nkt, a.boost) : 0.0;
    if (a.kf) a.kf[(size_t)b * F + f] = k ? 1 : 0;
  }
  __syncthreads();
  int M = 0;
  for (int f0 = 0; f0 < F; f0 += CUT_THREADS) {  // ordered compaction of the kept frequencies
    const int f = f0 + tid;
    const bool k = f < F && fflag[f];
    const unsigned m = __ballot_sync(0xffffffffu, k);
    if (lane == 0) sc[4 + warp] = __popc(m);
    __syncthreads();
    int off = M, tot = 0;
    for (int x = 0; x < CUT_WARPS; x++) {
      if (x < warp) off += sc[4 + x];
      tot += sc[4 + x];
    }
    if (k) flist[off + __popc(m & ((1u << lane) - 1u))] = f;
    M += tot;
    __syncthreads();
  }
  if (!any_weight || M == 0) {  // delay.py:545-546, 564-565
    if (tid == 0) {
      a.alive[b] = 0;
      a.Tout[b] = 0;
      a.Mout[b] = 0;
    }
    return;
  }

  // ---- mean removal (delay.py:570-574) and the all-zero exit (delay.py:576-578)
  double2* out = a.dout + (size_t)b * T * F;
  const double scl = 1.0 / (double)nkt;  // numpy divides the complex sum by n as (re * (1/n), im * (1/n))
  bool allzero = true;
  double sr_all = 0.0, si_all = 0.0;
  for (int m = tid; m < M; m += CUT_THREADS) {
    const int f = flist[m];
    D mr = (D)0, mi = (D)0;
    if (a.remove_mean) {
      const cplx_t<D> sm = pw_sum<cplx_t<D>, 4>([&](int i) {
        const D2 v = dat[(size_t)tlist[i] * F + f];
        return cplx_t<D>{v.x, v.y};
      }, nkt);
      mr = (D)((double)sm.x * scl);
      mi = (D)((double)sm.y * scl);
    }
    for (int i = 0; i < nkt; i++) {
      const D2 raw = dat[(size_t)tlist[i] * F + f];
      const D dx = raw.x - mr, dy = raw.y - mi;  // in the input precision, then widened
      double2 v = make_double2((double)dx, (double)dy);
      if (v.x != 0.0 || v.y != 0.0) allzero = false;
      out[(size_t)i * M + m] = v;
      sr_all += v.x;
      si_all += v.y;
    }
    a.Ni[(size_t)b * F + m] = wavg[f];
    a.fsel[(size_t)b * F + m] = a.cind[f];
  }
  const int dead = __syncthreads_and(allzero ? 1 : 0);
  if (tid == 0) {
    a.alive[b] = dead ? 0 : 1;
    a.Tout[b] = dead ? 0 : nkt;
    a.Mout[b] = dead ? 0 : M;
  }
  if (dead || !a.scale_freq) return;

  // ---- scale_freq (delay.py:580-587): data *= invert_no_zero(std over time / std over everything)
  const double cnt_all = (double)nkt * (double)M;
  const double scl_all = 1.0 / cnt_all;
  const double mra = cta_sum(sr_all, red) * scl_all, mia = cta_sum(si_all, red) * scl_all;
  double dev_all = 0.0;
  for (int m = tid; m < M; m += CUT_THREADS) {
    for (int i = 0; i < nkt; i++) {
      const double2 v = out[(size_t)i * M + m];
      const double dx = v.x - mra, dy = v.y - mia;
      dev_all += __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
    }
    const cplx sm = pw_sum<cplx, 4>([&](int i) {
      const double2 v = out[(size_t)i * M + m];
      return cplx{v.x, v.y};
    }, nkt);
    const double mr = sm.x * scl, mi = sm.y * scl;
    const double dv = pw_sum<double, 8>([&](int i) {
      const double2 v = out[(size_t)i * M + m];
      const double dx = v.x - mr, dy = v.y - mi;
      return __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
    }, nkt);
    wavg[m] = sqrt(dv / (double)nkt);  // std over time of column m (wavg is free: the weights are written)
  }
  const double std_all = sqrt(cta_sum(dev_all, red) / cnt_all);
  for (int m = tid; m < M; m += CUT_THREADS) {
    const double rel = wavg[m] / std_all;
    const double inv = rel != 0.0 ? 1.0 / rel : 0.0;
    for (int i = 0; i < nkt; i++) {
      double2 v = out[(size_t)i * M + m];
      v.x *= inv;
      v.y *= inv;
      out[(size_t)i * M + m] = v;
    }
  }
}

}  // namespace dg

using namespace dg;

#define KCU(call)                                                                                             \
  do {                                                                                                        \
    cudaError_t e_ = (call);                                                                                  \
    if (e_ != cudaSuccess) {                                                                                  \
      char msg_[512];                                                                                         \
      snprintf(msg_, sizeof msg_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      dg_set_error(h, msg_);                                                                                  \
      return (int)e_;                                                                                         \
    }                                                                                                         \
  } while (0)

// Replaces the per-baseline Python loop around DelayTransformBase._cut_data (delay.py:516-597, 1146-1167).
extern "C" int dgibbs_cut_data(dgibbs_handle h, const dgibbs_cut* c) {
  if (!h) return -1;
  auto bad = [&](const char* what) {
    dg_set_error(h, (std::string("bad argument: ") + what).c_str());
    return -2;
  };
  if (!c || c->struct_size != (int32_t)sizeof(dgibbs_cut)) return bad("cut struct");
  if (c->B <= 0 || c->T <= 0 || c->nfreq <= 0) return bad("B, T, nfreq must be positive");
  if (c->data_c64 && c->scale_freq)
    return bad("scale_freq on complex64 data is not served on the device (numpy's float32 std is not reproduced); cut on the host");
  if (!c->data || !c->weight || !c->channel_ind) return bad("data, weight and channel_ind are required");
  if (!c->data_out || !c->Ni_out || !c->fsel_out || !c->T_out || !c->M_out || !c->alive)
    return bad("data_out, Ni_out, fsel_out, T_out, M_out and alive are required");
  KCU(cudaSetDevice(dg_device(h)));
  cudaStream_t st = dg_stream(h);
  const size_t B = c->B, T = c->T, F = c->nfreq, wsz = c->weight_f32 ? 4 : 8, dsz = c->data_c64 ? 8 : 16;
  const bool dev = c->on_device != 0;

  CutArgs a{};
  a.B = c->B; a.T = c->T; a.F = c->nfreq;
  a.remove_mean = c->remove_mean; a.scale_freq = c->scale_freq;
  a.time_frac = c->time_frac; a.freq_frac = c->freq_frac; a.boost = c->weight_boost;
  void* q;
  auto in = [&](const char* name, const void* src, size_t bytes, const void** out) -> cudaError_t {
    if (dev) { *out = src; return cudaSuccess; }
    cudaError_t e = dg_pool_get(h, name, bytes, &q);
    if (e != cudaSuccess) return e;
    *out = q;
    return cudaMemcpyAsync(q, src, bytes, cudaMemcpyHostToDevice, st);
  };
  auto outbuf = [&](const char* name, void* dst, size_t bytes, void** out) -> cudaError_t {
    if (dev || !dst) { *out = dst; return cudaSuccess; }
    return dg_pool_get(h, name, bytes, out);
  };
  const void* pin;
  KCU(in("cut_data", c->data, B * T * F * dsz, &pin)); a.data = pin;
  KCU(in("cut_weight", c->weight, B * T * F * wsz, &pin)); a.weight = pin;
  KCU(in("cut_cind", c->channel_ind, F * 8, &pin)); a.cind = (const long long*)pin;
  void* po;
  KCU(outbuf("cut_dout", c->data_out, B * T * F * 16, &po)); a.dout = (double2*)po;
  KCU(outbuf("cut_Ni", c->Ni_out, B * F * 8, &po)); a.Ni = (double*)po;
  KCU(outbuf("cut_fsel", c->fsel_out, B * F * 8, &po)); a.fsel = (long long*)po;
  KCU(outbuf("cut_T", c->T_out, B * 4, &po)); a.Tout = (int*)po;
  KCU(outbuf("cut_M", c->M_out, B * 4, &po)); a.Mout = (int*)po;
  KCU(outbuf("cut_alive", c->alive, B, &po)); a.alive = (unsigned char*)po;
  KCU(outbuf("cut_kt", c->keep_t, B * T, &po)); a.kt = (unsigned char*)po;
  KCU(outbuf("cut_kf", c->keep_f, B * F, &po)); a.kf = (unsigned char*)po;

  const size_t smem = 8 * (F + 16) + 4 * (2 * T + 2 * F + 16);
  if (smem > 200 * 1024) return bad("T and nfreq too large for the cut kernel's shared memory");
  void (*kern)(CutArgs) = c->data_c64 ? (c->weight_f32 ? cut_data_kernel<float, float> : cut_data_kernel<double, float>)
                                      : (c->weight_f32 ? cut_data_kernel<float, double> : cut_data_kernel<double, double>);
  KCU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)B, CUT_THREADS, smem, st>>>(a);
  KCU(cudaGetLastError());
  if (!dev) {
    auto back = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
      return dst ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess;
    };
    KCU(back(c->data_out, a.dout, B * T * F * 16));
    KCU(back(c->Ni_out, a.Ni, B * F * 8));
    KCU(back(c->fsel_out, a.fsel, B * F * 8));
    KCU(back(c->T_out, a.Tout, B * 4));
    KCU(back(c->M_out, a.Mout, B * 4));
    KCU(back(c->alive, a.alive, B));
    KCU(back(c->keep_t, a.kt, B * T));
    KCU(back(c->keep_f, a.kf, B * F));
  }
  KCU(cudaStreamSynchronize(st));
  return 0;
}
