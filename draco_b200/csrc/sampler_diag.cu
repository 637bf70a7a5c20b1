// Diagonal(+rank-1) branch of the Gibbs sampler: a streaming, HBM-bound kernel.
//
// With uniform weights, no window and a complete channel set, FTNiF = a I + b u u^T (u_k = (-1)^k;
// b = 0 with the Nyquist channel present, b = -Ni without it: SURVEY 8d), so
//     Ci = D + b u u^T,  D = diag(a + 1/S)
// and the signal draw x ~ N(Ci^-1 yhat, Ci^-1), yhat = F^T N^-1/2 z (projected data, computed once),
// needs no factorisation:
//     x = D^-1 (yhat - beta u P + gam u E) + D^-1/2 e,      e ~ N(0, I)
//     P = u^T D^-1 yhat,  E = u^T D^-1/2 e,  beta = b / (1 + b q),  gam = (1/sqrt(1 + b q) - 1) / q,
//     q = u^T D^-1 u
// (Sherman-Morrison for the mean, a rank-1 corrected square root for the covariance).  Same
// posterior as _draw_signal_sample_f (delay.py:1778-1819); the random stream is native to this kernel
// (one normal per element instead of N + 2M), so it is used in device-RNG mode only -- injected-draw
// parity of the same solve runs through diag_solve_rows in sampler_dense.cu.
//
// Per baseline-iteration the kernel reads yhat (8 n T bytes) once and reads/writes S (16 N): the
// algorithmic traffic Q_d of SURVEY 8d.  One CTA per baseline chain, all iterations on the device.
#include <cooperative_groups.h>

#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "internal.h"

namespace cg = cooperative_groups;

namespace dg {

constexpr int DTHREADS = 1024;
constexpr int DWARPS = DTHREADS / 32;
enum { STREAM_DIAG_E = 4, STREAM_DIAG_CHI2 = 5 };

// four fp32 standard normals from one Philox block (Box-Muller with fast intrinsics)
__device__ __forceinline__ void normal4_f32(const uint4& r, float& z0, float& z1, float& z2, float& z3) {
  const float k = 2.3283064365386963e-10f;  // 2^-32
  float u1 = ((float)(r.x >> 8) + 0.5f) * 5.9604644775390625e-8f;  // (0,1), 24 bits
  float u3 = ((float)(r.z >> 8) + 0.5f) * 5.9604644775390625e-8f;
  float ra = sqrtf(-2.0f * __logf(u1)), rb = sqrtf(-2.0f * __logf(u3));
  float s, c;
  __sincosf(6.2831853071795865f * ((float)r.y * k), &s, &c);
  z0 = ra * c;
  z1 = ra * s;
  __sincosf(6.2831853071795865f * ((float)r.w * k), &s, &c);
  z2 = rb * c;
  z3 = rb * s;
}

// chi^2 with T degrees of freedom = 2 Gamma(T/2) (Marsaglia-Tsang); T == 1 is a squared normal.
__device__ double chi2_draw(const Philox& ph, unsigned it, unsigned long long k, int T) {
  if (T == 1) {
    double z0, z1;
    normal_pair(ph((unsigned)k, (unsigned)(k >> 32), it, STREAM_DIAG_CHI2), z0, z1);
    return z0 * z0;
  }
  const double alpha = 0.5 * T, dd = alpha - 1.0 / 3.0, cc = 1.0 / sqrt(9.0 * dd);
  for (unsigned attempt = 0; attempt < 64; attempt++) {
    uint4 r = ph((unsigned)k, (unsigned)(k >> 32) | (attempt << 16), it, STREAM_DIAG_CHI2);
    double u1 = u01(r.x, r.y);
    double s, c;
    sincospi(2.0 * ((double)r.z * 2.3283064365386963e-10), &s, &c);
    double x = sqrt(-2.0 * log(u1)) * c;
    double v = 1.0 + cc * x;
    if (v <= 0.0) continue;
    v = v * v * v;
    double u = ((double)r.w + 0.5) * 2.3283064365386963e-10;
    double x2 = x * x;
    if (u < 1.0 - 0.0331 * x2 * x2 || log(u) < 0.5 * x2 + dd * (1.0 - v + log(v))) return 2.0 * dd * v;
  }
  return 2.0 * dd;  // unreachable in practice (acceptance > 95 % per attempt)
}

// block-wide sums of NV values per thread; result broadcast to every thread (2 barriers)
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < NV; i++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
  }
  __syncthreads();  // previous readers of red are done
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NV; i++) red[warp * NV + i] = v[i];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < NV; i++) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < DWARPS; w++) s += red[w * NV + i];
    v[i] = s;
  }
}

// NPAIR: column pairs (k = 2p, 2p+1) owned by a thread, p = tid + 1024 i;  TC: samples per pass (even).
// 1024 threads x <= 64 registers: one CTA fills an SM (32 warps) so the yhat stream has the whole
// SM's memory-level parallelism behind it.
template <int NPAIR, int TC>
__global__ void __launch_bounds__(DTHREADS, 1) diag_sampler_kernel(DevBatch d, int it_begin, int it_end) {
  __shared__ double red[DWARPS * 2 * TC];
  const int tid = threadIdx.x;
  const int N = d.N, nsys = d.nsys, npairs = nsys >> 1;
  for (int li = blockIdx.x; li < d.nlist; li += gridDim.x) {
    const int b = d.blist[li];
    if (d.failed[b]) continue;
    const int Tb = d.T[b];
    const double a = d.diag_ab[2 * b], bc = d.diag_ab[2 * b + 1];
    const double* yh = d.yhat + (size_t)d.ycompact[b] * d.Tmax * d.ypitch;
    double* S = d.S + (size_t)b * N;
    const Philox ph(mix_key(d.seed, d.gid ? (unsigned long long)d.gid[b] : (unsigned long long)b));

    for (int it = it_begin; it < it_end; it++) {
      double invD[NPAIR][2], rD[NPAIR][2], sum[NPAIR][2], sq[NPAIR][2];
      double ql[1] = {0.0};
      bool pos = true;
#pragma unroll
      for (int i = 0; i < NPAIR; i++) {
        const int p = tid + DTHREADS * i;
#pragma unroll
        for (int e = 0; e < 2; e++) {
          invD[i][e] = rD[i][e] = sum[i][e] = sq[i][e] = 0.0;
          if (p < npairs) {
            const int k = 2 * p + e;
            double s = S[d.cplx ? p : k];
            double si = (s != 0.0) ? 1.0 / s : 0.0;  // invert_no_zero, delay.py:1785
            if (d.cplx) si *= 2.0;                   // delay.py:1794
            double D = a + si;
            if (!(D > 0.0)) pos = false;
            invD[i][e] = 1.0 / D;
            rD[i][e] = sqrt(invD[i][e]);
            ql[0] += invD[i][e];
          }
        }
      }
      block_sum<1>(ql, red);
      const int allpos = __syncthreads_and(pos ? 1 : 0);
      const double q = ql[0], den = 1.0 + bc * q;
      if (!allpos || !(den > 0.0)) {  // not positive definite: the chain stops here (delay.py:1891-1896)
        if (tid == 0) {
          d.failed[b] = 1;
          d.n_done[b] = it;
        }
        break;
      }
      const double beta = bc / den;
      const double gam = (bc != 0.0) ? (1.0 / sqrt(den) - 1.0) / q : 0.0;

      for (int t0 = 0; t0 < Tb; t0 += TC) {
        double y[NPAIR][2][TC];
        float ev[NPAIR][2][TC];
        double pe[2 * TC];
#pragma unroll
        for (int j = 0; j < 2 * TC; j++) pe[j] = 0.0;
#pragma unroll
        for (int i = 0; i < NPAIR; i++) {
          const int p = tid + DTHREADS * i;
          if (p < npairs) {
#pragma unroll
            for (int tt = 0; tt < TC; tt += 2) {
              unsigned long long blk = (unsigned long long)((t0 + tt) >> 1) * npairs + p;
              normal4_f32(ph((unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_DIAG_E), ev[i][0][tt], ev[i][1][tt],
                          ev[i][0][tt + 1], ev[i][1][tt + 1]);
            }
#pragma unroll
            for (int tt = 0; tt < TC; tt++) {
              const int t = t0 + tt;
              double2 v = make_double2(0.0, 0.0);
              if (t < Tb) v = *reinterpret_cast<const double2*>(yh + (size_t)t * d.ypitch + 2 * p);
              y[i][0][tt] = v.x;
              y[i][1][tt] = v.y;
              pe[tt] += invD[i][0] * v.x - invD[i][1] * v.y;                                            // P_t
              pe[TC + tt] += rD[i][0] * (double)ev[i][0][tt] - rD[i][1] * (double)ev[i][1][tt];         // E_t
            }
          } else {
#pragma unroll
            for (int tt = 0; tt < TC; tt++) {
              y[i][0][tt] = y[i][1][tt] = 0.0;
              ev[i][0][tt] = ev[i][1][tt] = 0.0f;
            }
          }
        }
        if (bc != 0.0) block_sum<2 * TC>(pe, red);  // uniform branch
#pragma unroll
        for (int tt = 0; tt < TC; tt++) {
          if (t0 + tt < Tb) {
            const double corr = gam * pe[TC + tt] - beta * pe[tt];
#pragma unroll
            for (int i = 0; i < NPAIR; i++) {
              double x0 = invD[i][0] * (y[i][0][tt] + corr) + rD[i][0] * (double)ev[i][0][tt];
              double x1 = invD[i][1] * (y[i][1][tt] - corr) + rD[i][1] * (double)ev[i][1][tt];
              sum[i][0] += x0; sq[i][0] += x0 * x0;
              sum[i][1] += x1; sq[i][1] += x1 * x1;
            }
          }
        }
      }

      // power-spectrum draw (delay.py:1862-1880)
      const double invT = 1.0 / (double)Tb;
#pragma unroll
      for (int i = 0; i < NPAIR; i++) {
        const int p = tid + DTHREADS * i;
        if (p < npairs) {
          double v0 = (sq[i][0] - sum[i][0] * sum[i][0] * invT) * invT;
          double v1 = (sq[i][1] - sum[i][1] * sum[i][1] * invT) * invT;
          if (d.cplx) {
            double snew = (v0 + v1) * (double)Tb / chi2_draw(ph, (unsigned)it, (unsigned long long)p, Tb);
            S[p] = snew;
            d.samples[((size_t)it * d.B + b) * N + p] = snew;
          } else {
            double s0 = v0 * (double)Tb / chi2_draw(ph, (unsigned)it, (unsigned long long)(2 * p), Tb);
            double s1 = v1 * (double)Tb / chi2_draw(ph, (unsigned)it, (unsigned long long)(2 * p + 1), Tb);
            *reinterpret_cast<double2*>(S + 2 * p) = make_double2(s0, s1);
            *reinterpret_cast<double2*>(d.samples + ((size_t)it * d.B + b) * N + 2 * p) = make_double2(s0, s1);
          }
        }
      }
      if (tid == 0) d.n_done[b] = it + 1;
      __syncthreads();  // S of this iteration is read by other threads in the next one (cplx: no; keep simple)
    }
  }
}


// ============================================================================ cluster-resident variant
// Two observations make the diagonal branch far cheaper than streaming yhat once per iteration:
//
// 1. Residency.  The projected data yhat of one baseline (8 n T bytes: 1 MiB at cfg3) does not fit one
//    SM's shared memory, but it fits the distributed shared memory of a thread-block cluster: C CTAs
//    each keep a column slice [T][n / C] resident for ALL iterations of the chain, so HBM is read once
//    per baseline instead of once per baseline-iteration.  Per iteration the only cross-CTA traffic is
//    an all-gather of the partial sums of P_t and q through DSMEM and one cluster barrier.
//
// 2. Collapsed moments.  The chain only uses x through  ssq_k = sum_t (x_tk - mean_t x_k)^2  (delay.py:1867).
//    For b <= 0 (the only values the reference's Fourier matrices produce: b = 0 with the Nyquist channel,
//    b = -Ni without it) the posterior covariance is  D^-1 + |beta| v v^T,  v = D^-1 u,  so
//        x_tk = m_tk + e_tk / sqrt(D_k),   m_tk = (y_tk + u_k corr_t) / D_k,
//        corr_t = -beta P_t + sqrt(|beta|) xi_t,   e_tk, xi_t ~ N(0, 1) independent,
//    and, with m~ = m - mean_t m (orthogonal to the ones vector),
//        ssq_k = |m~_k|^2 + 2 |m~_k| g_k / sqrt(D_k) + (g_k^2 + c_k) / D_k,   g_k ~ N(0,1), c_k ~ chi^2_{T-2}
//    EXACTLY in distribution (split e along 1, m~ and the rest).  One normal and one chi^2 per delay
//    replace T normals, and |m~_k|^2 follows from per-baseline constants (sum_t y, sum_t y^2) plus one
//    matrix-vector product  r_k = sum_t y_tk corr_t  with the resident slice.
//
// b > 0 falls back to the streaming kernel above.
constexpr int RMAXC = 16;      // cluster sizes tried: 1, 2, 4, 8 (portable) and 16 (opt-in)
constexpr int RBATCH = 4;      // iterations whose scalar draws are generated together (ILP)
enum { STREAM_DIAG_XI = 6, STREAM_DIAG_G = 7, STREAM_DIAG_C2 = 8 };

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  unsigned ok = 0;
  while (!ok) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  }
}
// shared::cluster address of `saddr` (a shared::cta address) in CTA `rank` of the cluster
__device__ __forceinline__ unsigned mapa_u32(unsigned saddr, unsigned rank) {
  unsigned r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// remote 8-byte store whose completion is counted (in bytes) on the DESTINATION CTA's mbarrier: the
// consumer waits on its own barrier, no cluster-wide fence or barrier is involved
__device__ __forceinline__ void st_async_f64(unsigned raddr, double v, unsigned rmbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(raddr),
               "l"(__double_as_longlong(v)), "r"(rmbar)
               : "memory");
}
// TMA bulk copy global -> this CTA's shared memory, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// one fp32 standard normal (Box-Muller, fast intrinsics) from two 32-bit words
__device__ __forceinline__ float normal1_f32(unsigned a, unsigned b) {
  float u1 = ((float)(a >> 8) + 0.5f) * 5.9604644775390625e-8f;
  float s, c;
  __sincosf(6.2831853071795865f * ((float)b * 2.3283064365386963e-10f), &s, &c);
  return sqrtf(-2.0f * __logf(u1)) * c;
}
// chi^2 with `dof` degrees of freedom (Marsaglia-Tsang on 2 Gamma(dof/2)); the proposal normal and the
// acceptance test run in fp32, the variate itself is formed in fp64.  dof <= 0 gives 0.
__device__ __forceinline__ double chi2_fast(const Philox& ph, unsigned it, unsigned key, unsigned stream, int dof,
                                            bool have_first = false, unsigned fa = 0, unsigned fb = 0, unsigned fu = 0) {
  if (dof <= 0) return 0.0;
  if (dof == 1) {
    uint4 r = ph(key, 0u, it, stream);
    double z = (double)normal1_f32(r.x, r.y);
    return z * z;
  }
  const float dd = 0.5f * (float)dof - (1.0f / 3.0f), cc = rsqrtf(9.0f * dd);
  for (unsigned attempt = 0; attempt < 64; attempt++) {
    unsigned a = fa, b = fb, uu = fu;
    if (attempt > 0 || !have_first) {
      uint4 r = ph(key, attempt, it, stream);
      a = r.x; b = r.y; uu = r.z;
    }
    const float x = normal1_f32(a, b);
    const float v1 = 1.0f + cc * x;
    if (v1 <= 0.0f) continue;
    const float v = v1 * v1 * v1;
    const float u = ((float)(uu >> 8) + 0.5f) * 5.9604644775390625e-8f;
    const float x2 = x * x;
    if (u < 1.0f - 0.0331f * x2 * x2 || __logf(u) < 0.5f * x2 + dd * (1.0f - v + __logf(v))) {
      const double w = 1.0 + (double)cc * (double)x;
      return 2.0 * (double)dd * w * w * w;
    }
  }
  return 2.0 * (double)dd;  // unreachable in practice (acceptance > 95 % per attempt)
}

// corr_t: even length, and at least the register-resident row count (rows past Tb are read as zeros there)
__host__ __device__ inline int corr_len(int Tm) { return Tm < 64 ? 64 : ((Tm + 1) & ~1); }

struct ResidentGeom {
  int C;        // cluster size (0: does not fit)
  int ncol;     // system columns per CTA
  int threads;  // ncol rounded up to a warp
  int XQp;      // padded width of one exchange vector: P_t[Tmax], then q and bad partials per warp
  size_t smem;
};

// shared memory: [ysm: Tmax*ncol][wsm: ncol][corr: Tmax][xi: RBATCH*Tmax][qw: 64][scw: 64][xch: 2*C*XQp][mbar: 3]
// Geometry of a cluster of C CTAs: an even number of columns per CTA, C * ncol >= nsys (the tail columns are the zero
// padding of yhat rows).  C need not be a power of two: on B200 clusters of 6 keep 132 SMs busy, clusters of 8 only 120.
static ResidentGeom resident_geom_for(const DevBatch& d, size_t smem_cap, int C) {
  ResidentGeom g{};
  const int ncol = 2 * ((d.nsys + 2 * C - 1) / (2 * C));
  if (ncol > 1024 || (long long)C * ncol > d.ypitch) return g;
  const int XQp = (d.Tmax + 2 + 1) & ~1;  // P_t[Tmax], q, bad
  size_t doubles = (size_t)d.Tmax * ncol + (size_t)ncol + (size_t)corr_len(d.Tmax) + (size_t)RBATCH * d.Tmax + 128 +
                   2 * (size_t)C * XQp + 4;
  if (doubles * 8 > smem_cap) return g;
  g.C = C; g.ncol = ncol; g.threads = (ncol + 31) / 32 * 32; g.XQp = XQp; g.smem = doubles * 8;
  return g;
}

// Partial P_t = sum_k u_k y_tk / D_k over the resident columns.  A warp takes NR sample rows t0 + j W at a time
// (independent accumulators); each lane takes column pairs (LDS.128); rows past Tb read row Tb-1 and are dropped at
// the store.  A transposing butterfly leaves the total of row j in lanes [j 32/NR, (j+1) 32/NR), which send it
// straight to every CTA of the cluster (st.async completes bytes on the destination's exchange barrier).
// phase profiler of the resident kernel (build with -DDGIBBS_DIAG_PROF, run with DGIBBS_PROFILE=1): thread 0 of
// CTA 0 accumulates clock64 deltas.  Compiled out by default: the clock reads pin the instruction schedule (7% slower).
#ifndef DGIBBS_DIAG_PROF
#define DPROF(idx) do { } while (0)
#else
#define DPROF(idx)                                                                       \
  do {                                                                                   \
    if (d.prof && blockIdx.x == 0 && threadIdx.x == 0) {                                 \
      long long t_ = clock64();                                                          \
      d.prof[idx] += t_ - tlast;                                                         \
      tlast = t_;                                                                        \
    }                                                                                    \
  } while (0)
#endif

template <int NR>
__device__ __forceinline__ void partial_P(const double* ysm, const double* wsm, double* xmine, unsigned long long* xb,
                                          int Tb, int ncol, int C, int warp, int lane, int W) {
  static_assert(NR == 4 || NR == 8, "rows per pass");
  const double2* w2 = reinterpret_cast<const double2*>(wsm);
  for (int t0 = warp; t0 < Tb; t0 += NR * W) {
    const double2* yr[NR];
    double acc[NR];
#pragma unroll
    for (int j = 0; j < NR; j++) {
      acc[j] = 0.0;
      yr[j] = reinterpret_cast<const double2*>(ysm + (size_t)min(t0 + j * W, Tb - 1) * ncol);
    }
    for (int c = lane; c < (ncol >> 1); c += 32) {
      const double2 wv = w2[c];
#pragma unroll
      for (int j = 0; j < NR; j++) {
        const double2 yv = yr[j][c];
        acc[j] = fma(wv.x, yv.x, fma(wv.y, yv.y, acc[j]));
      }
    }
    int bit = 16;
#pragma unroll
    for (int n = NR / 2; n >= 1; n >>= 1, bit >>= 1) {
      const bool hi = lane & bit;
#pragma unroll
      for (int j = 0; j < n; j++) {
        const double send = hi ? acc[j] : acc[j + n], keep = hi ? acc[j + n] : acc[j];
        acc[j] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
      }
    }
#pragma unroll
    for (int o = 16 / NR; o > 0; o >>= 1) acc[0] += __shfl_xor_sync(0xffffffffu, acc[0], o);
    constexpr int G = 32 / NR;  // lanes holding the same row total
    const int t = t0 + (lane / G) * W;
    if (t < Tb)
      for (int dst = lane % G; dst < C; dst += G)
        st_async_f64(mapa_u32(smem_u32(xmine + t), dst), acc[0], mapa_u32(smem_u32(xb), dst));
  }
}

// YREG > 0: every thread also keeps its own column of yhat (YREG >= Tb samples) in registers, so the moment pass reads
// no shared memory but the broadcast corr_t.
template <int MAXT, int YREG>
__global__ void __launch_bounds__(MAXT, 1) diag_resident_kernel(DevBatch d, int it_begin, int it_end, int C, int ncol,
                                                                 int XQp) {
  extern __shared__ __align__(16) double dsm[];
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int cid = blockIdx.x / C, ncl = gridDim.x / C;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, W = blockDim.x >> 5;
  const int N = d.N, nsys = d.nsys, Tm = d.Tmax;
  double* ysm = dsm;                                   // resident slice [Tb][ncol]
  double* wsm = ysm + (size_t)Tm * ncol;               // u_k / D_k of the local columns
  double* corr = wsm + ncol;                           // corr_t
  double* xi = corr + corr_len(Tm);                 // [RBATCH][Tm] xi_t of the current batch of iterations
  double* qw = xi + RBATCH * Tm;                       // [64] per-warp partials of q and of the non-positive count
  double* scw = qw + 64;                               // [64] per-warp partials of sum corr, sum corr^2
  double* xch = scw + 64;                               // [2][C][XQp] gathered partial vectors
  unsigned long long* mbar = reinterpret_cast<unsigned long long*>(xch + 2 * (size_t)C * XQp);
  const int col0 = rank * ncol;                        // first system column of this CTA (even)
  const int k = col0 + tid;                            // this thread's system column
  const bool owner = tid < ncol && k < nsys;
  const double sg = (tid & 1) ? -1.0 : 1.0;            // u_k = (-1)^k  (col0 is even)

  unsigned long long* xbar = mbar + 1;  // [2] exchange barriers, one per parity
  if (tid == 0) {
    mbar_init(mbar, 1);
    mbar_init(xbar, 1);
    mbar_init(xbar + 1, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  cluster.sync();  // every CTA's barriers exist before a peer signals them
  unsigned phase = 0, xphase = 0;  // xphase bit p: phase parity of xbar[p]
  int par = 0;

  for (int li = cid; li < d.nlist; li += ncl) {
    const int b = d.blist[li];
    if (d.failed[b]) continue;  // uniform across the cluster (set by an earlier launch only)
    const int Tb = d.T[b];
    const double a = d.diag_ab[2 * b], bc = d.diag_ab[2 * b + 1];
    const double* yh = d.yhat + (size_t)d.ycompact[b] * Tm * d.ypitch + col0;
    double* S = d.S + (size_t)b * N;
    const Philox ph(mix_key(d.seed, d.gid ? (unsigned long long)d.gid[b] : (unsigned long long)b));
    const unsigned dkey = (unsigned)(d.cplx ? (k >> 1) : k);  // delay index: key of the chi^2_T draw

    // ---- the column slice of yhat becomes resident (TMA bulk copies, one per sample row)
    if (tid == 0) {
      mbar_expect_tx(mbar, (unsigned)(Tb * ncol * 8));
      for (int t = 0; t < Tb; t++) bulk_g2s(ysm + (size_t)t * ncol, yh + (size_t)t * d.ypitch, (unsigned)(ncol * 8), mbar);
    }
    double Sk = owner ? S[dkey] : 0.0;
    mbar_wait(mbar, phase);
    phase ^= 1;
    // per-baseline constants of the owned column: sum_t y, sum_t y^2
    double sy = 0.0, syy = 0.0;
    [[maybe_unused]] double ycol[YREG > 0 ? YREG : 1];
    if constexpr (YREG > 0) {
#pragma unroll
      for (int t = 0; t < YREG; t++) {
        const double v = (owner && t < Tb) ? ysm[(size_t)t * ncol + tid] : 0.0;
        ycol[t] = v;
        sy += v;
        syy += v * v;
      }
      for (int t = Tb + tid; t < YREG; t += blockDim.x) corr[t] = 0.0;  // rows past Tb: finite, times ycol = 0
    } else if (owner) {
      for (int t = 0; t < Tb; t++) {
        const double v = ysm[(size_t)t * ncol + tid];
        sy += v;
        syy += v * v;
      }
    }
    const int nwc = min(W, (Tb + 31) >> 5);  // warps that own a sample index t < Tb (thread t forms corr_t)
    const double sfac = d.cplx ? 2.0 : 1.0, inv_a = 1.0 / a, inv_T = 1.0 / (double)Tb;  // delay.py:1794
    double iv = 0.0;
    // A: D = a + 1/S on the owned column; per-warp partials of q = sum 1/D and of the non-positive count
    auto stage_A = [&]() {
      double ql = 0.0, badl = 0.0;
      iv = 0.0;
      if (owner) {
        // D = a + s / S_k (invert_no_zero: 1/0 -> 0, delay.py:1785) and 1/D from one division: 1/D = S_k / (a S_k + s)
        const double den = fma(a, Sk, sfac);
        const bool zero = Sk == 0.0 || fabs(Sk) == INFINITY;  // s / S_k = 0 either way
        iv = zero ? inv_a : Sk / den;
        if (!(iv > 0.0) || !(a > 0.0 || !zero)) badl = 1.0;  // D > 0 <=> 1/D > 0
        ql = iv;
      }
      if (tid < ncol) wsm[tid] = sg * iv;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ql += __shfl_xor_sync(0xffffffffu, ql, o);
      const bool anybad = __any_sync(0xffffffffu, badl != 0.0);
      if (lane == 0) {
        qw[warp] = ql;
        qw[32 + warp] = anybad ? 1.0 : 0.0;
      }
    };
    stage_A();
    __syncthreads();

    // scalar draws of one iteration: g ~ N(0,1), c2 ~ chi^2_{T-2}, 1 / chi^2_T of the owned column, and xi_t (thread t).
    // Two Philox blocks feed g and the first attempts of both chi^2 draws (real mode; in complex mode the
    // chi^2_T draw is shared by the re/im columns of a delay and keyed on the delay).
    double g_cur = 0.0, c2_cur = 0.0, ict_cur = 0.0, g_nxt = 0.0, c2_nxt = 0.0, ict_nxt = 0.0;
    auto draw_scalars = [&](int itx, double& g, double& c2, double& ict) {
      if (owner) {
        const uint4 r = ph((unsigned)k, 0u, (unsigned)itx, STREAM_DIAG_G);
        const uint4 r2 = ph((unsigned)k, 1u, (unsigned)itx, STREAM_DIAG_G);
        g = (double)normal1_f32(r.x, r.y);
        c2 = chi2_fast(ph, (unsigned)itx, (unsigned)k, STREAM_DIAG_C2, Tb - 2, true, r.z, r.w, r2.z);
        ict = 1.0 / chi2_fast(ph, (unsigned)itx, dkey, STREAM_DIAG_CHI2, Tb, !d.cplx, r2.x, r2.y, r2.w);
      }
      if (bc != 0.0)
        for (int t = tid; t < Tb; t += blockDim.x) {
          uint4 r = ph((unsigned)t, 0u, (unsigned)itx, STREAM_DIAG_XI);
          xi[(itx & 1) * Tm + t] = (double)normal1_f32(r.x, r.y);
        }
    };
    draw_scalars(it_begin, g_cur, c2_cur, ict_cur);
#ifdef DGIBBS_DIAG_PROF
    long long tlast = clock64();
#endif
    for (int it = it_begin; it < it_end; it++) {
      DPROF(7);
      // ---- B: partial P_t = sum_k u_k y_tk / D_k over the local columns; a warp takes 8 sample rows at a
      // time (independent accumulators) and sends the sums straight to every CTA of the cluster
      double* xmine = xch + (size_t)(par * C + rank) * XQp;
      if (warp == 0 && lane < C) {
        double qs = 0.0, bs = 0.0;
        for (int w = 0; w < W; w++) {
          qs += qw[w];
          bs += qw[32 + w];
        }
        const unsigned rb = mapa_u32(smem_u32(xbar + par), lane), ra = mapa_u32(smem_u32(xmine + Tm), lane);
        st_async_f64(ra, qs, rb);
        st_async_f64(ra + 8, bs, rb);
      }
      if (tid == 32 % blockDim.x)  // arm this phase: q and bad from every CTA, plus T partial sums each if b != 0
        mbar_expect_tx(xbar + par, (unsigned)(C * ((bc != 0.0 ? Tb : 0) + 2) * 8));
      if (bc != 0.0) {
        if (4 * W >= Tb)
          partial_P<4>(ysm, wsm, xmine, xbar + par, Tb, ncol, C, warp, lane, W);
        else
          partial_P<8>(ysm, wsm, xmine, xbar + par, Tb, ncol, C, warp, lane, W);
      }
      DPROF(0);
      // the next iteration's scalar draws overlap the all-gather latency
      draw_scalars(it + 1, g_nxt, c2_nxt, ict_nxt);
      DPROF(1);
      mbar_wait(xbar + par, (xphase >> par) & 1u);
      DPROF(2);
      xphase ^= 1u << par;
      // ---- D: every thread forms q, the failure flag and beta; thread t forms corr_t
      const double* xb = xch + (size_t)par * C * XQp;
      double q = 0.0, bad = 0.0;
      for (int r = 0; r < C; r++) {
        q += xb[(size_t)r * XQp + Tm];
        bad += xb[(size_t)r * XQp + Tm + 1];
      }
      par ^= 1;
      const double den = 1.0 + bc * q;
      if (bad > 0.0 || !(den > 0.0)) {  // not positive definite: the chain stops here (delay.py:1891-1896); cluster-uniform
        if (rank == 0 && tid == 0) {
          d.failed[b] = 1;
          d.n_done[b] = it;
        }
        break;
      }
      if (bc != 0.0) {
        if (warp < nwc) {  // only the warps that hold a sample t < Tb form corr_t and its partial sums
          const double beta = bc / den, sb = sqrt(-beta);
          double scl = 0.0, sccl = 0.0;
          for (int t = tid; t < Tb; t += blockDim.x) {
            double P = 0.0;
            for (int r = 0; r < C; r++) P += xb[(size_t)r * XQp + t];
            const double cv = -beta * P + sb * xi[(it & 1) * Tm + t];
            corr[t] = cv;
            scl += cv;
            sccl += cv * cv;
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            scl += __shfl_xor_sync(0xffffffffu, scl, o);
            sccl += __shfl_xor_sync(0xffffffffu, sccl, o);
          }
          if (lane == 0) {
            scw[warp] = scl;
            scw[32 + warp] = sccl;
          }
        }
        __syncthreads();
      }
      DPROF(3);
      // ---- E + F: moments of the owned column, power-spectrum draw (delay.py:1862-1880), next D
      double ssq = 0.0;
      if (owner) {
        double r0 = 0.0, r1 = 0.0, sc = 0.0, scc = 0.0;
        if (bc != 0.0) {
          for (int w = 0; w < nwc; w++) {
            sc += scw[w];
            scc += scw[32 + w];
          }
          if constexpr (YREG > 0) {
            const double2* cp = reinterpret_cast<const double2*>(corr);
            double r2 = 0.0, r3 = 0.0;
#pragma unroll
            for (int t = 0; t < YREG; t += 4) {
              const double2 c0 = cp[t >> 1], c1 = cp[(t >> 1) + 1];
              r0 = fma(ycol[t], c0.x, r0);
              r1 = fma(ycol[t + 1], c0.y, r1);
              r2 = fma(ycol[t + 2], c1.x, r2);
              r3 = fma(ycol[t + 3], c1.y, r3);
            }
            r0 += r2;
            r1 += r3;
          } else {
            const double* yc = ysm + tid;
            const double2* cp = reinterpret_cast<const double2*>(corr);
            const int ncol2 = 2 * ncol;
            double r2 = 0.0, r3 = 0.0;
            int t = 0;
            for (; t + 7 < Tb; t += 8) {  // 8 rows per trip: all loads issued before the first FMA needs them
              double yv[8];
              double2 cv[4];
#pragma unroll
              for (int j = 0; j < 4; j++) {
                cv[j] = cp[j];
                yv[2 * j] = yc[(size_t)j * ncol2];
                yv[2 * j + 1] = yc[(size_t)j * ncol2 + ncol];
              }
              r0 = fma(yv[0], cv[0].x, r0);
              r1 = fma(yv[1], cv[0].y, r1);
              r2 = fma(yv[2], cv[1].x, r2);
              r3 = fma(yv[3], cv[1].y, r3);
              r0 = fma(yv[4], cv[2].x, r0);
              r1 = fma(yv[5], cv[2].y, r1);
              r2 = fma(yv[6], cv[3].x, r2);
              r3 = fma(yv[7], cv[3].y, r3);
              cp += 4;
              yc += 4 * (size_t)ncol2;
            }
            for (; t + 1 < Tb; t += 2) {
              const double2 c2 = *cp++;
              r0 = fma(yc[0], c2.x, r0);
              r1 = fma(yc[ncol], c2.y, r1);
              yc += ncol2;
            }
            r0 += r2;
            r1 += r3;
            if (t < Tb) r0 = fma(yc[0], corr[t], r0);
          }
        }
        const double sm1 = iv * (sy + sg * sc);                              // sum_t m
        const double sm2 = iv * iv * (syy + 2.0 * sg * (r0 + r1) + scc);     // sum_t m^2
        const double mm = fmax(sm2 - sm1 * sm1 * inv_T, 0.0);             // |m~|^2
        if (Tb > 1) ssq = mm + 2.0 * sqrt(iv * mm) * g_cur + iv * (g_cur * g_cur + c2_cur);
      }
      if (d.cplx) ssq += __shfl_xor_sync(0xffffffffu, ssq, 1);  // re + im columns of one delay (delay.py:1869-1875)
      if (owner) {
        Sk = ssq * ict_cur;
        if (!d.cplx || !(tid & 1)) {
          S[dkey] = Sk;
          d.samples[((size_t)it * d.B + b) * N + dkey] = Sk;
        }
      }
      g_cur = g_nxt;
      c2_cur = c2_nxt;
      ict_cur = ict_nxt;
      if (rank == 0 && tid == 0) d.n_done[b] = it + 1;
      DPROF(4);
      stage_A();
      DPROF(5);
      __syncthreads();
      DPROF(6);
#ifdef DGIBBS_DIAG_PROF
      if (d.prof && blockIdx.x == 0 && tid == 0) d.prof[8] += 1;
#endif
    }
    __syncthreads();  // every reader of ysm is done before the next baseline's copies land
  }
  cluster.sync();  // no CTA exits while a peer may still write its shared memory
}

typedef void (*ResidentKernel)(DevBatch, int, int, int, int, int);
static ResidentKernel resident_kernel_for(const DevBatch& d, const ResidentGeom& g) {
  static const bool no_yreg = getenv("DGIBBS_DIAG_NO_YREG") != nullptr;
  if (g.threads <= 256) return (d.Tmax <= 64 && !no_yreg) ? diag_resident_kernel<256, 64> : diag_resident_kernel<256, 0>;
  if (g.threads <= 384) return diag_resident_kernel<384, 0>;
  if (g.threads <= 512) return diag_resident_kernel<512, 0>;
  return diag_resident_kernel<1024, 0>;
}

// Cluster size: the candidate with the best modelled throughput = concurrent chains / cycles per Gibbs iteration, with
// cycles ~ 500 + 2500 [C > 1: the DSMEM exchange] + c ncol, c = 34 / 38 / 40 cycles per column for the register-column,
// <= 384-thread and larger instantiations (measured at cfg3 geometry: 8 x 256 columns 12.18 ms, 6 x 342 12.01 ms,
// 5 x 410 12.22 ms, 9 x 228 13.78 ms: an SM sustains ~35 cycles per column whatever the split, so what counts is how
// many SMs the cluster size lets the hardware fill: 132 with clusters of 6, 120 with clusters of 8).
// DGIBBS_DIAG_CLUSTER=C forces one.
static cudaError_t launch_resident(const DevBatch& d, int it_begin, int it_end, cudaStream_t st, bool* launched) {
  *launched = false;
  static const char* cap_env = getenv("DGIBBS_DIAG_SMEM_CAP_KB");  // experiment: smaller slices, more CTAs per SM
  static const char* c_env = getenv("DGIBBS_DIAG_CLUSTER");
  const size_t cap = (size_t)(cap_env ? atoi(cap_env) : 227) * 1024;
  ResidentGeom best{};
  int best_ncl = 0;
  double best_score = -1.0;
  for (int C = 1; C <= RMAXC; C++) {
    if (c_env && atoi(c_env) != C) continue;
    if (!c_env && C > 8 && best.C != 0) break;  // non-portable sizes only when no portable one fits
    ResidentGeom g = resident_geom_for(d, cap, C);
    if (g.C == 0) continue;
    ResidentKernel kern = resident_kernel_for(d, g);
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem) != cudaSuccess) {
      (void)cudaGetLastError();
      continue;
    }
    if (C > 8 && cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
      (void)cudaGetLastError();
      continue;
    }
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = C;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.blockDim = dim3(g.threads);
    cfg.dynamicSmemBytes = g.smem;
    cfg.gridDim = dim3(C);
    int ncl = 0;
    if (cudaOccupancyMaxActiveClusters(&ncl, kern, &cfg) != cudaSuccess || ncl <= 0) {
      (void)cudaGetLastError();
      continue;
    }
    if (ncl > d.nlist) ncl = d.nlist;
    const double cpc = (g.threads <= 256 && d.Tmax <= 64) ? 34.3 : (g.threads <= 384 ? 37.8 : 40.0);
    const double score = (double)ncl / (500.0 + (C > 1 ? 2500.0 : 0.0) + cpc * g.ncol);
    if (score > best_score) {
      best = g;
      best_ncl = ncl;
      best_score = score;
    }
  }
  if (best.C == 0) return cudaSuccess;  // not launchable on this device: the caller falls back to the streaming kernel
  ResidentKernel kern = resident_kernel_for(d, best);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)best.smem);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg{};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = best.C;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cfg.blockDim = dim3(best.threads);
  cfg.dynamicSmemBytes = best.smem;
  cfg.stream = st;
  cfg.gridDim = dim3(best_ncl * best.C);
  if (getenv("DGIBBS_VERBOSE"))
    fprintf(stderr, "[dgibbs] resident diagonal sampler: cluster %d x %d threads (%d columns), %d clusters, %zu B smem\n", best.C,
            best.threads, best.ncol, best_ncl, best.smem);
  e = cudaLaunchKernelEx(&cfg, kern, d, it_begin, it_end, best.C, best.ncol, best.XQp);
  if (e == cudaSuccess) *launched = true;
  return e;
}

cudaError_t launch_diag_sampler(const DevBatch& d, int it_begin, int it_end, int sm_count, cudaStream_t st,
                                bool allow_resident) {
  if (d.nlist <= 0) return cudaSuccess;
  if (allow_resident) {
    bool launched = false;
    cudaError_t e = launch_resident(d, it_begin, it_end, st, &launched);
    if (e != cudaSuccess || launched) return e;
  }
  int grid = d.nlist < sm_count ? d.nlist : sm_count;
  if (d.nsys <= 2048)
    diag_sampler_kernel<1, 4><<<grid, DTHREADS, 0, st>>>(d, it_begin, it_end);
  else if (d.nsys <= 4096)
    diag_sampler_kernel<2, 2><<<grid, DTHREADS, 0, st>>>(d, it_begin, it_end);
  else
    return cudaErrorInvalidValue;
  return cudaGetLastError();
}

}  // namespace dg
