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
#include "common.cuh"
#include "internal.h"

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
  const int N = d.N, nsys = d.nsys, npad = d.npad, npairs = nsys >> 1;
  for (int li = blockIdx.x; li < d.nlist; li += gridDim.x) {
    const int b = d.blist[li];
    if (d.failed[b]) continue;
    const int Tb = d.T[b];
    const double a = d.diag_ab[2 * b], bc = d.diag_ab[2 * b + 1];
    const double* yh = d.yhat + (size_t)d.ycompact[b] * d.Tmax * npad;
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
              if (t < Tb) v = *reinterpret_cast<const double2*>(yh + (size_t)t * npad + 2 * p);
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

cudaError_t launch_diag_sampler(const DevBatch& d, int it_begin, int it_end, int sm_count, cudaStream_t st) {
  if (d.nlist <= 0) return cudaSuccess;
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
