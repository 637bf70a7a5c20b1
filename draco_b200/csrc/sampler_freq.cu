// Frequency-space branch of the Gibbs sampler: _draw_signal_sample_t, draco/analysis/delay.py:1821-1860,
// chosen by the reference when few channels survive (M <= N / 4, delay.py:1884-1886).  Real time domain.
//
//   Sh = sqrt(S);  Rt = FTNih * Sh[:, None]  (N x 2M);  R = Rt^T
//   y  = data + w2 - R w1                      forward masked DFT of Sh w1[:, s]  (in-smem FFT, two samples per FFT)
//   Ci = R Rt + I   (2M x 2M)                  never formed by a GEMM: with Shat[m] = sum_t S_t e^{2 pi i m t / N}
//        = A[m] + i B[m] (ONE FFT of S per iteration) the entries are
//        Ci[(f,0),(f',0)] = h h' (A[f-f'] + A[f+f']) / 2      Ci[(f,1),(f',1)] = h h' (A[f-f'] - A[f+f']) / 2
//        Ci[(f,0),(f',1)] = h h' (B[f-f'] - B[f+f']) / 2      Ci[(f,1),(f',0)] = -h h' (B[f+f'] + B[f-f']) / 2
//        (+ identity), generated tile by tile inside the factorisation: the reference's 2 (2M)^2 N flop per
//        iteration for R @ Rt become O(N log N).
//   x  = Ci^-1 y                               fp64 DMMA Cholesky of dense_core.cuh on the 2M system
//   d  = Sh (Rt x + w1) = S idft(h x) + Sh w1  inverse masked DFT (in-smem FFT, two samples per FFT)
//   S' = var_s(d) T / chi2                     moments accumulated on the fly: d is never stored
// Only the 2M-sized system touches HBM: (2M)^3 / 3 flop instead of N^3 / 3 (>= 8x fewer).
#include "dense_core.cuh"
#include "tma_host.h"
#include "tma_pipe.cuh"

namespace dg {

constexpr int FMAXJ = 16;  // delays per thread: N <= THREADS * FMAXJ = 4096

// tile generator of Ci = R Rt + I on the compact (retained) channel set
struct FreqGen {
  const double2* Shat;     // [N]  A + iB
  const double* h;         // [Kpad] window * sqrt(Ni_r) on the full grid
  const long long* fsel;   // [M] channel of compact index
  int N, n2m;              // n2m = 2M
  __device__ __forceinline__ void operator()(double* sT, int ti, int tj, bool lower_only) const {
    const int R0 = ti * TILE, C0 = tj * TILE;
    for (int e = threadIdx.x; e < TILE * TILE; e += THREADS) {
      const int r = e >> 7, c = e & 127;
      if (lower_only && c > r) continue;
      const int R = R0 + r, C = C0 + c;
      double v;
      if (R < n2m && C < n2m) {
        const int f = (int)fsel[R >> 1], fp = (int)fsel[C >> 1], p = R & 1, pp = C & 1;
        int md = f - fp;
        if (md < 0) md += N;
        int ms = f + fp;
        if (ms >= N) ms -= N;
        const double2 sd = Shat[md], ss = Shat[ms];
        double g;
        if (p == 0 && pp == 0) g = sd.x + ss.x;
        else if (p == 1 && pp == 1) g = sd.x - ss.x;
        else if (p == 0) g = sd.y - ss.y;
        else g = -(ss.y + sd.y);
        v = 0.5 * g * h[2 * f + p] * h[2 * fp + pp];
        if (R == C) v += 1.0;
      } else {
        v = (R == C) ? 1.0 : 0.0;
      }
      sT[r * LDW + c] += v;
    }
  }
};

// Shared memory: [sWp: WP_DOUBLES][pipe: PIPE_DOUBLES][misc]
__global__ void __launch_bounds__(THREADS, 1)
    freq_sampler_kernel(DevBatch d, int it_begin, int it_end, int use_tma, const __grid_constant__ CUtensorMap tmap) {
  extern __shared__ __align__(16) double smem[];
  const Lane ln;
  const int slot = blockIdx.x;
  const int N = d.N, Kpad = d.Kpad;
  double* sWp = smem;
  double* pipe = sWp + WP_DOUBLES;
  double* misc = pipe + PIPE_DOUBLES;
  double2* bufA = reinterpret_cast<double2*>(pipe);
  double2* bufB = bufA + N;
  double2* tw_s = reinterpret_cast<double2*>(sWp);
  // operand ring of the factorisation's NT updates (tma_pipe.cuh); the host passes a tensor map only when every chain
  // of the list has the same padded system size (the map's row pitch)
  TmaRing ring;
  if (use_tma) tma_ring_init(ring, reinterpret_cast<unsigned long long*>(misc + MISC_DOUBLES));
  NtTma nt_tma{&tmap, &ring, pipe, slot};

  const size_t wrows_full = (size_t)d.npad + (size_t)d.RT * TILE;
  double* Wk = d.work + (size_t)slot * d.work_stride;  // the slot's delay-space workspace is large enough
  double* Wpg = d.Wp + (size_t)slot * d.nblk * WP_DOUBLES;
  double* Sig = d.Si + (size_t)slot * d.npad;               // per-slot scratch: Sh[N]
  double2* Shat = reinterpret_cast<double2*>(d.uT + (size_t)slot * d.RT * TILE * Kpad);  // [N] (uT is unused here)

  for (int li = slot; li < d.nlist; li += gridDim.x) {
    const int b = d.blist[li];
    if (d.failed[b]) continue;
    const int Tb = d.T[b], Mb = d.M[b], n2m = 2 * Mb;
    const double* h = d.h + (size_t)b * Kpad;
    const double* zT = d.zT + (size_t)b * d.Tmax * Kpad;
    const long long* fsel = d.fsel + d.chan_off[b];
    double* S = d.S + (size_t)b * N;
    const Philox ph(mix_key(d.seed, d.gid ? (unsigned long long)d.gid[b] : (unsigned long long)b));

    CholWS ws;
    ws.npad = (n2m + TILE - 1) / TILE * TILE;
    ws.nblk = ws.npad / TILE;
    ws.RT = d.RT;
    ws.rhs_frags = (Tb + 7) >> 3;
    ws.Wk = Wk;
    ws.Rh = Wk + (size_t)ws.npad * ws.npad;
    ws.Wpg = Wpg;
    ws.sWp = sWp;
    ws.pipe = pipe;
    ws.misc = misc;
    double* Rh = ws.Rh;
    const int npt = ws.npad;
    double* V1 = Wk + wrows_full * d.npad;         // [Tb][N]  Sh w1 of the current iteration (tail of the slot)
    FreqGen gen{Shat, h, fsel, N, n2m};
    // RHS rows / columns no phase writes: zero once per baseline (padding rows stay zero through the solves)
    for (long long e = threadIdx.x; e < (long long)d.RT * TILE * (npt / 2); e += THREADS)
      reinterpret_cast<double2*>(Rh)[e] = make_double2(0.0, 0.0);

    bool alive = true;
    for (int it = it_begin; it < it_end && alive; it++) {
      const double* w1 = (d.rng_mode == 0) ? d.w1 + d.w1_off[b] + (long long)it * N * Tb : nullptr;
      const double* w2 = (d.rng_mode == 0) ? d.w2 + d.w2_off[b] + (long long)it * 2 * Mb * Tb : nullptr;
      // w1[2p, s], w1[2p+1, s] for the sample pair (s0, s0+1): one Philox block, fp32 Box-Muller (native mode)
      auto w1_quad_at = [&](int p, int s0, bool two, double& a0, double& a1, double& b0, double& b1) {
        a0 = a1 = b0 = b1 = 0.0;
        if (d.rng_mode == 0) {
          a0 = w1[(long long)(2 * p) * Tb + s0];
          a1 = w1[(long long)(2 * p + 1) * Tb + s0];
          if (two) {
            b0 = w1[(long long)(2 * p) * Tb + s0 + 1];
            b1 = w1[(long long)(2 * p + 1) * Tb + s0 + 1];
          }
        } else if (d.rng_mode == 1) {
          const unsigned long long blk = (unsigned long long)(s0 >> 1) * (N >> 1) + p;
          normal_quad(ph, (unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_W1, STREAM_W1_ODD, true, two, a0, a1, b0, b1);
        }
      };
      __syncthreads();
      // ---- Sh = sqrt(S), Shat = FFT(S)   (delay.py:1825, 1851)
      for (int m = threadIdx.x; m < N / 2; m += THREADS) tw_s[m] = d.twtab[m];
      for (int t = threadIdx.x; t < N; t += THREADS) {
        const double s = S[t];
        Sig[t] = sqrt(s);
        bufA[t] = make_double2(s, 0.0);
      }
      __syncthreads();
      {
        const double2* res = fft_inverse_smem(bufA, bufB, tw_s, N);
        for (int m = threadIdx.x; m < N; m += THREADS) Shat[m] = res[m];
      }
      __syncthreads();
      // ---- y = data + w2 - R w1   (delay.py:1846-1849), two samples per complex FFT
      for (int s0 = 0; s0 < Tb; s0 += 2) {
        const bool two = s0 + 1 < Tb;
        for (int p = threadIdx.x; p < N / 2; p += THREADS) {
          const double sh0 = Sig[2 * p], sh1 = Sig[2 * p + 1];
          double a0, a1, b0, b1;
          w1_quad_at(p, s0, two, a0, a1, b0, b1);
          a0 *= sh0; a1 *= sh1; b0 *= sh0; b1 *= sh1;
          if (!two) b0 = b1 = 0.0;
          bufA[2 * p] = make_double2(a0, b0);
          bufA[2 * p + 1] = make_double2(a1, b1);
          // Sh w1 is needed again for d = S idft(h x) + Sh w1: keep it (sample-major) instead of redrawing
          *reinterpret_cast<double2*>(V1 + (size_t)s0 * N + 2 * p) = make_double2(a0, a1);
          if (two) *reinterpret_cast<double2*>(V1 + (size_t)(s0 + 1) * N + 2 * p) = make_double2(b0, b1);
        }
        __syncthreads();
        const double2* P = fft_inverse_smem(bufA, bufB, tw_s, N);
        // P[f] = Va[f] + i Vb[f],  V[f] = sum_t v_t e^{+2 pi i f t / N}:  Va = (P[f] + conj P[N-f]) / 2,
        // Vb = (P[f] - conj P[N-f]) / (2i)
        for (int ci = threadIdx.x; ci < Mb; ci += THREADS) {
          const int f = (int)fsel[ci];
          const double2 pf = P[f], pn = P[(N - f) & (N - 1)];
          const double var_ = 0.5 * (pf.x + pn.x), vai = 0.5 * (pf.y - pn.y);
          const double vbr = 0.5 * (pf.y + pn.y), vbi = -0.5 * (pf.x - pn.x);
          const double hre = h[2 * f], him = h[2 * f + 1];
          double e[4] = {0.0, 0.0, 0.0, 0.0};  // w2 of (s0: re, im), (s0+1: re, im)
          if (d.rng_mode == 1) {
            const unsigned long long blk = (unsigned long long)(s0 >> 1) * (Kpad >> 1) + f;
            normal_quad(ph, (unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_W2, STREAM_W2_ODD, true, s0 + 1 < Tb, e[0], e[1],
                        e[2], e[3]);
          }
#pragma unroll
          for (int k = 0; k < 2; k++) {
            const int s = s0 + k;
            if (s >= Tb) break;
            double n0 = e[2 * k], n1 = e[2 * k + 1];
            if (d.rng_mode == 0) {
              n0 = w2[(long long)(2 * ci) * Tb + s];
              n1 = w2[(long long)(2 * ci + 1) * Tb + s];
            }
            const double2 z = *reinterpret_cast<const double2*>(zT + (size_t)s * Kpad + 2 * f);
            const double vr = k ? vbr : var_, vi = k ? vbi : vai;
            // R w1: row (f,0) = h_re Re V, row (f,1) = h_im (-Im V)   (F rows cos, -sin; delay.py:1507-1508)
            *reinterpret_cast<double2*>(Rh + (size_t)s * npt + 2 * ci) = make_double2(z.x + n0 - hre * vr, z.y + n1 + him * vi);
          }
        }
        __syncthreads();
      }
      // ---- x = Ci^-1 y   (delay.py:1851-1858)
      alive = use_tma ? chol_factor_forward(ws, ln, gen, nt_tma) : chol_factor_forward(ws, ln, gen);
      if (!alive) {
        __syncthreads();
        if (threadIdx.x == 0) {
          d.failed[b] = 1;
          d.n_done[b] = it;
        }
        break;
      }
      chol_backward(ws, ln);
      __syncthreads();
      // ---- d = S idft(h x) + Sh w1  (delay.py:1860) and its moments over the samples, two samples per FFT
      double shift[FMAXJ], sum[FMAXJ], sq[FMAXJ];
#pragma unroll
      for (int j = 0; j < FMAXJ; j++) shift[j] = sum[j] = sq[j] = 0.0;
      for (int m = threadIdx.x; m < N / 2; m += THREADS) tw_s[m] = d.twtab[m];  // sWp was used by the solves
      for (int s0 = 0; s0 < Tb; s0 += 2) {
        const bool two = s0 + 1 < Tb;
        for (int k = threadIdx.x; k < N; k += THREADS) bufA[k] = make_double2(0.0, 0.0);
        __syncthreads();
        for (int ci = threadIdx.x; ci < Mb; ci += THREADS) {
          const int f = (int)fsel[ci];
          const double2 xa = *reinterpret_cast<const double2*>(Rh + (size_t)s0 * npt + 2 * ci);
          double2 xb = make_double2(0.0, 0.0);
          if (two) xb = *reinterpret_cast<const double2*>(Rh + (size_t)(s0 + 1) * npt + 2 * ci);
          const double2 U0 = make_double2(h[2 * f] * xa.x, h[2 * f + 1] * xa.y);
          const double2 U1 = make_double2(h[2 * f] * xb.x, h[2 * f + 1] * xb.y);
          if (f == 0 || f == N / 2) {
            bufA[f] = make_double2(U0.x, U1.x);
          } else {  // Hermitian extension of both half spectra, combined as A + iB (see project_rows_fft)
            bufA[f] = make_double2(0.5 * (U0.x - U1.y), 0.5 * (U0.y + U1.x));
            bufA[N - f] = make_double2(0.5 * (U0.x + U1.y), 0.5 * (U1.x - U0.y));
          }
        }
        __syncthreads();
        const double2* res = fft_inverse_smem(bufA, bufB, tw_s, N);
#pragma unroll
        for (int j = 0; j < FMAXJ; j++) {
          const int t = threadIdx.x + j * THREADS;
          if (t < N) {
            const double st = S[t];
            const double da = st * res[t].x + V1[(size_t)s0 * N + t];
            if (s0 == 0) shift[j] = da;
            const double ea = da - shift[j];
            sum[j] += ea;
            sq[j] += ea * ea;
            if (two) {
              const double eb = st * res[t].y + V1[(size_t)(s0 + 1) * N + t] - shift[j];
              sum[j] += eb;
              sq[j] += eb * eb;
            }
          }
        }
        __syncthreads();
      }
      // ---- power-spectrum draw (delay.py:1862-1880)
#pragma unroll
      for (int j = 0; j < FMAXJ; j++) {
        const int t = threadIdx.x + j * THREADS;
        if (t < N) {
          const double var = (sq[j] - sum[j] * sum[j] / (double)Tb) / (double)Tb;
          double chi2;
          if (d.rng_mode == 1) {
            double acc2 = 0.0;
            const unsigned long long base = (unsigned long long)t * ((Tb + 3) / 4 + 1);
            for (int p = 0; p < Tb / 2; p += 2) {
              unsigned long long blk = base + (p >> 1);
              uint4 r = ph((unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_CHI2);
              acc2 += -2.0 * log(u01(r.x, r.y));
              if (p + 1 < Tb / 2) acc2 += -2.0 * log(u01(r.z, r.w));
            }
            if (Tb & 1) {
              unsigned long long blk = base + (Tb + 3) / 4;
              double n0, n1;
              normal_pair(ph((unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_CHI2), n0, n1);
              acc2 += n0 * n0;
            }
            chi2 = acc2;
          } else {
            chi2 = d.chi2[((size_t)b * d.niter + it) * N + t];
          }
          const double snew = var * (double)Tb / chi2;
          S[t] = snew;
          d.samples[((size_t)it * d.B + b) * N + t] = snew;
        }
      }
      if (threadIdx.x == 0) d.n_done[b] = it + 1;
      __syncthreads();
    }
  }
}

// The frequency-space kernel borrows the slot workspace of the delay-space kernel (L and the RHS of the 2M
// system; the N x T buffer of Sh w1 is the slot's tail); N must be a power of two (FFT) within one CTA's reach.
bool freq_branch_fits(const DevBatch& d, int M, int T) {
  if (d.cplx || !d.use_fft || d.N > THREADS * FMAXJ) return false;
  (void)T;
  const size_t npt = ((size_t)2 * M + TILE - 1) / TILE * TILE;
  return npt <= (size_t)d.npad;  // L and the RHS of the 2M system fit the slot; Sh w1 has its own tail
}

cudaError_t launch_freq_sampler(const DevBatch& d, int it_begin, int it_end, cudaStream_t st, int uniform_npt) {
  if (d.nlist <= 0) return cudaSuccess;
  const size_t smem = sizeof(double) * (WP_DOUBLES + PIPE_DOUBLES + MISC_DOUBLES + 2 * TNST);
  cudaError_t e = cudaFuncSetAttribute(freq_sampler_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int grid = d.nlist < d.nslots ? d.nlist : d.nslots;
  CUtensorMap tm;
  int use_tma = 0;
  if (uniform_npt > 0)
    use_tma = tma_work_map(&tm, d.work, (size_t)uniform_npt + (size_t)d.RT * TILE, uniform_npt, d.work_stride, d.nslots,
                           sizeof(double) * WP_DOUBLES, TILE, TKC);
  else
    memset(&tm, 0, sizeof tm);
  freq_sampler_kernel<<<grid, THREADS, smem, st>>>(d, it_begin, it_end, use_tma, tm);
  return cudaGetLastError();
}

}  // namespace dg
