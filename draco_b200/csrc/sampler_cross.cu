// Cross-power-spectrum Gibbs sampler: delay_spectrum_gibbs_cross, draco/analysis/delay.py:1907-2099.
//
// nd datasets observed on the same channels share one signal covariance S_t (nd x nd, Hermitian) per
// delay t.  Per iteration the reference solves the complex Hermitian system of size nd N
//     Ci = blockdiag_d(F_d^H N_d^-1 F_d) + (delay-diagonal blocks S_t^-1)           delay.py:2021-2027
// by Cholesky (delay.py:2047-2058) and then draws every S_t from a complex inverse Wishart
// (delay.py:2065-2083).  Here:
//   * each F_d^H N_d^-1 F_d is Hermitian circulant, c_d[k] = sum_f h_df^2 e^{-2 pi i f k / N}: only the
//     first columns are kept and the tiles of Ci are generated on the fly;
//   * the complex system runs through the real fp64 DMMA Cholesky of dense_core.cuh in its real
//     embedding [[Re, -Im], [Im, Re]] (row 2 q + p for complex row q): 8 n^3 / 3 real flop, exactly the
//     cost of a complex factorisation, with the real/imaginary parts of a row adjacent so that a failing
//     leading minor is met at the same complex row as LAPACK's zpotrf;
//   * y = F^T h (z + w2) (delay.py:2033; the reference applies F^T, not F^H) is a forward DFT of the
//     zero-filled channel grid: the in-smem FFT of dense_core.cuh on conjugated input;
//   * the nd x nd work per delay (inverse, Cholesky, Bartlett factor, inverse Wishart) is one thread per
//     delay, in registers/local memory (nd <= 4).
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/dgibbs.h"
#include "dense_core.cuh"
#include "tma_host.h"
#include "tma_pipe.cuh"
#include "window.cuh"

namespace dg {

constexpr int NDMAX = 4;
enum { STREAM_X_W1 = 16, STREAM_X_W2 = 17, STREAM_X_WN = 18, STREAM_X_WG = 19 };

struct CrossBatch {
  int B, nd, N, T, M, niter, window_id, rng_mode;
  int nc, nsys, npad, nblk, RT, L, use_fft, nslots;
  unsigned long long seed;
  const double2* data;   // [B][nd][T][M]
  const double* Ni;      // [B][nd][M]
  const long long* fsel; // [M]
  const double2* S0;     // [B][nd][nd][N]
  const long long* gid;  // [B] or null
  const double2* w1;     // [B][niter][N][nd][T]
  const double2* w2;     // [B][niter][nd][M][T]
  const double* wn;      // [B][niter][N][2][L]
  const double* wg;      // [B][niter][N][nd]
  double* h;             // [B][nd][M]   window * sqrt(Ni)
  double2* z;            // [B][nd][T][M] windowed, whitened data (sample-major per dataset)
  double2* ctab;         // [B][nd][N]   first columns of the circulant blocks
  double2* S;            // [B][nd][nd][N] chain state
  double2* samples;      // [B][niter][nd][nd][N]
  int* n_done;
  int* failed;
  const double2* twtab;  // [N/2] twiddles when N is a power of two
  double* work;          // [nslots][(npad + RT*TILE)][npad]
  double* Wp;            // [nslots][nblk][WP_DOUBLES]
  double2* Si;           // [nslots][N][nd][nd]  S_t^-1
  double2* U;            // [nslots][N][nd][nd]  upper Cholesky factor of S_t (S = U^H U)
};

// ----------------------------------------------------------------------------- nd x nd complex helpers
struct CM {
  double2 a[NDMAX][NDMAX];
};
// cadd / csub come from dense_core.cuh
__device__ __forceinline__ double2 cmulc(double2 a, double2 b) {  // a * conj(b)
  return make_double2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}
__device__ __forceinline__ double2 cscale(double2 a, double s) { return make_double2(a.x * s, a.y * s); }

// Lower Cholesky factor of a Hermitian matrix (lower triangle read): A = L L^H.  False if not PD.
__device__ bool cm_chol_lower(const CM& A, int n, CM& Lm) {
  bool ok = true;
  for (int j = 0; j < n; j++) {
    double dj = A.a[j][j].x;
    for (int k = 0; k < j; k++) dj -= Lm.a[j][k].x * Lm.a[j][k].x + Lm.a[j][k].y * Lm.a[j][k].y;
    if (!(dj > 0.0)) ok = false;
    const double piv = sqrt(dj), r = 1.0 / piv;
    Lm.a[j][j] = make_double2(piv, 0.0);
    for (int i = j + 1; i < n; i++) {
      double2 v = A.a[i][j];
      for (int k = 0; k < j; k++) v = csub(v, cmulc(Lm.a[i][k], Lm.a[j][k]));
      Lm.a[i][j] = cscale(v, r);
    }
    for (int i = 0; i < j; i++) Lm.a[i][j] = make_double2(0.0, 0.0);
  }
  return ok;
}
// Inverse of a lower triangular matrix (real positive diagonal not required).
__device__ void cm_inv_lower(const CM& Lm, int n, CM& W) {
  for (int j = 0; j < n; j++) {
    for (int i = 0; i < j; i++) W.a[i][j] = make_double2(0.0, 0.0);
    const double2 dj = Lm.a[j][j];
    const double den = dj.x * dj.x + dj.y * dj.y;
    W.a[j][j] = make_double2(dj.x / den, -dj.y / den);
    for (int i = j + 1; i < n; i++) {
      double2 s = make_double2(0.0, 0.0);
      for (int k = j; k < i; k++) s = cadd(s, cmul(Lm.a[i][k], W.a[k][j]));
      const double2 di = Lm.a[i][i];
      const double dn = di.x * di.x + di.y * di.y;
      const double2 inv = make_double2(di.x / dn, -di.y / dn);
      W.a[i][j] = cmul(make_double2(-s.x, -s.y), inv);
    }
  }
}
// R = W^H W  (W lower triangular): the inverse of L L^H when W = L^-1
__device__ void cm_whw(const CM& W, int n, CM& R) {
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      double2 s = make_double2(0.0, 0.0);
      for (int k = (i > j ? i : j); k < n; k++) s = cadd(s, cmulc(W.a[k][j], W.a[k][i]));  // conj(W[k][i]) W[k][j]
      R.a[i][j] = s;
    }
}

// gamma(shape, 1) variate, shape >= 1 (Marsaglia-Tsang, fp64)
__device__ double gamma_draw(const Philox& ph, unsigned it, unsigned long long key, unsigned stream, double shape) {
  const double dd = shape - 1.0 / 3.0, cc = 1.0 / sqrt(9.0 * dd);
  for (unsigned attempt = 0; attempt < 64; attempt++) {
    uint4 r = ph((unsigned)key, (unsigned)(key >> 32) | (attempt << 24), it, stream);
    const double u1 = u01(r.x, r.y);
    double s, c;
    sincospi(2.0 * ((double)r.z * 2.3283064365386963e-10), &s, &c);
    const double x = sqrt(-2.0 * log(u1)) * c;
    double v = 1.0 + cc * x;
    if (v <= 0.0) continue;
    v = v * v * v;
    const double u = ((double)r.w + 0.5) * 2.3283064365386963e-10;
    const double x2 = x * x;
    if (u < 1.0 - 0.0331 * x2 * x2 || log(u) < 0.5 * x2 + dd * (1.0 - v + log(v))) return dd * v;
  }
  return dd;
}
// standard complex normal (unit total variance) number `idx` of a stream
__device__ __forceinline__ double2 cnormal(const Philox& ph, unsigned it, unsigned stream, unsigned long long idx) {
  double n0, n1;
  normal_pair(ph((unsigned)idx, (unsigned)(idx >> 32), it, stream), n0, n1);
  return make_double2(n0 * 0.70710678118654752440, n1 * 0.70710678118654752440);
}

// ----------------------------------------------------------------------------- per-problem setup
// window (x = fsel / N, delay.py:1982), whitening, first columns of the circulant blocks, initial state
__global__ void cross_setup_kernel(CrossBatch d) {
  const int b = blockIdx.x;
  const int nd = d.nd, N = d.N, T = d.T, M = d.M;
  for (int e = threadIdx.x; e < nd * M; e += blockDim.x) {
    const int f = e % M;
    double w = 1.0;
    if (d.window_id >= 0) {
      const double x = (double)d.fsel[f] * 1.0 / (double)N;
      w = window_value(d.window_id, x);
    }
    d.h[(size_t)b * nd * M + e] = w * sqrt(d.Ni[(size_t)b * nd * M + e]);
  }
  // z[d][s][f] = (data[d][s][f] * w_f) * sqrt(Ni[d][f])   (delay.py:1987, 2001)
  for (long long e = threadIdx.x; e < (long long)nd * T * M; e += blockDim.x) {
    const int f = (int)(e % M);
    const int dd = (int)(e / ((long long)T * M));
    double w = 1.0;
    if (d.window_id >= 0) w = window_value(d.window_id, (double)d.fsel[f] * 1.0 / (double)N);
    const double sq = sqrt(d.Ni[((size_t)b * nd + dd) * M + f]);
    const double2 v = d.data[(size_t)b * nd * T * M + e];
    d.z[(size_t)b * nd * T * M + e] = make_double2((v.x * w) * sq, (v.y * w) * sq);
  }
  __syncthreads();
  // c_d[k] = sum_f h_df^2 e^{-2 pi i f k / N}
  for (int e = threadIdx.x; e < nd * N; e += blockDim.x) {
    const int dd = e / N, k = e - dd * N;
    const double* h = d.h + ((size_t)b * nd + dd) * M;
    double cr = 0.0, ci = 0.0;
    for (int f = 0; f < M; f++) {
      const long long m = (d.fsel[f] * (long long)k) % N;
      double s, c;
      sincospi(2.0 * (double)m / (double)N, &s, &c);
      const double g = h[f] * h[f];
      cr += g * c;
      ci -= g * s;
    }
    d.ctab[(size_t)b * nd * N + e] = make_double2(cr, ci);
  }
  for (long long e = threadIdx.x; e < (long long)nd * nd * N; e += blockDim.x)
    d.S[(size_t)b * nd * nd * N + e] = d.S0[(size_t)b * nd * nd * N + e];
  if (threadIdx.x == 0) {
    d.n_done[b] = 0;
    d.failed[b] = 0;
  }
}

// ----------------------------------------------------------------------------- tile generator
// Adds tile (ti, tj) of the real embedding of Ci: complex row q = dataset * N + delay, real row 2 q + p.
struct CrossGen {
  const double2* ct;  // [nd][N]
  const double2* Si;  // [N][nd][nd]
  int N, nd, nsys;
  __device__ __forceinline__ void operator()(double* sT, int ti, int tj, bool lower_only) const {
    const int R0 = ti * TILE, C0 = tj * TILE;
    for (int e = threadIdx.x; e < TILE * TILE; e += THREADS) {
      const int r = e >> 7, c = e & 127;
      if (lower_only && c > r) continue;
      const int R = R0 + r, C = C0 + c;
      double v;
      if (R < nsys && C < nsys) {
        const int qr = R >> 1, pr = R & 1, qc = C >> 1, pc = C & 1;
        const int dr = qr / N, tr = qr - dr * N, dc = qc / N, tc = qc - dc * N;
        double2 zv = make_double2(0.0, 0.0);
        if (dr == dc) {
          int k = tr - tc;
          if (k < 0) k += N;
          zv = ct[dr * N + k];
        }
        if (tr == tc) zv = cadd(zv, Si[((size_t)tr * nd + dr) * nd + dc]);
        v = (pr == pc) ? zv.x : (pr ? zv.y : -zv.y);
      } else {
        v = (R == C) ? 1.0 : 0.0;
      }
      sT[r * LDW + c] += v;
    }
  }
};

// ----------------------------------------------------------------------------- the sampler
// Shared memory: [sWp: WP_DOUBLES][pipe: PIPE_DOUBLES][misc]
__global__ void __launch_bounds__(THREADS, 1)
    cross_sampler_kernel(CrossBatch d, int use_tma, const __grid_constant__ CUtensorMap tmap) {
  extern __shared__ __align__(16) double smem[];
  const Lane ln;
  const int slot = blockIdx.x;
  const int nd = d.nd, N = d.N, T = d.T, M = d.M, nsys = d.nsys, npad = d.npad, L = d.L;
  double* sWp = smem;
  double* pipe = sWp + WP_DOUBLES;
  double* misc = pipe + PIPE_DOUBLES;
  int* flag = reinterpret_cast<int*>(misc + 8);
  // operand ring of the factorisation's NT updates (tma_pipe.cuh); cp.async loaders when the host found no tensor map
  TmaRing ring;
  if (use_tma) tma_ring_init(ring, reinterpret_cast<unsigned long long*>(misc + MISC_DOUBLES));
  NtTma nt_tma{&tmap, &ring, pipe, slot};

  CholWS ws;
  const size_t wrows = (size_t)npad + (size_t)d.RT * TILE;
  ws.Wk = d.work + (size_t)slot * wrows * npad;
  ws.Rh = ws.Wk + (size_t)npad * npad;
  ws.Wpg = d.Wp + (size_t)slot * d.nblk * WP_DOUBLES;
  ws.sWp = sWp;
  ws.pipe = pipe;
  ws.misc = misc;
  ws.npad = npad;
  ws.nblk = d.nblk;
  ws.RT = d.RT;
  ws.rhs_frags = (T + 7) >> 3;
  double* Rh = ws.Rh;
  double2* Si = d.Si + (size_t)slot * N * nd * nd;
  double2* U = d.U + (size_t)slot * N * nd * nd;
  const int rows = ws.rhs_frags * 8;

  for (int b = slot; b < d.B; b += gridDim.x) {
    const Philox ph(mix_key(d.seed, d.gid ? (unsigned long long)d.gid[b] : (unsigned long long)b));
    double2* S = d.S + (size_t)b * nd * nd * N;
    const double* h = d.h + (size_t)b * nd * M;
    const double2* z = d.z + (size_t)b * nd * T * M;
    CrossGen gen{d.ctab + (size_t)b * nd * N, Si, N, nd, nsys};
    // RHS rows no phase ever writes: keep them finite
    for (long long e = (long long)rows * (npad / 2) + threadIdx.x; e < (long long)d.RT * TILE * (npad / 2); e += THREADS)
      reinterpret_cast<double2*>(Rh)[e] = make_double2(0.0, 0.0);

    for (int it = 0; it < d.niter; it++) {
      const size_t bi = (size_t)b * d.niter + it;
      if (threadIdx.x == 0) *flag = 0;
      __syncthreads();
      // ---- 1. per delay: S_t^-1 and the upper Cholesky factor of S_t (delay.py:2016-2019)
      for (int t = threadIdx.x; t < N; t += THREADS) {
        CM A, Lm, W, R;
        for (int i = 0; i < nd; i++)
          for (int j = 0; j < nd; j++) A.a[i][j] = S[((size_t)i * nd + j) * N + t];
        if (!cm_chol_lower(A, nd, Lm)) atomicOr(flag, 1);
        cm_inv_lower(Lm, nd, W);
        cm_whw(W, nd, R);
        for (int i = 0; i < nd; i++)
          for (int j = 0; j < nd; j++) {
            Si[((size_t)t * nd + i) * nd + j] = R.a[i][j];
            U[((size_t)t * nd + i) * nd + j] = make_double2(Lm.a[j][i].x, -Lm.a[j][i].y);  // U = L^H
          }
      }
      __syncthreads();
      if (*flag) break;
      // ---- 2a. Rh = U_t^-1 w1[t]  (delay.py:2035-2042), zero in the padding
      for (long long e = threadIdx.x; e < (long long)rows * (npad / 2); e += THREADS) {
        const int s = (int)(e / (npad / 2)), q = (int)(e % (npad / 2));
        if (s >= T || q >= d.nc) Rh[(size_t)s * npad + 2 * q] = Rh[(size_t)s * npad + 2 * q + 1] = 0.0;
      }
      for (long long e = threadIdx.x; e < (long long)T * N; e += THREADS) {
        const int s = (int)(e / N), t = (int)(e % N);
        double2 w[NDMAX], x[NDMAX];
        for (int i = 0; i < nd; i++) {
          if (d.rng_mode == 0) w[i] = d.w1[((bi * N + t) * nd + i) * T + s];
          else if (d.rng_mode == 1) w[i] = cnormal(ph, (unsigned)it, STREAM_X_W1, ((unsigned long long)t * nd + i) * T + s);
          else w[i] = make_double2(0.0, 0.0);
        }
        const double2* Ut = U + (size_t)t * nd * nd;
        for (int i = nd - 1; i >= 0; i--) {  // back substitution, U upper with real positive diagonal
          double2 v = w[i];
          for (int j = i + 1; j < nd; j++) v = csub(v, cmul(Ut[i * nd + j], x[j]));
          x[i] = cscale(v, 1.0 / Ut[i * nd + i].x);
        }
        for (int i = 0; i < nd; i++)
          *reinterpret_cast<double2*>(Rh + (size_t)s * npad + 2 * ((size_t)i * N + t)) = x[i];
      }
      __syncthreads();
      // ---- 2b. Rh += F^T h (z + w2): forward DFT of the zero-filled channel grid per dataset and sample
      for (int dd = 0; dd < nd; dd++)
        for (int s = 0; s < T; s++) {
          const double* hd = h + (size_t)dd * M;
          const double2* zr = z + ((size_t)dd * T + s) * M;
          double2* yrow = reinterpret_cast<double2*>(Rh + (size_t)s * npad) + (size_t)dd * N;
          auto uval = [&](int f) {
            double2 n = make_double2(0.0, 0.0);
            if (d.rng_mode == 0) n = d.w2[((bi * nd + dd) * M + f) * T + s];
            else if (d.rng_mode == 1) n = cnormal(ph, (unsigned)it, STREAM_X_W2, ((unsigned long long)dd * M + f) * T + s);
            const double2 zv = zr[f];
            return make_double2(hd[f] * (zv.x + n.x), hd[f] * (zv.y + n.y));
          };
          if (d.use_fft) {
            double2* bufA = reinterpret_cast<double2*>(pipe);
            double2* bufB = bufA + N;
            double2* tw_s = reinterpret_cast<double2*>(sWp);
            if (dd == 0 && s == 0)
              for (int m = threadIdx.x; m < N / 2; m += THREADS) tw_s[m] = d.twtab[m];
            for (int k = threadIdx.x; k < N; k += THREADS) bufA[k] = make_double2(0.0, 0.0);
            __syncthreads();
            for (int f = threadIdx.x; f < M; f += THREADS) {
              const double2 u = uval(f);
              bufA[d.fsel[f]] = make_double2(u.x, -u.y);  // conj: forward DFT through the inverse FFT
            }
            __syncthreads();
            const double2* res = fft_inverse_smem(bufA, bufB, tw_s, N);
            for (int t = threadIdx.x; t < N; t += THREADS) {
              double2 y = yrow[t];
              y.x += res[t].x;
              y.y -= res[t].y;
              yrow[t] = y;
            }
            __syncthreads();
          } else {  // any N: direct sum (slow path, small problems only)
            for (int t = threadIdx.x; t < N; t += THREADS) {
              double2 acc = make_double2(0.0, 0.0);
              for (int f = 0; f < M; f++) {
                const long long m = (d.fsel[f] * (long long)t) % N;
                double sn, cs;
                sincospi(2.0 * (double)m / (double)N, &sn, &cs);
                acc = cadd(acc, cmul(uval(f), make_double2(cs, -sn)));
              }
              yrow[t] = cadd(yrow[t], acc);
            }
          }
        }
      __syncthreads();
      // ---- 3. Cholesky solve of the real embedding (delay.py:2047-2058)
      const bool alive = use_tma ? chol_factor_forward(ws, ln, gen, nt_tma) : chol_factor_forward(ws, ln, gen);
      if (!alive) {
        __syncthreads();
        if (threadIdx.x == 0) *flag = 1;
        __syncthreads();
        break;
      }
      chol_backward(ws, ln);
      __syncthreads();
      // ---- 4. per delay: sample covariance over the samples, inverse Wishart draw (delay.py:2065-2083)
      for (int t = threadIdx.x; t < N; t += THREADS) {
        double2 mean[NDMAX];
        CM Cv, Lm, W, Q, G;
        for (int i = 0; i < nd; i++) {
          double2 m = make_double2(0.0, 0.0);
          for (int s = 0; s < T; s++) m = cadd(m, *reinterpret_cast<const double2*>(Rh + (size_t)s * npad + 2 * ((size_t)i * N + t)));
          mean[i] = cscale(m, 1.0 / (double)T);
        }
        for (int i = 0; i < nd; i++)
          for (int j = 0; j <= i; j++) {
            double2 acc = make_double2(0.0, 0.0);
            for (int s = 0; s < T; s++) {
              const double2 xi = csub(*reinterpret_cast<const double2*>(Rh + (size_t)s * npad + 2 * ((size_t)i * N + t)), mean[i]);
              const double2 xj = csub(*reinterpret_cast<const double2*>(Rh + (size_t)s * npad + 2 * ((size_t)j * N + t)), mean[j]);
              acc = cadd(acc, cmulc(xi, xj));
            }
            Cv.a[i][j] = cscale(acc, 1.0 / (double)T);
            Cv.a[j][i] = make_double2(Cv.a[i][j].x, -Cv.a[i][j].y);
          }
        bool ok = cm_chol_lower(Cv, nd, Lm);   // C = Lc Lc^H
        cm_inv_lower(Lm, nd, W);               // C^-1 = W^H W
        cm_whw(W, nd, Q);
        ok = cm_chol_lower(Q, nd, Lm) && ok;   // la.cholesky(C^-1, lower=True)   (random.py:158)
        // Bartlett factor (random.py:121-130)
        CM Tm;
        for (int i = 0; i < nd; i++)
          for (int j = 0; j < nd; j++) Tm.a[i][j] = make_double2(0.0, 0.0);
        int l = 0;
        for (int i = 1; i < nd; i++)
          for (int j = 0; j < i; j++, l++) {
            double re, im;
            if (d.rng_mode == 0) {
              re = d.wn[((bi * N + t) * 2 + 0) * L + l];
              im = d.wn[((bi * N + t) * 2 + 1) * L + l];
            } else {
              normal_pair(ph((unsigned)(t * L + l), 0u, (unsigned)it, STREAM_X_WN), re, im);
            }
            Tm.a[i][j] = make_double2(re / 1.4142135623730951, im / 1.4142135623730951);
          }
        for (int i = 0; i < nd; i++) {
          const double gm = (d.rng_mode == 0) ? d.wg[(bi * N + t) * nd + i]
                                              : gamma_draw(ph, (unsigned)it, (unsigned long long)t * nd + i, STREAM_X_WG, (double)(T - i));
          if (!(gm > 0.0)) ok = false;
          Tm.a[i][i] = make_double2(sqrt(gm), 0.0);
        }
        // G = L Tm (lower);  Si_samp = G G^H / T;  S_new = T G^-H G^-1
        for (int i = 0; i < nd; i++)
          for (int j = 0; j < nd; j++) {
            double2 acc = make_double2(0.0, 0.0);
            for (int k = j; k <= i; k++) acc = cadd(acc, cmul(Lm.a[i][k], Tm.a[k][j]));
            G.a[i][j] = acc;
          }
        cm_inv_lower(G, nd, W);
        cm_whw(W, nd, Q);
        if (!ok) atomicOr(flag, 1);
        for (int i = 0; i < nd; i++)
          for (int j = 0; j < nd; j++) {
            const double2 v = cscale(Q.a[i][j], (double)T);
            S[((size_t)i * nd + j) * N + t] = v;
            d.samples[((bi * nd + i) * nd + j) * N + t] = v;
          }
      }
      __syncthreads();
      if (*flag) break;
      if (threadIdx.x == 0) d.n_done[b] = it + 1;
    }
    __syncthreads();
    if (threadIdx.x == 0 && *flag) d.failed[b] = 1;
    __syncthreads();
  }
}

}  // namespace dg

// ----------------------------------------------------------------------------- C ABI entry point
using namespace dg;

#define XCU(call)                                                                                         \
  do {                                                                                                    \
    cudaError_t e_ = (call);                                                                              \
    if (e_ != cudaSuccess) {                                                                              \
      char msg_[512];                                                                                     \
      snprintf(msg_, sizeof msg_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      dg_set_error(h, msg_);                                                                              \
      return (int)e_;                                                                                     \
    }                                                                                                     \
  } while (0)

template <typename Tp>
static cudaError_t x_stage(dgibbs_handle h, const char* name, const void* src, size_t n, const Tp** out) {
  *out = nullptr;
  if (!src || n == 0) return cudaSuccess;
  void* p;
  cudaError_t e = dg_pool_get(h, name, n * sizeof(Tp), &p);
  if (e != cudaSuccess) return e;
  *out = (const Tp*)p;
  return cudaMemcpyAsync(p, src, n * sizeof(Tp), cudaMemcpyHostToDevice, dg_stream(h));
}

extern "C" int dgibbs_cross_run(dgibbs_handle h, const dgibbs_cross_problem* p, void* samples, int32_t* n_done,
                                uint8_t* success) {
  if (!h) return -1;
  auto bad = [&](int idx, const char* what) {
    dg_set_error(h, (std::string("bad argument: ") + what).c_str());
    return -idx;
  };
  if (!p || p->struct_size != (int32_t)sizeof(dgibbs_cross_problem)) return bad(2, "problem struct");
  if (p->B <= 0 || p->N < 1 || p->T < 1 || p->M < 1 || p->niter < 0) return bad(2, "B, N, T, M must be positive");
  if (p->nd < 1 || p->nd > DGIBBS_CROSS_MAX_DATASETS) return bad(2, "nd must be in [1, DGIBBS_CROSS_MAX_DATASETS]");
  if (!window_id_valid(p->window_id)) return bad(2, "unknown window_id");
  if (p->rng_mode < 0 || p->rng_mode > 2) return bad(2, "unknown rng_mode");
  if (!p->data || !p->Ni || !p->fsel || !p->S0) return bad(2, "a required input pointer is NULL");
  if (p->rng_mode == DGIBBS_RNG_INJECTED && p->niter > 0 && (!p->w1 || !p->w2 || !p->wg || (p->nd > 1 && !p->wn)))
    return bad(2, "injected rng mode needs w1, w2, wn, wg");
  if (!samples || !n_done || !success) return bad(3, "samples, n_done and success are required");
  for (int f = 0; f < p->M; f++)
    if (p->fsel[f] < 0 || p->fsel[f] >= p->N) return bad(2, "fsel out of range");
  XCU(cudaSetDevice(dg_device(h)));
  cudaStream_t st = dg_stream(h);

  CrossBatch d{};
  d.B = p->B; d.nd = p->nd; d.N = p->N; d.T = p->T; d.M = p->M; d.niter = p->niter;
  d.window_id = p->window_id; d.rng_mode = p->rng_mode; d.seed = p->seed;
  d.nc = d.nd * d.N;
  d.nsys = 2 * d.nc;
  d.npad = (d.nsys + TILE - 1) / TILE * TILE;
  d.nblk = d.npad / TILE;
  d.RT = (d.T + TILE - 1) / TILE;
  d.L = d.nd * (d.nd - 1) / 2;
  d.use_fft = (d.N >= 8 && (d.N & (d.N - 1)) == 0 && d.N <= 4096) ? 1 : 0;
  d.nslots = d.B < dg_sm_count(h) ? d.B : dg_sm_count(h);

  const size_t B = d.B, nd = d.nd, N = d.N, T = d.T, M = d.M, nit = d.niter, L = d.L;
  XCU(x_stage<double2>(h, "x_data", p->data, B * nd * T * M, &d.data));
  XCU(x_stage<double>(h, "x_Ni", p->Ni, B * nd * M, &d.Ni));
  XCU(x_stage<long long>(h, "x_fsel", p->fsel, M, &d.fsel));
  XCU(x_stage<double2>(h, "x_S0", p->S0, B * nd * nd * N, &d.S0));
  XCU(x_stage<long long>(h, "x_gid", p->global_id, B, &d.gid));
  if (d.rng_mode == DGIBBS_RNG_INJECTED && nit > 0) {
    XCU(x_stage<double2>(h, "x_w1", p->w1, B * nit * N * nd * T, &d.w1));
    XCU(x_stage<double2>(h, "x_w2", p->w2, B * nit * nd * M * T, &d.w2));
    XCU(x_stage<double>(h, "x_wn", p->wn, B * nit * N * 2 * L, &d.wn));
    XCU(x_stage<double>(h, "x_wg", p->wg, B * nit * N * nd, &d.wg));
  }
  void* q;
  XCU(dg_pool_get(h, "x_h", 8 * B * nd * M, &q)); d.h = (double*)q;
  XCU(dg_pool_get(h, "x_z", 16 * B * nd * T * M, &q)); d.z = (double2*)q;
  XCU(dg_pool_get(h, "x_ctab", 16 * B * nd * N, &q)); d.ctab = (double2*)q;
  XCU(dg_pool_get(h, "x_S", 16 * B * nd * nd * N, &q)); d.S = (double2*)q;
  XCU(dg_pool_get(h, "x_samples", 16 * B * (nit ? nit : 1) * nd * nd * N, &q)); d.samples = (double2*)q;
  XCU(dg_pool_get(h, "x_ndone", 4 * B, &q)); d.n_done = (int*)q;
  XCU(dg_pool_get(h, "x_failed", 4 * B, &q)); d.failed = (int*)q;
  const size_t wrows = (size_t)d.npad + (size_t)d.RT * TILE;
  XCU(dg_pool_get(h, "work", 8 * (size_t)d.nslots * wrows * d.npad, &q)); d.work = (double*)q;
  XCU(dg_pool_get(h, "Wp", 8 * (size_t)d.nslots * d.nblk * WP_DOUBLES, &q)); d.Wp = (double*)q;
  XCU(dg_pool_get(h, "x_Si", 16 * (size_t)d.nslots * N * nd * nd, &q)); d.Si = (double2*)q;
  XCU(dg_pool_get(h, "x_U", 16 * (size_t)d.nslots * N * nd * nd, &q)); d.U = (double2*)q;
  if (d.use_fft) {
    XCU(dg_pool_get(h, "x_tw", 16 * (N / 2), &q));
    XCU(launch_build_twtab(d.N, (double2*)q, st));
    d.twtab = (const double2*)q;
  }
  cross_setup_kernel<<<d.B, 256, 0, st>>>(d);
  XCU(cudaGetLastError());
  if (d.niter > 0) {
    const size_t smem = sizeof(double) * (WP_DOUBLES + PIPE_DOUBLES + MISC_DOUBLES + 2 * TNST);
    XCU(cudaFuncSetAttribute(cross_sampler_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CUtensorMap tm;
    const int use_tma = tma_work_map(&tm, d.work, wrows, d.npad, wrows * d.npad, d.nslots, sizeof(double) * WP_DOUBLES, TILE, TKC);
    cross_sampler_kernel<<<d.nslots, THREADS, smem, st>>>(d, use_tma, tm);
    XCU(cudaGetLastError());
  }
  std::vector<int> failed(B);
  if (nit > 0) XCU(cudaMemcpyAsync(samples, d.samples, 16 * B * nit * nd * nd * N, cudaMemcpyDeviceToHost, st));
  XCU(cudaMemcpyAsync(n_done, d.n_done, 4 * B, cudaMemcpyDeviceToHost, st));
  XCU(cudaMemcpyAsync(failed.data(), d.failed, 4 * B, cudaMemcpyDeviceToHost, st));
  XCU(cudaStreamSynchronize(st));
  for (size_t b = 0; b < B; b++) success[b] = failed[b] ? 0 : 1;
  return 0;
}
