// Per-batch setup: projection table, window/whiten/pack, circulant first columns.
// Reference behaviour: _compute_delay_spectrum_inputs, draco/analysis/delay.py:1657-1710,
// fourier_matrix_* delay.py:1480-1613, window_generalised draco/util/tools.py:547-601.
#include "dense_core.cuh"
#include "window.cuh"

namespace dg {

// Ftab[k][col]: the (zero padded) matrix F of the FULL channel grid, stored channel-row major
// so that  y[t, col] = sum_k u[t, k] * Ftab[k, col]  is the masked inverse DFT.
__global__ void build_ftab_kernel(double* __restrict__ F, int N, int nsys, int npad, int K, int Kpad, int cplx) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long total = (long long)Kpad * npad;
  if (idx >= total) return;
  int k = (int)(idx / npad), col = (int)(idx % npad);
  double v = 0.0;
  if (k < K && col < nsys) {
    int i = k >> 1, pr = k & 1;
    double c, s;
    if (!cplx) {
      twiddle(((long long)i * col) % N, N, c, s);
      v = pr ? -s : c;  // delay.py:1507-1508
    } else {
      int t = col >> 1, pc = col & 1;
      twiddle(((long long)i * t) % N, N, c, s);
      v = (pr == pc) ? c : (pr ? -s : s);  // delay.py:1580-1583
    }
  }
  F[idx] = v;
}

cudaError_t launch_build_ftab(const DevBatch& d, double* Ftab, cudaStream_t st) {
  long long total = (long long)d.Kpad * d.npad;
  int blocks = (int)((total + 255) / 256);
  build_ftab_kernel<<<blocks, 256, 0, st>>>(Ftab, d.N, d.nsys, d.npad, d.K, d.Kpad, d.cplx);
  return cudaGetLastError();
}

// tw[m] = e^{2 pi i m / N}, m < N/2: the twiddles of the in-smem inverse FFT (sampler_dense.cu)
__global__ void build_twtab_kernel(double2* __restrict__ tw, int N) {
  int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= N / 2) return;
  double c, s;
  twiddle(m, N, c, s);
  tw[m] = make_double2(c, s);
}

cudaError_t launch_build_twtab(int N, double2* tw, cudaStream_t st) {
  build_twtab_kernel<<<(N / 2 + 255) / 256, 256, 0, st>>>(tw, N);
  return cudaGetLastError();
}

// One CTA (THREADS threads) per baseline.  FFT: the first column of the circulant comes from one in-smem inverse FFT
// of the squared channel weights (N log N) instead of the N * nch direct sum.
template <bool FFT>
__global__ void __launch_bounds__(THREADS) setup_baseline_kernel(DevBatch d, const double* __restrict__ S0) {
  extern __shared__ __align__(16) double sm[];
  const int b = blockIdx.x;
  const int N = d.N, nch = d.nch;
  // FFT: [bufA: N double2][bufB: N double2]; direct: [tabc: N][tabs: N if complex]
  double* tabc = sm;
  double* tabs = d.cplx ? sm + N : nullptr;
  double* hsq = sm + (FFT ? 4 * N : (d.cplx ? 2 * N : N));  // [nch]
  double* s_w = hsq + nch;                    // [nch] window
  double* s_re = s_w + nch;                   // [nch] sqrt(Ni_r) of the real part
  double* s_im = s_re + nch;                  // [nch] sqrt(Ni_r) of the imaginary part
  const int Tb = d.T[b], Mb = d.M[b];
  const long long coff = d.chan_off[b], doff = d.data_off[b];
  int* cidx = d.cidx + (size_t)b * nch;
  double* h = d.h + (size_t)b * d.Kpad;
  double* zT = d.zT + (size_t)b * d.Tmax * d.Kpad;

  for (int i = threadIdx.x; i < nch; i += blockDim.x) cidx[i] = -1;
  if (!FFT)
    for (int m = threadIdx.x; m < N; m += blockDim.x) {
      double c, s;
      twiddle(m, N, c, s);
      tabc[m] = c;
      if (tabs) tabs[m] = s;
    }
  __syncthreads();
  bool inorder = true;
  for (int ci = threadIdx.x; ci < Mb; ci += blockDim.x) {
    long long f = d.fsel[coff + ci];
    if (f >= 0 && f < nch) {
      // two entries on one channel (a mis-set freq_zero / freq_spacing / nfreq aliasing channels): reported, not raced
      if (atomicExch(&cidx[f], ci) != -1) atomicOr(d.setup_err, 2);
    } else {
      atomicOr(d.setup_err, 1);
    }
    inorder = inorder && f == ci;
  }
  const int ident = __syncthreads_and(inorder ? 1 : 0);
  if (threadIdx.x == 0) d.cident[b] = ident ? 1 : 0;

  // window * sqrt(Ni_r) on the full grid (delay.py:1679-1703)
  for (int i = threadIdx.x; i < d.Kpad / 2; i += blockDim.x) {
    double hre = 0.0, him = 0.0;
    double2 wre = make_double2(0.0, 0.0);
    if (i < nch) {
      int ci = cidx[i];
      if (ci >= 0) {
        double w = window_value(d.window_id, (double)i / (double)nch);
        double ni = d.Ni[coff + ci];
        bool purely_real = !d.cplx && (i == 0 || i == N / 2);  // delay.py:1692
        double nre = purely_real ? ni : 2.0 * ni;
        double nim = purely_real ? 0.0 : 2.0 * ni;
        double sre = sqrt(nre), sim = sqrt(nim);
        hre = w * sre;
        him = w * sim;
        s_w[i] = w;
        s_re[i] = sre;
        s_im[i] = sim;
        wre = make_double2(w, sre);
      }
      hsq[i] = hre * hre;
    }
    h[2 * i] = hre;
    h[2 * i + 1] = him;
    d.wsre[(size_t)b * (d.Kpad / 2) + i] = wre;
  }
  __syncthreads();

  // first column of the circulant  FTNiF[t,t'] = c[(t-t') mod N] = sum_i hsq_i e^{2 pi i (i k) / N}  (real part in
  // real mode)
  double* ct = d.ctab + (size_t)b * d.nsys;
  if (FFT) {
    double2* bufA = reinterpret_cast<double2*>(sm);
    for (int i = threadIdx.x; i < N; i += blockDim.x) bufA[i] = make_double2(i < nch ? hsq[i] : 0.0, 0.0);
    __syncthreads();
    const double2* res = fft_inverse_smem(bufA, bufA + N, d.twtab, N);
    for (int k = threadIdx.x; k < N; k += blockDim.x) {
      ct[k] = res[k].x;
      if (d.cplx) ct[N + k] = res[k].y;
    }
  } else {
    for (int k = threadIdx.x; k < N; k += blockDim.x) {
      double accr = 0.0, acci = 0.0;
      int m = 0;
      for (int i = 0; i < nch; i++) {
        double g = hsq[i];
        accr += g * tabc[m];
        if (tabs) acci += g * tabs[m];
        m += k;
        if (m >= N) m -= N;
      }
      ct[k] = accr;
      if (tabs) ct[N + k] = acci;
    }
  }

  // Is FTNiF = a I + b u u^T (u_t = (-1)^t)?  True for uniform weights, no window and a complete
  // channel set (b = 0), or with only the Nyquist channel missing (b = -Ni): SURVEY 8d.  Gate on
  // the computed first column, not on configuration flags.
  __syncthreads();
  bool isdiag = false;
  {
    const double tol = 1e-12;
    const double c0 = ct[0];
    double bc = 0.0;
    bool ok = c0 > 0.0;
    if (!d.cplx) {
      bc = (N > 1) ? -ct[1] : 0.0;
      if (fabs(bc) <= tol * c0) bc = 0.0;
      for (int k = 1 + threadIdx.x; k < N; k += blockDim.x)
        if (fabs(ct[k] - ((k & 1) ? -bc : bc)) > tol * c0) ok = false;
    } else {
      for (int k = threadIdx.x; k < N; k += blockDim.x)
        if ((k > 0 && fabs(ct[k]) > tol * c0) || fabs(ct[N + k]) > tol * c0) ok = false;
    }
    int all = __syncthreads_and(ok ? 1 : 0);
    isdiag = all && d.force_branch != 1 && (c0 - bc) > 0.0;  // CTA-uniform
    if (threadIdx.x == 0) {
      d.branch[b] = isdiag ? 3 : 1;
      d.diag_ab[2 * b] = c0 - bc;
      d.diag_ab[2 * b + 1] = bc;
    }
  }

  // A diagonal-branch baseline whose projection reads the raw data (project_fft.cu) needs no whitened copy.
  if (!(isdiag && d.skip_zT_diag)) {
    // whitened, windowed data, sample-major (delay.py:1676,1687,1707); a thread keeps its channels' factors
    const int halfK = d.Kpad / 2;
    for (int i = threadIdx.x; i < halfK; i += blockDim.x) {
      const int ci = i < nch ? cidx[i] : -1;
      const double w = ci >= 0 ? s_w[i] : 0.0, fre = ci >= 0 ? s_re[i] : 0.0, fim = ci >= 0 ? s_im[i] : 0.0;
      const long long src = doff + (ci >= 0 ? ci : 0);
      double* dst = zT + 2 * i;
  #pragma unroll 4
      for (int t = 0; t < d.Tmax; t++) {
        double2 z = make_double2(0.0, 0.0);
        if (t < Tb && ci >= 0) {
          const double2 v = ld_vis(d.data, d.data_c64, src + (long long)t * Mb);
          z = make_double2((v.x * w) * fre, (v.y * w) * fim);
        }
        *reinterpret_cast<double2*>(dst + (size_t)t * d.Kpad) = z;
      }
    }
  }

  for (int k = threadIdx.x; k < N; k += blockDim.x) d.S[(size_t)b * N + k] = S0[(size_t)b * N + k];
  if (threadIdx.x == 0) {
    d.n_done[b] = 0;
    d.failed[b] = 0;
  }
}

cudaError_t launch_setup_baselines(const DevBatch& d, const double* S0, cudaStream_t st) {
  bool fft = d.use_fft && d.twtab != nullptr;
  if (fft && sizeof(double) * (4 * (size_t)d.N + 4 * (size_t)d.nch) > 200 * 1024) fft = false;
  size_t smem = sizeof(double) * ((fft ? 4 : (d.cplx ? 2 : 1)) * (size_t)d.N + 4 * (size_t)d.nch);
  auto kern = fft ? setup_baseline_kernel<true> : setup_baseline_kernel<false>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  kern<<<d.B, THREADS, smem, st>>>(d, S0);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ debug / parity
__global__ void materialize_kernel(DevBatch d, int b, double* data_w, double* FTNih, double* FTNiF) {
  const int Tb = d.T[b], Mb = d.M[b];
  const long long coff = d.chan_off[b];
  const double* h = d.h + (size_t)b * d.Kpad;
  const double* zT = d.zT + (size_t)b * d.Tmax * d.Kpad;
  const double* ct = d.ctab + (size_t)b * d.nsys;
  long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  if (data_w)
    for (long long e = tid; e < 2LL * Mb * Tb; e += nth) {
      int row = (int)(e / Tb), t = (int)(e % Tb);
      long long f = d.fsel[coff + (row >> 1)];
      data_w[e] = zT[(size_t)t * d.Kpad + 2 * f + (row & 1)];
    }
  if (FTNih)
    for (long long e = tid; e < (long long)d.nsys * 2 * Mb; e += nth) {
      int r = (int)(e / (2 * Mb)), col = (int)(e % (2 * Mb));
      long long f = d.fsel[coff + (col >> 1)];
      long long k = 2 * f + (col & 1);
      FTNih[e] = d.Ftab[(size_t)k * d.npad + r] * h[k];
    }
  if (FTNiF)
    for (long long e = tid; e < (long long)d.nsys * d.nsys; e += nth) {
      int r = (int)(e / d.nsys), c = (int)(e % d.nsys);
      FTNiF[e] = circ_entry(ct, d.N, d.cplx, r, c);
    }
}

cudaError_t launch_materialize_inputs(const DevBatch& d, int b, double* data_w, double* FTNih, double* FTNiF,
                                      cudaStream_t st) {
  materialize_kernel<<<296, 256, 0, st>>>(d, b, data_w, FTNih, FTNiF);
  return cudaGetLastError();
}

// kind: 0 r2c (2M,N)  1 c2r (N,2M)  2 c2c (2M,2N)  3 complex (M,N) interleaved
__global__ void fourier_matrix_kernel(int kind, int N, const long long* fsel, int M, double* out) {
  long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  long long total = (kind == 2) ? 4LL * M * N : 2LL * M * N;
  for (long long e = tid; e < total; e += nth) {
    int row, col;
    double c, s, v;
    if (kind == 0) {  // delay.py:1480-1510
      row = (int)(e / N); col = (int)(e % N);
      long long f = fsel[row >> 1];
      twiddle((f * col) % N, N, c, s);
      v = (row & 1) ? -s : c;
    } else if (kind == 1) {  // delay.py:1513-1546
      row = (int)(e / (2 * M)); col = (int)(e % (2 * M));
      long long f = fsel[col >> 1];
      twiddle((f * row) % N, N, c, s);
      double mul = ((f == 0 || f == N / 2) ? 1.0 : 2.0) / (double)N;
      v = ((col & 1) ? -s : c) * mul;
    } else if (kind == 2) {  // delay.py:1549-1585
      row = (int)(e / (2 * N)); col = (int)(e % (2 * N));
      long long f = fsel[row >> 1];
      twiddle((f * (col >> 1)) % N, N, c, s);
      int pr = row & 1, pc = col & 1;
      v = (pr == pc) ? c : (pr ? -s : s);
    } else {  // delay.py:1588-1613  exp(-2 pi i t f / N)
      row = (int)(e / (2 * N)); col = (int)(e % (2 * N));
      long long f = fsel[row];
      twiddle((f * (col >> 1)) % N, N, c, s);
      v = (col & 1) ? -s : c;
    }
    out[e] = v;
  }
}

cudaError_t launch_fourier_matrix(int kind, int N, const long long* fsel, int M, double* out, cudaStream_t st) {
  fourier_matrix_kernel<<<296, 256, 0, st>>>(kind, N, fsel, M, out);
  return cudaGetLastError();
}

}  // namespace dg
