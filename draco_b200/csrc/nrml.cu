// NRML / maximum-posterior likelihood object on the GPU: the linear algebra of LogLikePS,
// draco/analysis/delayopt.py:61-236 (C ABI in include/dgnrml.h).  Built into libdgnrml.so, separate from
// libdgibbs.so, because it links cuBLAS / cuSOLVER for the plain library operations (ZHERK, ZGEMM, ZPOTRF, ZPOTRS on
// one M x M / N x N problem that fills the GPU by itself); parameter clipping, column scaling, the Hermitian
// fill + X - C, log-determinant / trace, the gradient contraction and the Hessian assembly are the kernels below.
//
// Device matrices are column-major complex128.  A numpy row-major [R, C] array read as column-major is its
// transpose: MF is transposed on upload, X (Hermitian) is conjugated.
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <cusolverDn.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/dgnrml.h"
#include "nrml_own.h"

namespace {

typedef cuDoubleComplex zc;

struct Libs {
  cublasHandle_t blas = nullptr;
  cusolverDnHandle_t solver = nullptr;
};
Libs g_libs[64];

// ------------------------------------------------------------------------------------------------ kernels
__global__ void conj_copy_kernel(const zc* __restrict__ in, zc* __restrict__ out, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = make_cuDoubleComplex(in[i].x, -in[i].y);
}

// row-major [R, C] -> column-major [R, C] (32 x 32 tiles through shared memory)
__global__ void transpose_kernel(const zc* __restrict__ in, zc* __restrict__ out, int R, int C) {
  __shared__ zc tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int r = r0 + j, c = c0 + threadIdx.x;
    if (r < R && c < C) tile[j][threadIdx.x] = in[(size_t)r * C + c];
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int c = c0 + j, r = r0 + threadIdx.x;
    if (r < R && c < C) out[(size_t)c * R + r] = tile[threadIdx.x][j];
  }
}

// sq_t = sqrt(exp(clip(x_t)))  (delayopt.py:137-140, 153)
__global__ void prep_kernel(const double* __restrict__ x, double* __restrict__ sq, int N, double lo, double hi) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= N) return;
  const double s = fmin(fmax(x[t], lo), hi);
  sq[t] = sqrt(exp(s));
}

// U[:, t] = sqrt(S_t) MF[:, t]  (delayopt.py:153)
__global__ void scale_cols_kernel(const zc* __restrict__ MF, const double* __restrict__ sq, zc* __restrict__ U, int M,
                                  int N) {
  const size_t n = (size_t)M * N;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    const double s = sq[e / M];
    U[e] = make_cuDoubleComplex(MF[e].x * s, MF[e].y * s);
  }
}

// the lower triangle of C = U U^H comes from ZHERK: add the noise variances to the diagonal, mirror it, and form
// XC = X - C  (delayopt.py:143-144, 151)
__global__ void finish_cov_kernel(zc* __restrict__ Cm, const zc* __restrict__ X, const double* __restrict__ noise,
                                  zc* __restrict__ XC, int M) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;  // row i (fast), column j
  if (i >= M || i < j) return;
  zc c = Cm[(size_t)j * M + i];
  if (i == j) c = make_cuDoubleComplex(c.x + noise[i], 0.0);
  Cm[(size_t)j * M + i] = c;
  const zc xl = X[(size_t)j * M + i];
  XC[(size_t)j * M + i] = make_cuDoubleComplex(xl.x - c.x, xl.y - c.y);
  if (i != j) {
    Cm[(size_t)i * M + j] = make_cuDoubleComplex(c.x, -c.y);
    const zc xu = X[(size_t)i * M + j];
    XC[(size_t)i * M + j] = make_cuDoubleComplex(xu.x - c.x, xu.y + c.y);
  }
}

// out[0] = sum_i f(A_ii) with f = 2 log Re (log-determinant from the Cholesky factor) or Re (trace); one CTA
__global__ void diag_reduce_kernel(const zc* __restrict__ A, int M, int want_log, double* __restrict__ out) {
  __shared__ double red[32];
  double s = 0.0;
  for (int i = threadIdx.x; i < M; i += blockDim.x) {
    const double v = A[(size_t)i * M + i].x;
    s += want_log ? 2.0 * log(v) : v;
  }
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += red[w];
    out[0] = t;
  }
}

// t_a = -sum_i Re(conj(Ut[i, a]) W[i, a])  (delayopt.py:205, 232): one warp per column
__global__ void col_dot_kernel(const zc* __restrict__ Ut, const zc* __restrict__ W, double* __restrict__ t, int M, int N) {
  const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (a >= N) return;
  const zc* u = Ut + (size_t)a * M;
  const zc* w = W + (size_t)a * M;
  double s = 0.0;
  for (int i = lane; i < M; i += 32) s += u[i].x * w[i].x + u[i].y * w[i].y;
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) t[a] = -s;
}

// H[a, b] = nsamp (Re(G_ab conj(G_ba)) [+ 2 Re(K_ab G_ba) + delta_ab t_a])  (delayopt.py:221-236), row-major out
__global__ void hessian_kernel(const zc* __restrict__ G, const zc* __restrict__ K, const double* __restrict__ t,
                               double* __restrict__ H, int N, int exact, double nsamp, int ld) {
  constexpr int TS = 16;  // blockDim = (TS, TS)
  __shared__ zc gab[TS][TS + 1], kab[TS][TS + 1], gba[TS][TS + 1];
  const int A0 = blockIdx.y * TS, B0 = blockIdx.x * TS, tx = threadIdx.x;
  {
    const int ty = threadIdx.y;
    const int a = A0 + tx, b = B0 + ty;  // a fast: column-major reads of G[a, b], K[a, b]
    if (a < N && b < N) {
      gab[ty][tx] = G[(size_t)b * ld + a];
      if (exact) kab[ty][tx] = K[(size_t)b * ld + a];
    }
    const int bb = B0 + tx, aa = A0 + ty;  // b fast: G[b, a]
    if (aa < N && bb < N) gba[ty][tx] = G[(size_t)aa * ld + bb];
  }
  __syncthreads();
  {
    const int ty = threadIdx.y;
    const int a = A0 + ty, b = B0 + tx;  // b fast: row-major writes
    if (a >= N || b >= N) return;
    const zc g1 = gab[tx][ty], g2 = gba[ty][tx];
    double h = g1.x * g2.x + g1.y * g2.y;  // Re(g1 conj(g2))
    if (exact) {
      const zc k = kab[tx][ty];
      h += 2.0 * (k.x * g2.x - k.y * g2.y);
      if (a == b) h += t[a];
    }
    H[(size_t)a * N + b] = nsamp * h;
  }
}

// rows of Z (M x T, column-major) scaled by the pseudo-inverse spectrum: 1 / lambda_i above the cut, else 0
__global__ void scale_rows_pinv_kernel(zc* __restrict__ Z, const double* __restrict__ lam, int M, int T, double cut) {
  const size_t n = (size_t)M * T;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    const double l = lam[e % M];
    const double f = l > cut ? 1.0 / l : 0.0;
    Z[e] = make_cuDoubleComplex(Z[e].x * f, Z[e].y * f);
  }
}

// var over the T columns of Z (N x T, column-major): out[n] = mean_t |z_nt - mean_t z_nt|^2
__global__ void var_cols_kernel(const zc* __restrict__ Z, int N, int T, double* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  double sr = 0.0, si = 0.0;
  for (int t = 0; t < T; t++) {
    sr += Z[(size_t)t * N + n].x;
    si += Z[(size_t)t * N + n].y;
  }
  const double mr = sr / T, mi = si / T;
  double v = 0.0;
  for (int t = 0; t < T; t++) {
    const double dx = Z[(size_t)t * N + n].x - mr, dy = Z[(size_t)t * N + n].y - mi;
    v += dx * dx + dy * dy;
  }
  out[n] = v / T;
}

}  // namespace

struct dgnrml_s {
  int device = 0, M = 0, N = 0, nsamp = 0, exact = 1;
  double lo = 0, hi = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  zc *X = nullptr, *MF = nullptr, *U = nullptr, *Ut = nullptr, *W = nullptr, *C = nullptr, *XC = nullptr, *CiX = nullptr,
     *G = nullptr, *K = nullptr, *work = nullptr;
  double *noise = nullptr, *x = nullptr, *sq = nullptr, *t = nullptr, *H = nullptr, *scal = nullptr;
  int* info = nullptr;
  int lwork = 0;
  std::vector<double> x_at;
  bool factored = false, have_hess_terms = false;
  bool own = true;   // every contraction / factorisation / solve on this repository's kernels (nrml_own.cu)
  dgn::Own o;
  double last_ms = 0, last_flop = 0;
  long long last_launches = 0;
  std::string err;
};

#define NCU(call)                                                                                  \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      s->err = std::string(#call) + " failed: " + cudaGetErrorString(e_);                          \
      return 100 + (int)e_;                                                                        \
    }                                                                                              \
  } while (0)
#define NBL(call)                                                                                  \
  do {                                                                                             \
    int e_ = (int)(call);                                                                          \
    if (e_ != 0) {                                                                                 \
      s->err = std::string(#call) + " failed with library status " + std::to_string(e_);           \
      return 1000 + e_;                                                                            \
    }                                                                                              \
  } while (0)

static int nrml_free(dgnrml s) {
  cudaSetDevice(s->device);
  dgn::own_free(s->o);
  for (void* p : {(void*)s->X, (void*)s->MF, (void*)s->U, (void*)s->Ut, (void*)s->W, (void*)s->C, (void*)s->XC,
                  (void*)s->CiX, (void*)s->G, (void*)s->K, (void*)s->work, (void*)s->noise, (void*)s->x, (void*)s->sq,
                  (void*)s->t, (void*)s->H, (void*)s->scal, (void*)s->info})
    if (p) cudaFree(p);
  if (s->ev0) cudaEventDestroy(s->ev0);
  if (s->ev1) cudaEventDestroy(s->ev1);
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
  return 0;
}

extern "C" int dgnrml_create(int32_t device_id, int32_t M, int32_t N, int32_t nsamp, int32_t exact_hessian,
                             double log_lo, double log_hi, const void* X, const void* MF, const double* noise,
                             dgnrml* out) {
  if (!out) return -1;
  *out = nullptr;
  if (M <= 0 || N <= 0 || nsamp <= 0 || !X || !MF || !noise || device_id < 0 || device_id >= 64) return -2;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device_id >= ndev) return -3;  // no CPU fallback
  dgnrml s = new dgnrml_s;
  *out = s;  // returned even on failure so that dgnrml_last_error can be read; the caller destroys it
  s->device = device_id; s->M = M; s->N = N; s->nsamp = nsamp; s->exact = exact_hessian ? 1 : 0;
  s->lo = log_lo < log_hi ? log_lo : log_hi;
  s->hi = log_lo < log_hi ? log_hi : log_lo;
  NCU(cudaSetDevice(device_id));
  Libs& L = g_libs[device_id];
  if (!L.blas) NBL(cublasCreate(&L.blas));
  if (!L.solver) NBL(cusolverDnCreate(&L.solver));
  NCU(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  NCU(cudaEventCreate(&s->ev0));
  NCU(cudaEventCreate(&s->ev1));
  const size_t MM = (size_t)M * M, MN = (size_t)M * N, NN = (size_t)N * N;
  NCU(cudaMalloc(&s->X, MM * 16)); NCU(cudaMalloc(&s->C, MM * 16)); NCU(cudaMalloc(&s->XC, MM * 16));
  NCU(cudaMalloc(&s->CiX, MM * 16));
  NCU(cudaMalloc(&s->MF, MN * 16)); NCU(cudaMalloc(&s->U, MN * 16)); NCU(cudaMalloc(&s->Ut, MN * 16));
  NCU(cudaMalloc(&s->W, MN * 16));
  NCU(cudaMalloc(&s->G, NN * 16)); NCU(cudaMalloc(&s->K, NN * 16)); NCU(cudaMalloc(&s->H, NN * 8));
  NCU(cudaMalloc(&s->noise, (size_t)M * 8)); NCU(cudaMalloc(&s->x, (size_t)N * 8)); NCU(cudaMalloc(&s->sq, (size_t)N * 8));
  NCU(cudaMalloc(&s->t, (size_t)N * 8)); NCU(cudaMalloc(&s->scal, 4 * 8)); NCU(cudaMalloc(&s->info, 4 * sizeof(int)));
  NBL(cusolverDnSetStream(L.solver, s->stream));
  NBL(cusolverDnZpotrf_bufferSize(L.solver, CUBLAS_FILL_MODE_LOWER, M, s->C, M, &s->lwork));
  NCU(cudaMalloc(&s->work, (size_t)(s->lwork > 0 ? s->lwork : 1) * 16));
  // uploads: X conjugated (row-major Hermitian = transpose), MF transposed to column-major; W / CiX are scratch
  NCU(cudaMemcpyAsync(s->CiX, X, MM * 16, cudaMemcpyHostToDevice, s->stream));
  conj_copy_kernel<<<1024, 256, 0, s->stream>>>(s->CiX, s->X, MM);
  NCU(cudaMemcpyAsync(s->W, MF, MN * 16, cudaMemcpyHostToDevice, s->stream));
  transpose_kernel<<<dim3((N + 31) / 32, (M + 31) / 32), dim3(32, 8), 0, s->stream>>>(s->W, s->MF, M, N);
  s->own = getenv("DGNRML_LIBRARY") == nullptr;  // DGNRML_LIBRARY=1: cuBLAS / cuSOLVER path (A/B)
  if (s->own) {
    int sm_count = 148;
    NCU(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device_id));
    NCU(dgn::own_alloc(s->o, M, N, sm_count, s->stream));
    NCU(dgn::own_set_mf(s->o, s->W, s->stream));  // W still holds the row-major MF
  }
  NCU(cudaMemcpyAsync(s->noise, noise, (size_t)M * 8, cudaMemcpyHostToDevice, s->stream));
  NCU(cudaGetLastError());
  NCU(cudaStreamSynchronize(s->stream));
  return 0;
}

extern "C" int dgnrml_eval(dgnrml s, const double* x, int32_t what, double* value, double* gradient, double* hessian) {
  if (!s) return -1;
  if (!x || ((what & DGNRML_VALUE) && !value) || ((what & DGNRML_GRADIENT) && !gradient) ||
      ((what & DGNRML_HESSIAN) && !hessian))
    return -2;
  NCU(cudaSetDevice(s->device));
  Libs& L = g_libs[s->device];
  NBL(cublasSetStream(L.blas, s->stream));
  NBL(cusolverDnSetStream(L.solver, s->stream));
  const int M = s->M, N = s->N;
  const size_t MM = (size_t)M * M, MN = (size_t)M * N;
  const zc one = make_cuDoubleComplex(1.0, 0.0), zero = make_cuDoubleComplex(0.0, 0.0);
  const double done = 1.0, dzero = 0.0;
  double flop = 0.0;
  NCU(cudaEventRecord(s->ev0, s->stream));
  const bool same = s->factored && s->x_at.size() == (size_t)N && std::memcmp(s->x_at.data(), x, (size_t)N * 8) == 0;
  if (!same) {  // _precompute, delayopt.py:120-164
    s->factored = false;
    s->have_hess_terms = false;
    NCU(cudaMemcpyAsync(s->x, x, (size_t)N * 8, cudaMemcpyHostToDevice, s->stream));
    prep_kernel<<<(N + 255) / 256, 256, 0, s->stream>>>(s->x, s->sq, N, s->lo, s->hi);
    scale_cols_kernel<<<2048, 256, 0, s->stream>>>(s->MF, s->sq, s->U, M, N);
    if (s->own) {
      int info = 0;
      long long nl = 0;
      NCU(dgn::own_precompute(s->o, s->stream, s->sq, s->U, s->X, s->noise, s->t, s->scal, &info, (int64_t*)&nl));
      s->last_launches = nl + 3;
      if (info != 0) {
        s->err = "covariance not positive definite (block column " + std::to_string(info - 1) + ")";
        return 1;
      }
      s->x_at.assign(x, x + N);
      s->factored = true;
      flop += 4.0 * M * (double)M * N + 4.0 / 3.0 * M * (double)M * M + 16.0 * M * (double)M * N + 8.0 * M * (double)M * M;
    } else {
    NBL(cublasZherk(L.blas, CUBLAS_FILL_MODE_LOWER, CUBLAS_OP_N, M, N, &done, s->U, M, &dzero, s->C, M));
    finish_cov_kernel<<<dim3((M + 127) / 128, M), 128, 0, s->stream>>>(s->C, s->X, s->noise, s->XC, M);
    NBL(cusolverDnZpotrf(L.solver, CUBLAS_FILL_MODE_LOWER, M, s->C, M, s->work, s->lwork, s->info));
    int info = 0;
    NCU(cudaMemcpyAsync(&info, s->info, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
    NCU(cudaStreamSynchronize(s->stream));
    if (info != 0) {
      s->err = "covariance not positive definite (leading minor " + std::to_string(info) + ")";
      return 1;
    }
    NCU(cudaMemcpyAsync(s->Ut, s->U, MN * 16, cudaMemcpyDeviceToDevice, s->stream));
    NBL(cusolverDnZpotrs(L.solver, CUBLAS_FILL_MODE_LOWER, M, N, s->C, M, s->Ut, M, s->info));
    NBL(cublasZgemm(L.blas, CUBLAS_OP_N, CUBLAS_OP_N, M, N, M, &one, s->XC, M, s->Ut, M, &zero, s->W, M));
    col_dot_kernel<<<(N + 7) / 8, 256, 0, s->stream>>>(s->Ut, s->W, s->t, M, N);
    s->x_at.assign(x, x + N);
    s->factored = true;
    flop += 4.0 * M * (double)M * N + 4.0 / 3.0 * M * (double)M * M + 16.0 * M * (double)M * N;
    }  // library path
  }
  double scal[2] = {0.0, 0.0};
  if ((what & DGNRML_VALUE) && s->own) {  // log det C and trace(C^-1 X) came out of the factorisation (X rode along)
    NCU(cudaMemcpyAsync(scal, s->scal, 16, cudaMemcpyDeviceToHost, s->stream));
  } else if (what & DGNRML_VALUE) {  // delayopt.py:166-189
    diag_reduce_kernel<<<1, 256, 0, s->stream>>>(s->C, M, 1, s->scal);
    NCU(cudaMemcpyAsync(s->CiX, s->X, MM * 16, cudaMemcpyDeviceToDevice, s->stream));
    NBL(cusolverDnZpotrs(L.solver, CUBLAS_FILL_MODE_LOWER, M, M, s->C, M, s->CiX, M, s->info));
    diag_reduce_kernel<<<1, 256, 0, s->stream>>>(s->CiX, M, 0, s->scal + 1);
    NCU(cudaMemcpyAsync(scal, s->scal, 16, cudaMemcpyDeviceToHost, s->stream));
    flop += 8.0 * M * (double)M * M;
  }
  if (what & DGNRML_GRADIENT) NCU(cudaMemcpyAsync(gradient, s->t, (size_t)N * 8, cudaMemcpyDeviceToHost, s->stream));
  if (what & DGNRML_HESSIAN) {  // delayopt.py:209-236
    if (!s->have_hess_terms) {
      const zc *Gp = s->G, *Kp = s->K;
      int ldg = N;
      if (s->own) {
        long long nl = 0;
        NCU(dgn::own_hessian_terms(s->o, s->stream, s->exact, (int64_t*)&nl));
        s->last_launches += nl;
        Gp = reinterpret_cast<const zc*>(s->o.G);
        Kp = reinterpret_cast<const zc*>(s->o.K);
        ldg = s->o.ldG();
      } else {
        NBL(cublasZgemm(L.blas, CUBLAS_OP_C, CUBLAS_OP_N, N, N, M, &one, s->U, M, s->Ut, M, &zero, s->G, N));
        if (s->exact) NBL(cublasZgemm(L.blas, CUBLAS_OP_C, CUBLAS_OP_N, N, N, M, &one, s->Ut, M, s->W, M, &zero, s->K, N));
      }
      flop += 8.0 * M * (double)N * N * (s->exact ? 2.0 : 1.0);
      hessian_kernel<<<dim3((N + 15) / 16, (N + 15) / 16), dim3(16, 16), 0, s->stream>>>(Gp, Kp, s->t, s->H, N, s->exact,
                                                                                        (double)s->nsamp, ldg);
      s->have_hess_terms = true;
    }
    NCU(cudaEventRecord(s->ev1, s->stream));
    NCU(cudaMemcpyAsync(hessian, s->H, (size_t)N * N * 8, cudaMemcpyDeviceToHost, s->stream));
  } else {
    NCU(cudaEventRecord(s->ev1, s->stream));
  }
  NCU(cudaGetLastError());
  NCU(cudaStreamSynchronize(s->stream));
  if (what & DGNRML_VALUE) *value = (double)s->nsamp * (scal[0] + scal[1]);
  if (what & DGNRML_GRADIENT)
    for (int a = 0; a < N; a++) gradient[a] *= (double)s->nsamp;
  float ms = 0.f;
  cudaEventElapsedTime(&ms, s->ev0, s->ev1);
  s->last_ms = ms;
  s->last_flop = flop;
  return 0;
}

// Starting point of the optimiser, delayopt.py:505-506: (data @ pinv(MF^T, rtol)).var(axis=0).  With the Hermitian
// eigendecomposition MF MF^H = V diag(lambda) V^H the pseudo-inverse is pinv(MF^T) = conj(V) diag(1/lambda) V^T conj(MF)
// over the eigenvalues above (rtol sigma_max)^2 = rtol^2 lambda_max, and the variance does not see the conjugation:
// three small GEMMs after one ZHERK + ZHEEVD instead of an SVD of the N x M matrix on the host.
extern "C" int dgnrml_initial_guess(dgnrml s, const void* data, int32_t T, double rtol, double* S0) {
  if (!s) return -1;
  if (!data || T <= 0 || !S0 || !(rtol >= 0.0)) return -2;
  NCU(cudaSetDevice(s->device));
  Libs& L = g_libs[s->device];
  NBL(cublasSetStream(L.blas, s->stream));
  NBL(cusolverDnSetStream(L.solver, s->stream));
  const int M = s->M, N = s->N;
  const zc one = make_cuDoubleComplex(1.0, 0.0), zero = make_cuDoubleComplex(0.0, 0.0);
  const double done = 1.0, dzero = 0.0;
  zc *V = s->CiX, *D = nullptr, *Z1 = nullptr, *Z2 = nullptr, *Z3 = nullptr, *work = nullptr;
  double* lam = nullptr;
  int lwork = 0, info = 0;
  auto cleanup = [&]() {
    for (void* p : {(void*)D, (void*)Z1, (void*)Z2, (void*)Z3, (void*)work, (void*)lam})
      if (p) cudaFree(p);
  };
  NCU(cudaMalloc(&D, (size_t)M * T * 16)); NCU(cudaMalloc(&Z1, (size_t)M * T * 16)); NCU(cudaMalloc(&Z2, (size_t)M * T * 16));
  NCU(cudaMalloc(&Z3, (size_t)N * T * 16)); NCU(cudaMalloc(&lam, (size_t)M * 8));
  // row-major data [T, M] is column-major [M, T]: column t = sample t
  NCU(cudaMemcpyAsync(D, data, (size_t)M * T * 16, cudaMemcpyHostToDevice, s->stream));
  NBL(cublasZherk(L.blas, CUBLAS_FILL_MODE_LOWER, CUBLAS_OP_N, M, N, &done, s->MF, M, &dzero, V, M));
  NBL(cusolverDnZheevd_bufferSize(L.solver, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_LOWER, M, V, M, lam, &lwork));
  NCU(cudaMalloc(&work, (size_t)(lwork > 0 ? lwork : 1) * 16));
  NBL(cusolverDnZheevd(L.solver, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_LOWER, M, V, M, lam, work, lwork, s->info));
  NCU(cudaMemcpyAsync(&info, s->info, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
  double lmax = 0.0;  // eigenvalues ascend
  NCU(cudaMemcpyAsync(&lmax, lam + (M - 1), 8, cudaMemcpyDeviceToHost, s->stream));
  NCU(cudaStreamSynchronize(s->stream));
  if (info != 0) {
    cleanup();
    s->err = "eigendecomposition for the pseudo-inverse start did not converge (info " + std::to_string(info) + ")";
    return 1;
  }
  NBL(cublasZgemm(L.blas, CUBLAS_OP_C, CUBLAS_OP_N, M, T, M, &one, V, M, D, M, &zero, Z1, M));       // V^H D
  scale_rows_pinv_kernel<<<256, 256, 0, s->stream>>>(Z1, lam, M, T, rtol * rtol * lmax);
  NBL(cublasZgemm(L.blas, CUBLAS_OP_N, CUBLAS_OP_N, M, T, M, &one, V, M, Z1, M, &zero, Z2, M));      // V (.)
  NBL(cublasZgemm(L.blas, CUBLAS_OP_C, CUBLAS_OP_N, N, T, M, &one, s->MF, M, Z2, M, &zero, Z3, N));  // MF^H (.)
  var_cols_kernel<<<(N + 127) / 128, 128, 0, s->stream>>>(Z3, N, T, s->sq);
  NCU(cudaMemcpyAsync(S0, s->sq, (size_t)N * 8, cudaMemcpyDeviceToHost, s->stream));
  NCU(cudaGetLastError());
  NCU(cudaStreamSynchronize(s->stream));
  cleanup();
  s->factored = false;  // CiX / sq were used as scratch
  return 0;
}

extern "C" int dgnrml_destroy(dgnrml s) { return s ? nrml_free(s) : 0; }
extern "C" const char* dgnrml_last_error(dgnrml s) { return s ? s->err.c_str() : "null handle"; }
extern "C" int dgnrml_last_timing(dgnrml s, double* ms, double* flop) {
  if (!s) return -1;
  if (ms) *ms = s->last_ms;
  if (flop) *flop = s->last_flop;
  return 0;
}
