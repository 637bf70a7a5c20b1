// Internal (non-ABI) declarations shared by the .cu files of libdgibbs.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dg {

// Device-side description of one batch.  All pointers are device pointers.
struct DevBatch {
  // geometry
  int B;       // baselines
  int N;       // delays
  int nsys;    // system size of the delay-space branch: N (real) or 2N (complex time domain)
  int npad;    // nsys rounded up to TILE
  int nblk;    // npad / TILE
  int nch;     // channels of the FULL grid: N/2+1 (real) or N (complex)
  int K;       // 2*nch  rows of the projection table
  int Kpad;    // K rounded up to KC
  int Tmax;    // max_b T[b]
  int RT;      // RHS row tiles: ceil(Tmax / TILE)
  int niter;
  int window_id;
  int cplx;
  int rng_mode;
  unsigned long long seed;
  int nslots;  // concurrent CTAs (= workspace slots)
  int wiener;  // 1: stop after the solve of iteration 0 and keep x (Wiener filter)
  int ctab_smem;  // circulant table fits in shared memory
  int force_branch;  // 0 auto; DGIBBS_BRANCH_DELAY_SPACE (1) = never take the diagonal shortcut
  int project_only;  // dense kernel: stop after phase 0 and export y^T = (F^T h z)^T to yhat

  // caller inputs
  const int* T;
  const int* M;
  const long long* data_off;
  const long long* chan_off;
  const double2* data;   // visibilities: complex128, or complex64 (float2) when data_c64 (read through ld_vis)
  int data_c64;
  const double* Ni;
  const long long* fsel;
  const long long* gid;  // may be null
  const double* w1;
  const long long* w1_off;
  const double* w2;
  const long long* w2_off;
  const double* chi2;

  // derived per-baseline tables
  int* cidx;     // [B][nch]   compact channel index on the full grid, -1 = masked
  double* h;     // [B][Kpad]  window * sqrt(Ni_r) on the full grid (0 where masked)
  double* zT;    // [B][Tmax][Kpad] whitened data, transposed (sample-major)
  double2* wsre; // [B][Kpad/2] (window, sqrt(Ni_r) of the real part) per channel of the full grid, 0 where masked
  int* setup_err;         // [1] set by the setup kernel: bit 0 = an fsel entry outside the channel grid, bit 1 = a duplicate
  unsigned char* cident;  // [B] 1: the kept channels are 0 .. M-1 in order (cidx[i] = i for i < M): no index look-up needed
  int skip_zT_diag;  // 1: diagonal-branch baselines get no zT (their projection reads the raw data, project_fft.cu)
  double* ctab;  // [B][nsys]  first column of the circulant FTNiF (cplx: cr[N] then ci[N])
  double* S;     // [B][N]     chain state
  double* samples;  // [niter][B][N]
  int* n_done;      // [B] completed iterations
  int* failed;      // [B] 1 once a factorisation met a non-positive pivot
  unsigned char* branch;  // [B] 1 = dense delay-space system, 3 = diagonal(+rank-1) system
  double* diag_ab;        // [B][2]  FTNiF = a I + b u u^T (u = (-1)^t) when branch == 3
  const int* blist;       // baselines this launch works on (null = all B)
  int nlist;
  int li_begin, li_end;   // dense kernel: positions [li_begin, li_end) of the list this launch covers (li_end 0 = to the end)
  double* yhat;           // [ndiag][Tmax][ypitch] projected data of the diagonal-branch baselines (zero beyond nsys)
  int ypitch;             // row pitch of yhat: npad + 128, so that any even column split of a cluster stays inside a row
  const int* ycompact;    // [B] row block of a baseline in yhat, or -1

  // shared tables
  const double* Ftab;  // [Kpad][npad] projection table  F[k][col]
  const double2* twtab;  // [N/2] exact twiddles e^{2 pi i m / N} (N a power of two), else null
  const double2* stwtab; // per-stage twiddle tables of fft_inverse_fused (fft_fused.cuh), else null
  int use_fft;           // 1: masked inverse DFT of phase 0 runs as an in-smem FFT (N = 2^m <= 4096)
  int use_tma;           // dense kernel: NT updates fed by the TMA ring (tma_pipe.cuh); set per launch

  // per-slot workspace
  size_t work_stride;  // doubles per slot of `work`: (npad + RT*TILE) * npad for L + RHS, then N * Tmax (Sh w1, _t branch)
  double* work;  // [nslots][(npad + RT*TILE)][npad]  L (lower) with W=inv(L_jj) in the diagonal
                 //                                   blocks, then the RHS rows (sample-major)
  double* Wp;    // [nslots][nblk][8704]  packed inv(L_jj) per block column (for the backward solve)
  double* uT;    // [nslots][RT*TILE][Kpad]
  double* Si;    // [nslots][npad]
  double* xout;  // wiener: [sum T][nsys] output; else null
  const long long* xoff;  // [B] row offsets into xout
  long long* prof;        // optional [16] phase cycle counters (debug), else null
};

// handle services for entry points implemented outside dgibbs_api.cu (sampler_cross.cu)
}  // namespace dg
struct dgibbs_handle_s;
cudaError_t dg_pool_get(dgibbs_handle_s* h, const char* name, size_t bytes, void** out);
cudaStream_t dg_stream(dgibbs_handle_s* h);
int dg_device(dgibbs_handle_s* h);
int dg_sm_count(dgibbs_handle_s* h);
void dg_set_error(dgibbs_handle_s* h, const char* msg);
// per-handle 64-bit tags (0 when new): lets an entry point remember what a pooled table holds between calls
long long* dg_tag(dgibbs_handle_s* h, const char* name);
cudaEvent_t dg_event(dgibbs_handle_s* h, int i);  // the handle's pre-created timing events, i in [0, 8)
namespace dg {

// setup_kernels.cu
cudaError_t launch_build_ftab(const DevBatch& d, double* Ftab, cudaStream_t st);
cudaError_t launch_build_twtab(int N, double2* tw, cudaStream_t st);
cudaError_t launch_build_stage_twiddles(int N, double2* tw, cudaStream_t st);  // project_fft.cu; 2 N entries
cudaError_t launch_setup_baselines(const DevBatch& d, const double* S0, cudaStream_t st);
cudaError_t launch_materialize_inputs(const DevBatch& d, int b, double* data_w, double* FTNih,
                                      double* FTNiF, cudaStream_t st);
cudaError_t launch_fourier_matrix(int kind, int N, const long long* fsel, int M, double* out,
                                  cudaStream_t st);

// sampler_dense.cu
size_t dense_smem_bytes(const DevBatch& d);
size_t dense_wp_doubles_per_slot(const DevBatch& d);
cudaError_t dense_configure(const DevBatch& d);
// How a dense launch is divided: the full waves run one CTA per chain, the remainder on clusters of C_tail CTAs per chain.
struct DensePlan {
  int n_main, grid_main;          // chains / CTAs of the C = 1 launch (0: none)
  int n_tail, grid_tail, C_tail;  // chains / clusters / cluster size of the second launch (n_tail 0: none)
};
DensePlan plan_dense(const DevBatch& d, int nchains, int sm_count, int force_cluster);
// force_cluster: 0 = plan as above; 1, 2, 4 = every chain on clusters of that size.  *launches is incremented per launch.
cudaError_t launch_dense_sampler(const DevBatch& d, int it_begin, int it_end, cudaStream_t st, int sm_count = 0,
                                 int force_cluster = 0, int* launches = nullptr);

// sampler_diag.cu
// allow_resident: every listed baseline has b <= 0, so the cluster-resident kernel may be used
cudaError_t launch_diag_sampler(const DevBatch& d, int it_begin, int it_end, int sm_count, cudaStream_t st,
                                bool allow_resident);

// sampler_freq.cu
bool diag_project_fft_fits(const DevBatch& d);
cudaError_t launch_diag_project_fft(const DevBatch& d, int sm_count, cudaStream_t st);
bool freq_branch_fits(const DevBatch& d, int M, int T);
// uniform_npt: the padded 2M system size shared by every chain of d.blist (feeds the TMA operand ring), or 0
cudaError_t launch_freq_sampler(const DevBatch& d, int it_begin, int it_end, cudaStream_t st, int uniform_npt);

// finish_kernels.cu
cudaError_t launch_finish(const DevBatch& d, const double* S0, double median_frac, double* spectrum,
                          double* samples_out, int* n_done_out, unsigned char* success_out,
                          double* final_S, cudaStream_t st);

}  // namespace dg
