// C ABI of libdgibbs (see include/dgibbs.h).  Host-side orchestration only: staging, workspace,
// kernel launches and timing.  There is NO CPU fallback: every entry point that computes
// requires a CUDA device and returns the CUDA error otherwise.
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/dgibbs.h"
#include "common.cuh"
#include "internal.h"

using namespace dg;

struct Buf {
  void* p = nullptr;
  size_t cap = 0;
};

struct dgibbs_handle_s {
  int device = 0;
  int sm_count = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  std::string err;
  dgibbs_timing timing{};
  std::map<std::string, Buf> pool;  // grow-only device workspace, freed in destroy
  std::map<std::string, long long> tags;
  cudaEvent_t ev[8]{};
  // pipelined host-pointer path: uploads / downloads on their own streams, two buffer sets
  cudaStream_t up_stream = nullptr, down_stream = nullptr;
  cudaEvent_t up_done[2]{}, down_done[2]{}, up_t0 = nullptr, up_t1 = nullptr;
  // cached projection table
  int ft_N = -1, ft_cplx = -1, tw_N = -1;
};

#define CU(call)                                                                      \
  do {                                                                                \
    cudaError_t e_ = (call);                                                          \
    if (e_ != cudaSuccess) {                                                          \
      char msg_[512];                                                                 \
      snprintf(msg_, sizeof msg_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      h->err = msg_;                                                                  \
      return (int)e_;                                                                 \
    }                                                                                 \
  } while (0)

static int bad_arg(dgibbs_handle h, int idx, const char* what) {
  if (h) h->err = std::string("bad argument: ") + what;
  return -idx;
}

static cudaError_t pool_get(dgibbs_handle h, const char* name, size_t bytes, void** out) {
  Buf& b = h->pool[name];
  if (bytes > b.cap) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
    size_t want = bytes + (bytes >> 4) + 256;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) {
      want = bytes;
      e = cudaMalloc(&b.p, want);
      if (e != cudaSuccess) return e;
    }
    b.cap = want;
  }
  *out = b.p;
  return cudaSuccess;
}

cudaError_t dg_pool_get(dgibbs_handle h, const char* name, size_t bytes, void** out) { return pool_get(h, name, bytes, out); }
cudaStream_t dg_stream(dgibbs_handle h) { return h->stream; }
int dg_device(dgibbs_handle h) { return h->device; }
int dg_sm_count(dgibbs_handle h) { return h->sm_count; }
void dg_set_error(dgibbs_handle h, const char* msg) { h->err = msg; }
long long* dg_tag(dgibbs_handle h, const char* name) { return &h->tags[name]; }
cudaEvent_t dg_event(dgibbs_handle h, int i) { return h->ev[i & 7]; }

extern "C" int dgibbs_version(void) { return DGIBBS_VERSION; }

extern "C" const char* dgibbs_rng_info(void) {
#if DGIBBS_FAST_NORMALS
  return "device Philox4x32-10 keyed on (seed, global baseline id), counter (element, iteration, stream); N(0,1): fp32 Box-Muller "
         "widened to fp64 (-DDGIBBS_FAST_NORMALS); chi^2_T: -2 sum ln u, fp64; diagonal branch: fp32 Marsaglia-Tsang chi^2";
#else
  return "device Philox4x32-10 keyed on (seed, global baseline id), counter (element, iteration, stream); N(0,1): fp64 Box-Muller "
         "(53-bit uniforms, fp64 log / sincospi); chi^2_T: -2 sum ln u, fp64; diagonal branch: collapsed-moment draws, fp32 "
         "Marsaglia-Tsang chi^2";
#endif
}

extern "C" int dgibbs_sizeof(int which) {
  switch (which) {
    case 0: return (int)sizeof(dgibbs_problem);
    case 1: return (int)sizeof(dgibbs_result);
    case 2: return (int)sizeof(dgibbs_timing);
    case 3: return (int)sizeof(dgibbs_cross_problem);
    case 4: return (int)sizeof(dgibbs_cut);
    default: return -1;
  }
}

extern "C" int dgibbs_create(int device_id, uint64_t stream, dgibbs_handle* out) {
  if (!out) return -3;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess) return (int)e;
  if (device_id < 0 || device_id >= ndev) return -1;
  e = cudaSetDevice(device_id);
  if (e != cudaSuccess) return (int)e;
  dgibbs_handle h = new dgibbs_handle_s();
  h->device = device_id;
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, device_id);
  if (e != cudaSuccess) { delete h; return (int)e; }
  h->sm_count = prop.multiProcessorCount;
  if (stream) {
    h->stream = (cudaStream_t)(uintptr_t)stream;
  } else {
    e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete h; return (int)e; }
    h->own_stream = true;
  }
  for (auto& ev : h->ev) cudaEventCreate(&ev);
  cudaStreamCreateWithFlags(&h->up_stream, cudaStreamNonBlocking);
  cudaStreamCreateWithFlags(&h->down_stream, cudaStreamNonBlocking);
  for (int i = 0; i < 2; i++) {
    cudaEventCreateWithFlags(&h->up_done[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->down_done[i], cudaEventDisableTiming);
  }
  cudaEventCreate(&h->up_t0);
  cudaEventCreate(&h->up_t1);
  *out = h;
  return 0;
}

extern "C" int dgibbs_destroy(dgibbs_handle h) {
  if (!h) return -1;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  for (auto& kv : h->pool)
    if (kv.second.p) cudaFree(kv.second.p);
  for (auto& ev : h->ev) cudaEventDestroy(ev);
  for (int i = 0; i < 2; i++) {
    cudaEventDestroy(h->up_done[i]);
    cudaEventDestroy(h->down_done[i]);
  }
  cudaEventDestroy(h->up_t0);
  cudaEventDestroy(h->up_t1);
  cudaStreamSynchronize(h->up_stream);
  cudaStreamSynchronize(h->down_stream);
  cudaStreamDestroy(h->up_stream);
  cudaStreamDestroy(h->down_stream);
  if (h->own_stream) cudaStreamDestroy(h->stream);
  delete h;
  return 0;
}

extern "C" const char* dgibbs_last_error(dgibbs_handle h) { return h ? h->err.c_str() : "null handle"; }

extern "C" int dgibbs_get_timing(dgibbs_handle h, dgibbs_timing* t) {
  if (!h) return -1;
  if (!t) return -2;
  *t = h->timing;
  return 0;
}

// ------------------------------------------------------------------------------ batch staging
struct Staged {
  DevBatch d{};
  std::vector<int> T, M;
  std::vector<long long> data_off, chan_off;
  const double* S0 = nullptr;  // device
  long long sumT = 0;
  std::vector<long long> xoff;
};

static int round_up(int x, int m) { return (x + m - 1) / m * m; }

// Host copies of a batch's small arrays (the pipelined path has them already: no device -> host round trip)
struct HostSmall {
  std::vector<int> T, M;
  std::vector<long long> data_off, chan_off;
};

template <typename Tp>
static cudaError_t fetch(std::vector<Tp>& dst, const Tp* src, size_t n, bool on_device, cudaStream_t st) {
  dst.resize(n);
  if (n == 0) return cudaSuccess;
  if (on_device) {
    cudaError_t e = cudaMemcpyAsync(dst.data(), src, n * sizeof(Tp), cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess) return e;
    return cudaStreamSynchronize(st);
  }
  memcpy(dst.data(), src, n * sizeof(Tp));
  return cudaSuccess;
}

// copy (or alias) an input array onto the device
template <typename Tp>
static cudaError_t stage_in(dgibbs_handle h, const char* name, const Tp* src, size_t n, bool on_device, const Tp** out) {
  if (!src || n == 0) { *out = nullptr; return cudaSuccess; }
  if (on_device) { *out = src; return cudaSuccess; }
  void* p;
  cudaError_t e = pool_get(h, name, n * sizeof(Tp), &p);
  if (e != cudaSuccess) return e;
  e = cudaMemcpyAsync(p, src, n * sizeof(Tp), cudaMemcpyHostToDevice, h->stream);
  *out = (const Tp*)p;
  return e;
}

static int check_problem(dgibbs_handle h, const dgibbs_problem* p) {
  if (!h) return -1;
  if (!p) return bad_arg(h, 2, "problem is NULL");
  if (p->struct_size != (int32_t)sizeof(dgibbs_problem)) return bad_arg(h, 2, "problem.struct_size mismatch");
  if (p->B <= 0) return bad_arg(h, 2, "B must be positive");
  if (p->N < 2 || (!p->complex_timedomain && (p->N & 1))) return bad_arg(h, 2, "N must be even and >= 2");
  if (p->niter < 0) return bad_arg(h, 2, "niter must be >= 0");
  if (!window_id_valid(p->window_id)) return bad_arg(h, 2, "unknown window_id");
  if (p->rng_mode < 0 || p->rng_mode > 2) return bad_arg(h, 2, "unknown rng_mode");
  if (p->cluster_size != 0 && p->cluster_size != 1 && p->cluster_size != 2 && p->cluster_size != 4)
    return bad_arg(h, 2, "cluster_size must be 0 (auto), 1, 2 or 4");
  if (!p->T || !p->M || !p->data_off || !p->chan_off || !p->data || !p->Ni || !p->fsel || !p->S0)
    return bad_arg(h, 2, "a required input pointer is NULL");
  if (p->rng_mode == DGIBBS_RNG_INJECTED && p->niter > 0 && (!p->w1 || !p->w2 || !p->chi2 || !p->w1_off || !p->w2_off))
    return bad_arg(h, 2, "injected rng mode needs w1, w2, chi2 and their offsets");
  return 0;
}

// Builds the device-side batch: geometry, staged inputs, derived tables (setup kernels enqueued).
static int stage_batch(dgibbs_handle h, const dgibbs_problem* p, Staged& s, bool wiener, const HostSmall* hs = nullptr) {
  DevBatch& d = s.d;
  const bool dev = p->on_device != 0;
  const int B = p->B;
  CU(cudaSetDevice(h->device));
  if (hs) {
    s.T = hs->T; s.M = hs->M; s.data_off = hs->data_off; s.chan_off = hs->chan_off;
  } else {
    CU(fetch(s.T, p->T, B, dev, h->stream));
    CU(fetch(s.M, p->M, B, dev, h->stream));
    CU(fetch(s.data_off, (const long long*)p->data_off, (size_t)B + 1, dev, h->stream));
    CU(fetch(s.chan_off, (const long long*)p->chan_off, (size_t)B + 1, dev, h->stream));
  }

  d.B = B;
  d.N = p->N;
  d.cplx = p->complex_timedomain ? 1 : 0;
  d.nsys = d.cplx ? 2 * d.N : d.N;
  d.npad = round_up(d.nsys, TILE);
  d.nblk = d.npad / TILE;
  d.nch = d.cplx ? d.N : d.N / 2 + 1;
  d.K = 2 * d.nch;
  d.Kpad = round_up(d.K, KC);
  d.niter = wiener ? 1 : p->niter;
  d.window_id = p->window_id;
  d.rng_mode = wiener ? DGIBBS_RNG_ZERO : p->rng_mode;
  d.seed = p->seed;
  d.wiener = wiener ? 1 : 0;
  int Tmax = 1;
  s.sumT = 0;
  s.xoff.resize(B);
  for (int b = 0; b < B; b++) {
    if (s.T[b] <= 0) return bad_arg(h, 2, "T[b] must be positive");
    if (s.M[b] <= 0 || s.M[b] > d.nch) return bad_arg(h, 2, "M[b] out of range");
    // blocks may leave gaps (the device-side cut writes into fixed slots) but must not overlap
    if (s.chan_off[b] < 0 || s.chan_off[b + 1] - s.chan_off[b] < s.M[b]) return bad_arg(h, 2, "chan_off inconsistent with M");
    if (s.data_off[b] < 0 || s.data_off[b + 1] - s.data_off[b] < (long long)s.T[b] * s.M[b])
      return bad_arg(h, 2, "data_off inconsistent with T*M");
    Tmax = s.T[b] > Tmax ? s.T[b] : Tmax;
    s.xoff[b] = s.sumT;
    s.sumT += s.T[b];
  }
  d.Tmax = Tmax;
  d.RT = (Tmax + TILE - 1) / TILE;
  d.nslots = B < h->sm_count ? B : h->sm_count;
  d.ctab_smem = 1;
  if (dense_smem_bytes(d) > 227 * 1024) d.ctab_smem = 0;

  // --- inputs
  const long long nchan = s.chan_off[B], ndata = s.data_off[B];
  CU(stage_in(h, "in_T", (const int*)p->T, B, dev, &d.T));
  CU(stage_in(h, "in_M", (const int*)p->M, B, dev, &d.M));
  CU(stage_in(h, "in_doff", (const long long*)p->data_off, (size_t)B + 1, dev, &d.data_off));
  CU(stage_in(h, "in_coff", (const long long*)p->chan_off, (size_t)B + 1, dev, &d.chan_off));
  d.data_c64 = p->data_c64 ? 1 : 0;
  {
    const char* dptr = nullptr;  // complex128 (16 B) or complex64 (8 B) elements
    CU(stage_in(h, "in_data", (const char*)p->data, (size_t)ndata * (d.data_c64 ? 8 : 16), dev, &dptr));
    d.data = (const double2*)dptr;
  }
  CU(stage_in(h, "in_Ni", p->Ni, (size_t)nchan, dev, &d.Ni));
  CU(stage_in(h, "in_fsel", (const long long*)p->fsel, (size_t)nchan, dev, &d.fsel));
  CU(stage_in(h, "in_S0", p->S0, (size_t)B * d.N, dev, &s.S0));
  CU(stage_in(h, "in_gid", (const long long*)p->global_id, B, dev, &d.gid));
  if (d.rng_mode == DGIBBS_RNG_INJECTED && d.niter > 0) {
    std::vector<long long> w1o, w2o;
    CU(fetch(w1o, (const long long*)p->w1_off, (size_t)B + 1, dev, h->stream));
    CU(fetch(w2o, (const long long*)p->w2_off, (size_t)B + 1, dev, h->stream));
    CU(stage_in(h, "in_w1off", (const long long*)p->w1_off, (size_t)B + 1, dev, &d.w1_off));
    CU(stage_in(h, "in_w2off", (const long long*)p->w2_off, (size_t)B + 1, dev, &d.w2_off));
    CU(stage_in(h, "in_w1", p->w1, (size_t)w1o[B], dev, &d.w1));
    CU(stage_in(h, "in_w2", p->w2, (size_t)w2o[B], dev, &d.w2));
    CU(stage_in(h, "in_chi2", p->chi2, (size_t)B * d.niter * d.N, dev, &d.chi2));
  }
  if (wiener) {
    void* px;
    CU(pool_get(h, "xoff", sizeof(long long) * B, &px));
    CU(cudaMemcpyAsync(px, s.xoff.data(), sizeof(long long) * B, cudaMemcpyHostToDevice, h->stream));
    d.xoff = (const long long*)px;
  }
  CU(cudaEventRecord(h->ev[1], h->stream));

  // --- derived tables and workspace
  void* q;
  CU(pool_get(h, "cidx", sizeof(int) * (size_t)B * d.nch, &q)); d.cidx = (int*)q;
  CU(pool_get(h, "h", sizeof(double) * (size_t)B * d.Kpad, &q)); d.h = (double*)q;
  CU(pool_get(h, "zT", sizeof(double) * (size_t)B * d.Tmax * d.Kpad, &q)); d.zT = (double*)q;
  CU(pool_get(h, "cident", (size_t)B, &q)); d.cident = (unsigned char*)q;
  CU(pool_get(h, "wsre", sizeof(double2) * (size_t)B * (d.Kpad / 2), &q)); d.wsre = (double2*)q;
  CU(pool_get(h, "ctab", sizeof(double) * (size_t)B * d.nsys, &q)); d.ctab = (double*)q;
  CU(pool_get(h, "S", sizeof(double) * (size_t)B * d.N, &q)); d.S = (double*)q;
  CU(pool_get(h, "n_done", sizeof(int) * (size_t)B, &q)); d.n_done = (int*)q;
  CU(pool_get(h, "failed", sizeof(int) * (size_t)B, &q)); d.failed = (int*)q;
  CU(pool_get(h, "branch", (size_t)B, &q)); d.branch = (unsigned char*)q;
  CU(pool_get(h, "diag_ab", sizeof(double) * 2 * (size_t)B, &q)); d.diag_ab = (double*)q;
  d.force_branch = p->force_branch;
  const size_t wrows = (size_t)d.npad + (size_t)d.RT * TILE;
  d.work_stride = wrows * d.npad + (size_t)d.N * d.Tmax;
  CU(pool_get(h, "work", sizeof(double) * (size_t)d.nslots * d.work_stride, &q)); d.work = (double*)q;
  // Wp and Si are indexed per CTA (a chain may be shared by a cluster of CTAs, each with its own copy)
  CU(pool_get(h, "Wp", sizeof(double) * (size_t)h->sm_count * dense_wp_doubles_per_slot(d), &q)); d.Wp = (double*)q;
  CU(pool_get(h, "uT", sizeof(double) * (size_t)d.nslots * d.RT * TILE * d.Kpad, &q)); d.uT = (double*)q;
  CU(pool_get(h, "Si", sizeof(double) * (size_t)h->sm_count * d.npad, &q)); d.Si = (double*)q;
  {
    size_t ftbytes = sizeof(double) * (size_t)d.Kpad * d.npad;
    bool fresh = h->pool["Ftab"].cap < ftbytes || h->ft_N != d.N || h->ft_cplx != d.cplx;
    CU(pool_get(h, "Ftab", ftbytes, &q));
    d.Ftab = (const double*)q;
    if (fresh) {
      CU(launch_build_ftab(d, (double*)q, h->stream));
      h->timing.launches++;
      h->ft_N = d.N;
      h->ft_cplx = d.cplx;
    }
  }
  // in-smem FFT for the masked inverse DFT when N is a power of two that fits the pipeline buffers
  d.use_fft = (d.N >= 8 && (d.N & (d.N - 1)) == 0 && d.N <= 4096 && !getenv("DGIBBS_NO_FFT")) ? 1 : 0;
  d.twtab = nullptr;
  d.stwtab = nullptr;
  if (d.use_fft) {
    bool fresh = h->tw_N != d.N;
    CU(pool_get(h, "twtab", sizeof(double2) * (size_t)(d.N / 2), &q));
    d.twtab = (const double2*)q;
    void* q2;
    CU(pool_get(h, "stwtab", sizeof(double2) * 2 * (size_t)d.N, &q2));
    d.stwtab = (const double2*)q2;
    if (fresh) {
      CU(launch_build_twtab(d.N, (double2*)q, h->stream));
      CU(launch_build_stage_twiddles(d.N, (double2*)q2, h->stream));
      h->timing.launches += 2;
      h->tw_N = d.N;
    }
  }
  // diagonal-branch baselines that will be projected by diag_project_fft_kernel (same conditions as in dgibbs_run)
  d.skip_zT_diag = (d.rng_mode == DGIBBS_RNG_PHILOX && d.nsys <= 4096 && d.niter > 0 && !d.wiener && diag_project_fft_fits(d) &&
                    !getenv("DGIBBS_NO_PROJECT_KERNEL")) ? 1 : 0;
  CU(pool_get(h, "setup_err", sizeof(int), &q)); d.setup_err = (int*)q;
  CU(cudaMemsetAsync(d.setup_err, 0, sizeof(int), h->stream));
  CU(launch_setup_baselines(d, s.S0, h->stream));
  h->timing.launches++;
  {
    // channel_ind out of range or aliased (the reference would sum the aliased channels; here it is an error, never a race)
    int err = 0;
    CU(cudaMemcpyAsync(&err, d.setup_err, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (err & 1) return bad_arg(h, 2, d.cplx ? "an fsel entry lies outside [0, N)" : "an fsel entry lies outside [0, N/2]");
    if (err & 2) return bad_arg(h, 2, "fsel holds the same channel twice for one baseline (check freq_zero / freq_spacing / nfreq)");
  }
  CU(dense_configure(d));
  return 0;
}

// CTAs per chain of the dense kernel: the problem's cluster_size, else DGIBBS_DENSE_CLUSTER, else 0 (planned per launch)
static int cluster_request(const dgibbs_problem* p) {
  if (p->cluster_size > 0) return p->cluster_size;
  const char* e = getenv("DGIBBS_DENSE_CLUSTER");
  return e ? atoi(e) : 0;
}

static double dense_flop_per_iter(const DevBatch& d, int T, int M) {
  // W_f = n^3/3 + 2 n^2 T + 4 n M T  (n = system size)
  double n = d.nsys;
  return n * n * n / 3.0 + 2.0 * n * n * T + 4.0 * n * M * (double)T;
}

static double freq_flop_per_iter(const DevBatch& d, int T, int M) {
  // W_t = (2M)^3/3 + 2 (2M)^2 T + 2 (2M)^2 N + 8 N M T  (the reference's arithmetic for this branch;
  // the device generates R Rt + I from one FFT of S instead of the 2 (2M)^2 N GEMM)
  double m2 = 2.0 * M, n = d.N;
  return m2 * m2 * m2 / 3.0 + 2.0 * m2 * m2 * T + 2.0 * m2 * m2 * n + 8.0 * n * M * (double)T;
}

extern "C" int dgibbs_workspace_bytes(dgibbs_handle h, const dgibbs_problem* p, int64_t* bytes) {
  int rc = check_problem(h, p);
  if (rc) return rc;
  if (!bytes) return -3;
  if (p->on_device) return bad_arg(h, 2, "workspace query needs host-side T/M arrays");
  int cplx = p->complex_timedomain ? 1 : 0;
  size_t nsys = cplx ? 2 * (size_t)p->N : p->N, npad = round_up((int)nsys, TILE);
  size_t nch = cplx ? p->N : p->N / 2 + 1, Kpad = round_up((int)(2 * nch), KC);
  size_t Tmax = 1;
  for (int b = 0; b < p->B; b++) Tmax = (size_t)p->T[b] > Tmax ? p->T[b] : Tmax;
  size_t RT = (Tmax + TILE - 1) / TILE, B = p->B;
  size_t nslots = B < (size_t)h->sm_count ? B : h->sm_count;
  size_t tot = 8 * (B * Kpad + B * Tmax * Kpad + B * nsys + 2 * B * p->N + Kpad * npad) + 4 * B * nch +
               8 * nslots * ((npad + RT * TILE) * npad + (size_t)p->N * Tmax + RT * TILE * Kpad + npad) +
               8 * (size_t)p->niter * B * p->N + 8 * B * p->N;
  *bytes = (int64_t)(tot + tot / 16);
  return 0;
}

// One batch, start to finish, on the handle's stream (synchronous).  `hs`: host copies of T / M / offsets when the
// caller has them (device-resident problem built by the pipelined path).
static int run_core(dgibbs_handle h, const dgibbs_problem* p, dgibbs_result* r, const HostSmall* hs) {
  int rc;
  const bool dev = p->on_device != 0;
  h->timing = dgibbs_timing{};
  CU(cudaSetDevice(h->device));
  CU(cudaEventRecord(h->ev[0], h->stream));

  Staged s;
  rc = stage_batch(h, p, s, false, hs);
  if (rc) return rc;
  DevBatch& d = s.d;
  const size_t BN = (size_t)d.B * d.N;

  // outputs: write straight into the caller's buffers when they are device buffers
  void* q;
  double* d_spec; double* d_samples; int* d_ndone; unsigned char* d_succ; double* d_final = nullptr;
  if (dev) {
    d_spec = r->spectrum; d_ndone = r->n_done; d_succ = r->success; d_final = r->final_S;
  } else {
    CU(pool_get(h, "out_spec", sizeof(double) * BN, &q)); d_spec = (double*)q;
    CU(pool_get(h, "out_ndone", sizeof(int) * d.B, &q)); d_ndone = (int*)q;
    CU(pool_get(h, "out_succ", d.B, &q)); d_succ = (unsigned char*)q;
    if (r->final_S) { CU(pool_get(h, "out_final", sizeof(double) * BN, &q)); d_final = (double*)q; }
  }
  if (dev && r->samples) {
    d_samples = r->samples;
  } else {
    CU(pool_get(h, "samples", sizeof(double) * (size_t)(d.niter > 0 ? d.niter : 1) * BN, &q));
    d_samples = (double*)q;
  }
  d.samples = d_samples;
  CU(cudaEventRecord(h->ev[2], h->stream));

  if (getenv("DGIBBS_PROFILE")) {
    CU(pool_get(h, "prof", 16 * sizeof(long long), &q));
    CU(cudaMemsetAsync(q, 0, 16 * sizeof(long long), h->stream));
    d.prof = (long long*)q;
  }
  // --- classify: which baselines have a diagonal(+rank-1) system (decided by the setup kernel)
  std::vector<unsigned char> br(d.B);
  CU(cudaMemcpyAsync(br.data(), d.branch, d.B, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  const bool fast_diag = d.rng_mode == DGIBBS_RNG_PHILOX && d.nsys <= 4096 && d.niter > 0;
  // frequency-space branch (_draw_signal_sample_t): the reference's rule len(fsel) <= 0.25 N (delay.py:1884-1886)
  std::vector<int> dense_list, diag_list, freq_list, ycompact(d.B, -1);
  for (int b = 0; b < d.B; b++) {
    const bool few = (double)s.M[b] <= 0.25 * (double)d.N;
    const bool want_t = p->force_branch == DGIBBS_BRANCH_FREQ_SPACE || (p->force_branch == 0 && few);
    if (br[b] == DGIBBS_BRANCH_DIAGONAL && fast_diag) {
      ycompact[b] = (int)diag_list.size();
      diag_list.push_back(b);
    } else if (br[b] != DGIBBS_BRANCH_DIAGONAL && want_t && d.niter > 0 && freq_branch_fits(d, s.M[b], s.T[b]) &&
               !getenv("DGIBBS_NO_FREQ_BRANCH")) {
      br[b] = DGIBBS_BRANCH_FREQ_SPACE;
      freq_list.push_back(b);
    } else {
      dense_list.push_back(b);
    }
  }
  int *d_dense = nullptr, *d_diag = nullptr, *d_freq = nullptr, *d_yc = nullptr;
  const bool split = !diag_list.empty() || !freq_list.empty();
  if (split) {
    CU(pool_get(h, "list_dense", sizeof(int) * (dense_list.size() + 1), &q)); d_dense = (int*)q;
    if (!dense_list.empty())
      CU(cudaMemcpyAsync(d_dense, dense_list.data(), sizeof(int) * dense_list.size(), cudaMemcpyHostToDevice, h->stream));
  }
  if (!freq_list.empty()) {
    CU(pool_get(h, "list_freq", sizeof(int) * freq_list.size(), &q)); d_freq = (int*)q;
    CU(cudaMemcpyAsync(d_freq, freq_list.data(), sizeof(int) * freq_list.size(), cudaMemcpyHostToDevice, h->stream));
  }
  if (!diag_list.empty()) {
    CU(pool_get(h, "list_diag", sizeof(int) * diag_list.size(), &q)); d_diag = (int*)q;
    CU(pool_get(h, "ycompact", sizeof(int) * d.B, &q)); d_yc = (int*)q;
    CU(cudaMemcpyAsync(d_diag, diag_list.data(), sizeof(int) * diag_list.size(), cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemcpyAsync(d_yc, ycompact.data(), sizeof(int) * d.B, cudaMemcpyHostToDevice, h->stream));
    d.ypitch = d.npad + 128;
    CU(pool_get(h, "yhat", sizeof(double) * diag_list.size() * (size_t)d.Tmax * d.ypitch, &q));
    d.yhat = (double*)q;
    d.ycompact = d_yc;
  }
  h->timing.n_dense = (int64_t)dense_list.size();
  h->timing.n_diag = (int64_t)diag_list.size();
  CU(cudaEventRecord(h->ev[5], h->stream));
  if (!diag_list.empty()) {
    // projection yhat = F^T N^-1/2 z with the dense kernel's phase 0 (zero draws), then the streaming chains
    DevBatch dp = d;
    dp.blist = d_diag;
    dp.nlist = (int)diag_list.size();
    dp.project_only = 1;
    dp.rng_mode = DGIBBS_RNG_ZERO;
    dp.nslots = dp.nlist < d.nslots ? dp.nlist : d.nslots;
    static const bool no_pk = getenv("DGIBBS_NO_PROJECT_KERNEL") != nullptr;
    if (!no_pk && diag_project_fft_fits(dp)) CU(launch_diag_project_fft(dp, h->sm_count, h->stream));
    else CU(launch_dense_sampler(dp, 0, 1, h->stream, h->sm_count, 1));
    h->timing.launches++;
    CU(cudaEventRecord(h->ev[6], h->stream));
    DevBatch dd = d;
    dd.blist = d_diag;
    dd.nlist = (int)diag_list.size();
    // b > 0 never comes out of the reference's Fourier matrices; if it does, use the streaming kernel
    std::vector<double> ab(2 * (size_t)d.B);
    CU(cudaMemcpyAsync(ab.data(), d.diag_ab, sizeof(double) * ab.size(), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    bool resident_ok = !getenv("DGIBBS_DIAG_STREAMING");
    for (int b : diag_list) resident_ok = resident_ok && !(ab[2 * (size_t)b + 1] > 0.0);
    CU(launch_diag_sampler(dd, 0, d.niter, h->sm_count, h->stream, resident_ok));
    h->timing.launches++;
    h->timing.sampler_launches++;
  } else {
    CU(cudaEventRecord(h->ev[6], h->stream));
  }
  CU(cudaEventRecord(h->ev[7], h->stream));

  // --- the dense chains: one persistent launch covers every iteration of every baseline
  double flop = 0.0;
  if (d.niter > 0 && !dense_list.empty()) {
    DevBatch dn = d;
    if (split) {
      dn.blist = d_dense;
      dn.nlist = (int)dense_list.size();
      dn.nslots = dn.nlist < d.nslots ? dn.nlist : d.nslots;
    }
    int nl = 0;
    CU(launch_dense_sampler(dn, 0, d.niter, h->stream, h->sm_count, cluster_request(p), &nl));
    h->timing.launches += nl;
    h->timing.sampler_launches += nl;
  }
  if (d.niter > 0 && !freq_list.empty()) {
    DevBatch df = d;
    df.blist = d_freq;
    df.nlist = (int)freq_list.size();
    // one tensor map serves the list when every chain pads its 2M system to the same size
    int npt = 0;
    for (int b : freq_list) {
      const int n = (2 * s.M[b] + TILE - 1) / TILE * TILE;
      npt = (npt == 0 || npt == n) ? n : -1;
      if (npt < 0) break;
    }
    CU(launch_freq_sampler(df, 0, d.niter, h->stream, npt > 0 ? npt : 0));
    h->timing.launches++;
    h->timing.sampler_launches++;
  }
  CU(cudaEventRecord(h->ev[3], h->stream));

  CU(launch_finish(d, s.S0, p->median_frac, d_spec, r->samples ? d_samples : nullptr, d_ndone, d_succ, d_final,
                   h->stream));
  h->timing.launches++;
  if (!dev) {
    CU(cudaMemcpyAsync(r->spectrum, d_spec, sizeof(double) * BN, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(r->n_done, d_ndone, sizeof(int) * d.B, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(r->success, d_succ, d.B, cudaMemcpyDeviceToHost, h->stream));
    if (r->final_S) CU(cudaMemcpyAsync(r->final_S, d_final, sizeof(double) * BN, cudaMemcpyDeviceToHost, h->stream));
    if (r->samples && d.niter > 0)
      CU(cudaMemcpyAsync(r->samples, d_samples, sizeof(double) * (size_t)d.niter * BN, cudaMemcpyDeviceToHost, h->stream));
  }
  CU(cudaEventRecord(h->ev[4], h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (d.prof) {
    long long pc[16];
    CU(cudaMemcpy(pc, d.prof, sizeof pc, cudaMemcpyDeviceToHost));
    if (!diag_list.empty() && dense_list.empty() && freq_list.empty() && pc[8] > 0) {
      static const char* dn[8] = {"P_partial+send", "draws", "wait_allgather", "corr+sync", "moments+draw_S", "stage_D", "sync", "loop_head"};
      for (int i = 0; i < 8; i++) fprintf(stderr, "[dgibbs prof diag] %-16s %8.0f cyc/iteration\n", dn[i], (double)pc[i] / (double)pc[8]);
      for (int i = 0; i < 16; i++) pc[i] = 0;
    }
    static const char* names[12] = {"prepass", "phase0_gemm", "diag_gemm", "diag_stage_factor", "diag_storeW", "row_stage_addG",
                                    "row_gemm", "diag_stage_gen", "row_trimul_store", "back_gemm", "back_trimul", "draw"};
    double tot = 0;
    for (int i = 0; i < 12; i++) tot += (double)pc[i];
    for (int i = 0; i < 12; i++)
      fprintf(stderr, "[dgibbs prof] %-18s %14lld cyc  %5.1f%%\n", names[i], pc[i], 100.0 * pc[i] / (tot > 0 ? tot : 1));
    fprintf(stderr, "[dgibbs prof] inside prepass (fused FFT projection): load %lld  fft %lld  epilogue %lld cyc\n", pc[12], pc[13],
            pc[14]);
  }

  // bookkeeping that needs n_done
  std::vector<int> nd(d.B);
  if (dev) CU(cudaMemcpy(nd.data(), d_ndone, sizeof(int) * d.B, cudaMemcpyDeviceToHost));
  else memcpy(nd.data(), r->n_done, sizeof(int) * d.B);
  double bytes = 0.0;
  for (int b = 0; b < d.B; b++) {
    if (ycompact[b] >= 0) bytes += (double)nd[b] * (8.0 * d.nsys * s.T[b] + 16.0 * d.N);  // Q_d
    else if (br[b] == DGIBBS_BRANCH_FREQ_SPACE) flop += (double)nd[b] * freq_flop_per_iter(d, s.T[b], s.M[b]);
    else flop += (double)nd[b] * dense_flop_per_iter(d, s.T[b], s.M[b]);
  }
  if (r->branch) {
    for (int b = 0; b < d.B; b++)
      if (br[b] != DGIBBS_BRANCH_DIAGONAL && br[b] != DGIBBS_BRANCH_FREQ_SPACE) br[b] = DGIBBS_BRANCH_DELAY_SPACE;
    if (dev) CU(cudaMemcpy(r->branch, br.data(), d.B, cudaMemcpyHostToDevice));
    else memcpy(r->branch, br.data(), d.B);
  }
  float ms;
  cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]); h->timing.ms_h2d = ms;
  cudaEventElapsedTime(&ms, h->ev[1], h->ev[5]); h->timing.ms_setup = ms;
  cudaEventElapsedTime(&ms, h->ev[5], h->ev[6]); h->timing.ms_diag_project = ms;
  cudaEventElapsedTime(&ms, h->ev[6], h->ev[7]); h->timing.ms_sampler_diag = ms;
  cudaEventElapsedTime(&ms, h->ev[7], h->ev[3]); h->timing.ms_sampler = ms;
  cudaEventElapsedTime(&ms, h->ev[3], h->ev[4]); h->timing.ms_finish = ms;
  h->timing.flop_sampler = flop;
  h->timing.bytes_sampler = bytes;
  return 0;
}

// ------------------------------------------------------------------------------ pipelined host-pointer path
// Host inputs of a large batch go up in chunks on their own stream, double-buffered: the upload of chunk c+1 runs
// under the set-up / projection / sampler kernels of chunk c, and the results of chunk c come down on a third stream
// under chunk c+1.  Chunks are multiples of the SM count, so the dense branch keeps full waves.  Every chain is keyed
// on its global baseline id, so the results are those of the single-shot call, bit for bit.
static void add_timing(dgibbs_timing& a, const dgibbs_timing& b) {
  a.ms_setup += b.ms_setup; a.ms_sampler += b.ms_sampler; a.ms_finish += b.ms_finish;
  a.launches += b.launches; a.sampler_launches += b.sampler_launches;
  a.flop_sampler += b.flop_sampler; a.bytes_sampler += b.bytes_sampler;
  a.ms_diag_project += b.ms_diag_project; a.ms_sampler_diag += b.ms_sampler_diag;
  a.n_dense += b.n_dense; a.n_diag += b.n_diag;
}

static int pipeline_chunk(dgibbs_handle h, const dgibbs_problem* p) {
  // baselines per chunk: 0 = never split.  Default 8 waves of the dense kernel (1184 on a B200).
  const char* e = getenv("DGIBBS_PIPELINE_CHUNK");
  int c = e ? atoi(e) : 8 * h->sm_count;
  if (c > 0) c = (c + h->sm_count - 1) / h->sm_count * h->sm_count;
  return c;
}

static int run_pipelined(dgibbs_handle h, const dgibbs_problem* p, dgibbs_result* r, int chunk) {
  const int B = p->B, N = p->N, niter = p->niter;
  const size_t esz = p->data_c64 ? 8 : 16;
  CU(cudaSetDevice(h->device));
  const int nchunks = (B + chunk - 1) / chunk;
  const long long* doff = (const long long*)p->data_off;
  const long long* coff = (const long long*)p->chan_off;
  // capacity of one buffer set
  size_t max_data = 0, max_chan = 0;
  for (int c = 0; c < nchunks; c++) {
    const int lo = c * chunk, hi = lo + chunk < B ? lo + chunk : B;
    if (doff[hi] < doff[lo] || coff[hi] < coff[lo]) return bad_arg(h, 2, "data_off / chan_off must be non-decreasing");
    max_data = std::max(max_data, (size_t)(doff[hi] - doff[lo]));
    max_chan = std::max(max_chan, (size_t)(coff[hi] - coff[lo]));
  }
  const size_t cb = (size_t)chunk;
  struct Set {
    int *T, *M; long long *doff, *coff, *fsel, *gid; char* data; double *Ni, *S0;
    double *spec, *samples, *finalS; int* ndone; unsigned char *succ, *branch;
  } set[2];
  for (int k = 0; k < 2; k++) {
    char nm[32];
    void* q;
#define PGET(field, type, count)                                       \
    snprintf(nm, sizeof nm, "pl%d_" #field, k);                         \
    CU(pool_get(h, nm, sizeof(type) * (size_t)(count), &q));            \
    set[k].field = (type*)q;
    PGET(T, int, cb) PGET(M, int, cb) PGET(doff, long long, cb + 1) PGET(coff, long long, cb + 1)
    PGET(fsel, long long, max_chan) PGET(gid, long long, cb) PGET(data, char, max_data * esz) PGET(Ni, double, max_chan)
    PGET(S0, double, cb * N) PGET(spec, double, cb * N) PGET(ndone, int, cb) PGET(succ, unsigned char, cb)
    PGET(branch, unsigned char, cb)
    set[k].samples = nullptr;
    set[k].finalS = nullptr;
    if (r->samples) { PGET(samples, double, (size_t)(niter > 0 ? niter : 1) * cb * N) }
    if (r->final_S) { PGET(finalS, double, cb * N) }
#undef PGET
  }
  std::vector<HostSmall> hs(2);
  std::vector<long long> gid_host[2];

  auto upload = [&](int c) -> int {
    const int k = c & 1, lo = c * chunk, hi = lo + chunk < B ? lo + chunk : B, nb = hi - lo;
    HostSmall& s = hs[k];
    s.T.assign(p->T + lo, p->T + hi);
    s.M.assign(p->M + lo, p->M + hi);
    s.data_off.resize(nb + 1);
    s.chan_off.resize(nb + 1);
    for (int i = 0; i <= nb; i++) {
      s.data_off[i] = doff[lo + i] - doff[lo];
      s.chan_off[i] = coff[lo + i] - coff[lo];
    }
    gid_host[k].resize(nb);
    for (int i = 0; i < nb; i++) gid_host[k][i] = p->global_id ? p->global_id[lo + i] : (long long)(lo + i);
    cudaStream_t us = h->up_stream;
    CU(cudaMemcpyAsync(set[k].T, s.T.data(), sizeof(int) * nb, cudaMemcpyHostToDevice, us));
    CU(cudaMemcpyAsync(set[k].M, s.M.data(), sizeof(int) * nb, cudaMemcpyHostToDevice, us));
    CU(cudaMemcpyAsync(set[k].doff, s.data_off.data(), sizeof(long long) * (nb + 1), cudaMemcpyHostToDevice, us));
    CU(cudaMemcpyAsync(set[k].coff, s.chan_off.data(), sizeof(long long) * (nb + 1), cudaMemcpyHostToDevice, us));
    CU(cudaMemcpyAsync(set[k].gid, gid_host[k].data(), sizeof(long long) * nb, cudaMemcpyHostToDevice, us));
    CU(cudaMemcpyAsync(set[k].data, (const char*)p->data + (size_t)doff[lo] * esz, (size_t)(doff[hi] - doff[lo]) * esz,
                       cudaMemcpyHostToDevice, us));
    CU(cudaMemcpyAsync(set[k].Ni, p->Ni + coff[lo], sizeof(double) * (size_t)(coff[hi] - coff[lo]), cudaMemcpyHostToDevice, us));
    CU(cudaMemcpyAsync(set[k].fsel, p->fsel + coff[lo], sizeof(long long) * (size_t)(coff[hi] - coff[lo]), cudaMemcpyHostToDevice, us));
    CU(cudaMemcpyAsync(set[k].S0, p->S0 + (size_t)lo * N, sizeof(double) * (size_t)nb * N, cudaMemcpyHostToDevice, us));
    CU(cudaEventRecord(h->up_done[k], us));
    return 0;
  };

  dgibbs_timing total{};
  CU(cudaEventRecord(h->up_t0, h->up_stream));
  int rc = upload(0);
  if (rc) return rc;
  for (int c = 0; c < nchunks; c++) {
    const int k = c & 1, lo = c * chunk, hi = lo + chunk < B ? lo + chunk : B, nb = hi - lo;
    // the other input set was last read by chunk c-1, whose run_core has returned (stream synchronised)
    if (c + 1 < nchunks && (rc = upload(c + 1))) return rc;
    if (c + 1 == nchunks) CU(cudaEventRecord(h->up_t1, h->up_stream));
    CU(cudaStreamWaitEvent(h->stream, h->up_done[k], 0));
    if (c >= 2) CU(cudaStreamWaitEvent(h->stream, h->down_done[k], 0));  // result set k drained
    dgibbs_problem pc = *p;
    pc.on_device = 1;
    pc.B = nb;
    pc.T = set[k].T; pc.M = set[k].M;
    pc.data_off = (const int64_t*)set[k].doff; pc.chan_off = (const int64_t*)set[k].coff;
    pc.data = set[k].data; pc.Ni = set[k].Ni; pc.fsel = (const int64_t*)set[k].fsel; pc.S0 = set[k].S0;
    pc.global_id = (const int64_t*)set[k].gid;
    dgibbs_result rcn{};
    rcn.struct_size = (int32_t)sizeof(dgibbs_result);
    rcn.on_device = 1;
    rcn.spectrum = set[k].spec; rcn.samples = set[k].samples; rcn.n_done = set[k].ndone;
    rcn.success = set[k].succ; rcn.branch = r->branch ? set[k].branch : nullptr; rcn.final_S = set[k].finalS;
    rc = run_core(h, &pc, &rcn, &hs[k]);
    if (rc) return rc;
    add_timing(total, h->timing);
    cudaStream_t ds = h->down_stream;  // run_core has synchronised the compute stream: the results are complete
    CU(cudaMemcpyAsync(r->spectrum + (size_t)lo * N, set[k].spec, sizeof(double) * (size_t)nb * N, cudaMemcpyDeviceToHost, ds));
    CU(cudaMemcpyAsync(r->n_done + lo, set[k].ndone, sizeof(int) * nb, cudaMemcpyDeviceToHost, ds));
    CU(cudaMemcpyAsync(r->success + lo, set[k].succ, nb, cudaMemcpyDeviceToHost, ds));
    if (r->branch) CU(cudaMemcpyAsync(r->branch + lo, set[k].branch, nb, cudaMemcpyDeviceToHost, ds));
    if (r->final_S)
      CU(cudaMemcpyAsync(r->final_S + (size_t)lo * N, set[k].finalS, sizeof(double) * (size_t)nb * N, cudaMemcpyDeviceToHost, ds));
    if (r->samples && niter > 0)  // [niter][B][N] on the host, [niter][nb][N] on the device
      CU(cudaMemcpy2DAsync(r->samples + (size_t)lo * N, sizeof(double) * (size_t)B * N, set[k].samples,
                           sizeof(double) * (size_t)nb * N, sizeof(double) * (size_t)nb * N, niter, cudaMemcpyDeviceToHost, ds));
    CU(cudaEventRecord(h->down_done[k], ds));
  }
  CU(cudaStreamSynchronize(h->down_stream));
  CU(cudaStreamSynchronize(h->up_stream));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, h->up_t0, h->up_t1);
  total.ms_h2d = ms;  // span of the uploads (they overlap the kernels of the previous chunk)
  h->timing = total;
  return 0;
}

extern "C" int dgibbs_run(dgibbs_handle h, const dgibbs_problem* p, dgibbs_result* r) {
  int rc = check_problem(h, p);
  if (rc) return rc;
  if (!r || r->struct_size != (int32_t)sizeof(dgibbs_result)) return bad_arg(h, 3, "result struct");
  if (!r->spectrum || !r->n_done || !r->success) return bad_arg(h, 3, "result.spectrum/n_done/success required");
  if ((r->on_device != 0) != (p->on_device != 0)) return bad_arg(h, 3, "problem and result must live on the same side");
  if (!p->on_device && p->rng_mode != DGIBBS_RNG_INJECTED) {
    const int chunk = pipeline_chunk(h, p);
    if (chunk > 0 && p->B >= 2 * chunk) return run_pipelined(h, p, r, chunk);
  }
  return run_core(h, p, r, nullptr);
}

extern "C" int dgibbs_setup_inputs(dgibbs_handle h, const dgibbs_problem* p, int32_t b, double* data_w,
                                   double* FTNih, double* FTNiF) {
  int rc = check_problem(h, p);
  if (rc) return rc;
  if (b < 0 || b >= p->B) return bad_arg(h, 3, "baseline index out of range");
  h->timing = dgibbs_timing{};
  CU(cudaSetDevice(h->device));
  CU(cudaEventRecord(h->ev[0], h->stream));
  Staged s;
  dgibbs_problem pp = *p;
  pp.rng_mode = DGIBBS_RNG_ZERO;  // draws are irrelevant here
  rc = stage_batch(h, &pp, s, false);
  if (rc) return rc;
  DevBatch& d = s.d;
  const bool dev = p->on_device != 0;
  const size_t n_dw = 2 * (size_t)s.M[b] * s.T[b], n_h = (size_t)d.nsys * 2 * s.M[b], n_f = (size_t)d.nsys * d.nsys;
  double *o_dw = data_w, *o_h = FTNih, *o_f = FTNiF;
  void* q;
  if (!dev) {
    if (data_w) { CU(pool_get(h, "dbg_dw", 8 * n_dw, &q)); o_dw = (double*)q; }
    if (FTNih) { CU(pool_get(h, "dbg_h", 8 * n_h, &q)); o_h = (double*)q; }
    if (FTNiF) { CU(pool_get(h, "dbg_f", 8 * n_f, &q)); o_f = (double*)q; }
  }
  CU(launch_materialize_inputs(d, b, o_dw, o_h, o_f, h->stream));
  if (!dev) {
    if (data_w) CU(cudaMemcpyAsync(data_w, o_dw, 8 * n_dw, cudaMemcpyDeviceToHost, h->stream));
    if (FTNih) CU(cudaMemcpyAsync(FTNih, o_h, 8 * n_h, cudaMemcpyDeviceToHost, h->stream));
    if (FTNiF) CU(cudaMemcpyAsync(FTNiF, o_f, 8 * n_f, cudaMemcpyDeviceToHost, h->stream));
  }
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}

extern "C" int dgibbs_wiener(dgibbs_handle h, const dgibbs_problem* p, double* out, uint8_t* success) {
  int rc = check_problem(h, p);
  if (rc) return rc;
  if (!out) return bad_arg(h, 3, "out is NULL");
  h->timing = dgibbs_timing{};
  CU(cudaSetDevice(h->device));
  CU(cudaEventRecord(h->ev[0], h->stream));
  Staged s;
  rc = stage_batch(h, p, s, true);
  if (rc) return rc;
  DevBatch& d = s.d;
  const bool dev = p->on_device != 0;
  const size_t nout = (size_t)s.sumT * d.nsys;
  void* q;
  double* d_out = out;
  if (!dev) { CU(pool_get(h, "wiener_out", 8 * nout, &q)); d_out = (double*)q; }
  d.xout = d_out;
  CU(pool_get(h, "samples", sizeof(double) * (size_t)d.B * d.N, &q));
  d.samples = (double*)q;
  CU(cudaEventRecord(h->ev[2], h->stream));
  int nl = 0;
  CU(launch_dense_sampler(d, 0, 1, h->stream, h->sm_count, cluster_request(p), &nl));
  h->timing.launches += nl;
  h->timing.sampler_launches += nl;
  CU(cudaEventRecord(h->ev[3], h->stream));
  if (!dev) CU(cudaMemcpyAsync(out, d_out, 8 * nout, cudaMemcpyDeviceToHost, h->stream));
  std::vector<int> failed(d.B);
  CU(cudaMemcpyAsync(failed.data(), d.failed, sizeof(int) * d.B, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaEventRecord(h->ev[4], h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (success) {
    std::vector<unsigned char> ok(d.B);
    for (int b = 0; b < d.B; b++) ok[b] = failed[b] ? 0 : 1;
    if (dev) CU(cudaMemcpy(success, ok.data(), d.B, cudaMemcpyHostToDevice));
    else memcpy(success, ok.data(), d.B);
  }
  float ms;
  cudaEventElapsedTime(&ms, h->ev[2], h->ev[3]); h->timing.ms_sampler = ms;
  return 0;
}

extern "C" int dgibbs_fourier_matrix(dgibbs_handle h, int32_t kind, int32_t N, const int64_t* fsel, int32_t M,
                                     double* out) {
  if (!h) return -1;
  if (kind < 0 || kind > 3) return bad_arg(h, 2, "kind");
  if (N < 1 || M < 1 || !fsel || !out) return bad_arg(h, 3, "N/M/fsel/out");
  CU(cudaSetDevice(h->device));
  size_t n = (kind == 2 ? 4 : 2) * (size_t)M * N;
  void *dsel, *dout;
  CU(pool_get(h, "fm_sel", sizeof(long long) * M, &dsel));
  CU(pool_get(h, "fm_out", 8 * n, &dout));
  CU(cudaMemcpyAsync(dsel, fsel, sizeof(long long) * M, cudaMemcpyHostToDevice, h->stream));
  CU(launch_fourier_matrix(kind, N, (const long long*)dsel, M, (double*)dout, h->stream));
  CU(cudaMemcpyAsync(out, dout, 8 * n, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}
