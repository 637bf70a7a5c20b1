/*
 * dgibbs.h -- C ABI of the B200-native delay power-spectrum Gibbs sampler.
 *
 * The reference (radiocosmology/draco) has no FFI: the path is Python calling
 * numpy / scipy.linalg.  The entry points below are what a ctypes binding inside
 * draco's own task would call instead of the per-baseline Python loop
 *   DelayPowerSpectrumBase._evaluate          draco/analysis/delay.py:1117-1185
 *     -> DelayPowerSpectrumGibbs._estimator   draco/analysis/delay.py:1242-1267
 *        -> delay_power_spectrum_gibbs        draco/analysis/delay.py:1713-1904
 *           -> _compute_delay_spectrum_inputs draco/analysis/delay.py:1657-1710
 *              fourier_matrix_r2c / _c2c      draco/analysis/delay.py:1480-1585
 *              tools.window_generalised       draco/util/tools.py:547-601
 *   delay_spectrum_wiener_filter              draco/analysis/delay.py:2132-2201
 * (see INTEGRATION.md for the binding a draco maintainer would add).
 *
 * Conventions
 *   - plain pointers and sizes only; every buffer is allocated and owned by the caller;
 *     nothing is retained after a call returns.  `on_device` says whether the pointers of
 *     a problem/result are CUDA device pointers (on the handle's device) or host pointers.
 *   - return value: 0 ok; <0 = -(index of the offending argument/field); >0 = a CUDA
 *     error code (text via dgibbs_last_error).  A numerically non-positive-definite
 *     system is NOT an error: it is reported per baseline through success[]/n_done[],
 *     reproducing delay.py:1891-1896 and 1261-1265.
 *   - one handle per process/rank, bound to one CUDA device; calls on one handle are
 *     synchronous with respect to the handle's stream and not re-entrant.
 */
#ifndef DGIBBS_H
#define DGIBBS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DGIBBS_VERSION 100 /* 0.1.0 */

/* window_id: index into the cosine-sum table of draco/util/tools.py:584-595, DGIBBS_WINDOW_TRIANGULAR
 * (tools.py:566-567) or DGIBBS_WINDOW_TUKEY(r) for "tukey-r" (tools.py:569-581; r = the tapered fraction in units of
 * 1e-6, 0 .. 1000000: "tukey-0.5" is DGIBBS_WINDOW_TUKEY(500000)) */
#define DGIBBS_WINDOW_TUKEY_BASE (1 << 20)
#define DGIBBS_WINDOW_TUKEY(r_millionths) (DGIBBS_WINDOW_TUKEY_BASE + (r_millionths))
enum {
  DGIBBS_WINDOW_NONE = -1,
  DGIBBS_WINDOW_UNIFORM = 0,
  DGIBBS_WINDOW_HANN = 1,
  DGIBBS_WINDOW_HANNING = 2,
  DGIBBS_WINDOW_HAMMING = 3,
  DGIBBS_WINDOW_BLACKMAN = 4,
  DGIBBS_WINDOW_NUTTALL = 5,
  DGIBBS_WINDOW_BLACKMAN_NUTTALL = 6,
  DGIBBS_WINDOW_BLACKMAN_HARRIS = 7,
  DGIBBS_WINDOW_TRIANGULAR = 8
};

/* rng_mode */
enum {
  DGIBBS_RNG_INJECTED = 0, /* caller supplies w1, w2, chi2 in the reference's draw order */
  DGIBBS_RNG_PHILOX = 1,   /* device Philox4x32-10 keyed on (seed, global baseline id, iteration) */
  DGIBBS_RNG_ZERO = 2      /* no perturbation: the signal draw becomes the Wiener mean */
};

/* branch[] values (which solver ran for a baseline) */
enum {
  DGIBBS_BRANCH_SKIPPED = 0,
  DGIBBS_BRANCH_DELAY_SPACE = 1, /* _draw_signal_sample_f, n_sys x n_sys system, delay.py:1778-1819 */
  DGIBBS_BRANCH_FREQ_SPACE = 2,  /* _draw_signal_sample_t, 2M x 2M system,     delay.py:1821-1860 */
  DGIBBS_BRANCH_DIAGONAL = 3     /* diagonal(+rank-1) special case of the _f system */
};

typedef struct dgibbs_handle_s* dgibbs_handle;

/*
 * One batch of baselines sharing N (number of delays), window and mode.
 * Ragged per-baseline shapes are described by offset arrays, i.e. each baseline is stored
 * exactly as DelayTransformBase._cut_data (delay.py:516-597) returns it:
 *   data  : complex128 (complex64 with data_c64), baseline b occupies data[data_off[b] .. data_off[b+1]) (complex
 *           elements), row-major (T[b], M[b])
 *   Ni    : float64 inverse noise variance, baseline b at Ni[chan_off[b] .. chan_off[b+1])
 *   fsel  : int64 channel indices (values in [0, N/2] real mode, [0, N) complex mode),
 *           same offsets as Ni
 */
typedef struct {
  int32_t struct_size; /* sizeof(dgibbs_problem), for forward compatibility */
  int32_t on_device;   /* 1: array pointers below are device pointers; 0: host pointers */
  int32_t B;           /* baselines in the batch */
  int32_t N;           /* number of delays (even) */
  int32_t niter;       /* Gibbs iterations (nsamp) */
  int32_t window_id;
  int32_t complex_timedomain;
  int32_t rng_mode;
  uint64_t seed;
  double median_frac;         /* spectrum = median of the last int(niter*median_frac) samples */
  const int32_t* T;           /* [B] samples per baseline */
  const int32_t* M;           /* [B] retained channels per baseline */
  const int64_t* data_off;    /* [B+1] start of each (T[b], M[b]) block; gaps allowed, no overlap */
  const int64_t* chan_off;    /* [B+1] */
  const void* data;           /* complex128 */
  const double* Ni;
  const int64_t* fsel;
  const double* S0;           /* [B, N] initial power spectrum (prior) */
  const int64_t* global_id;   /* [B] global baseline index: Philox key, makes results
                                 independent of sharding; may be NULL (= 0..B-1) */
  /* injected draws (rng_mode == DGIBBS_RNG_INJECTED), in the reference's order and shapes:
   *   w1   : baseline b, iteration i at w1[w1_off[b] + i*n_sys*T[b] ...], (n_sys, T[b]) row-major
   *   w2   : baseline b, iteration i at w2[w2_off[b] + i*2*M[b]*T[b] ...], (2*M[b], T[b])
   *   chi2 : [B, niter, N]
   * n_sys = N (real) or 2N (complex_timedomain). */
  const double* w1;
  const int64_t* w1_off; /* [B+1] offsets in doubles */
  const double* w2;
  const int64_t* w2_off; /* [B+1] */
  const double* chi2;
  int32_t force_branch;  /* 0 = choose as the reference does (delay.py:1884-1886); else a
                            DGIBBS_BRANCH_* value to force where mathematically valid */
  int32_t cluster_size;  /* CTAs (one thread-block cluster) that share one chain of the dense branch: 0 = chosen per
                            launch (full waves one CTA per chain, the remainder on clusters so that the last wave
                            fills the GPU); 1, 2, 4 = every chain on clusters of that size.  The samples do not
                            depend on it (bit-identical). */
  int32_t data_c64;      /* element type of `data`: 0 = complex128, 1 = complex64 (what draco's containers store,
                            core/containers.py:503): widened exactly on the device, half the host->device bytes */
  int32_t reserved;      /* keep 0 */
} dgibbs_problem;

typedef struct {
  int32_t struct_size;
  int32_t on_device;
  double* spectrum;   /* [B, N] fftshifted median (delay.py:1256-1260); prior if the chain is empty */
  double* samples;    /* [niter, B, N] unshifted samples, zero where a chain stopped early
                         (front-padded like delay.py:1175-1178); may be NULL */
  int32_t* n_done;    /* [B] completed iterations */
  uint8_t* success;   /* [B] 1 = full chain */
  uint8_t* branch;    /* [B] DGIBBS_BRANCH_* ; may be NULL */
  double* final_S;    /* [B, N] last state of the chain (to continue it); may be NULL */
} dgibbs_result;

/* timings of the last dgibbs_run on a handle, CUDA events on the handle's stream */
typedef struct {
  double ms_h2d;      /* host->device staging (0 when on_device) */
  double ms_setup;    /* window/whiten/projection-table kernels */
  double ms_sampler;  /* all Gibbs iterations of the dense-branch baselines */
  double ms_finish;   /* median + fftshift (+ device->host) */
  int64_t launches;   /* kernels launched by the call */
  int64_t sampler_launches;
  double flop_sampler; /* algorithmic fp64 flop of the sampler launches (SURVEY 8d W_f/W_t) */
  double bytes_sampler; /* algorithmic HBM bytes of the diagonal-branch launch (SURVEY 8d Q_d = 8nT + 16N per baseline-iteration) */
  double ms_diag_project; /* one-off projection yhat = F^T N^-1/2 z of the diagonal-branch baselines */
  double ms_sampler_diag; /* all Gibbs iterations of the diagonal-branch baselines (streaming kernel) */
  int64_t n_dense;        /* baselines that ran the dense system */
  int64_t n_diag;         /* baselines that ran the diagonal(+rank-1) system */
} dgibbs_timing;

int dgibbs_version(void);
/* Which generator DGIBBS_RNG_PHILOX uses for the N(0,1) / chi^2 variates in this build (static string). */
const char* dgibbs_rng_info(void);
/* sizeof of the ABI structs as compiled (0: dgibbs_problem, 1: dgibbs_result, 2: dgibbs_timing, 3: dgibbs_cross_problem);
 * lets a binding verify its struct layout without a GPU. */
int dgibbs_sizeof(int which);

/* Create a handle on CUDA device `device_id`.  `stream` is a cudaStream_t passed as an
 * integer (0 = the library creates its own non-blocking stream). */
int dgibbs_create(int device_id, uint64_t stream, dgibbs_handle* out);
int dgibbs_destroy(dgibbs_handle h);
const char* dgibbs_last_error(dgibbs_handle h);

/* Upper bound of device workspace (bytes) dgibbs_run would allocate for this problem. */
int dgibbs_workspace_bytes(dgibbs_handle h, const dgibbs_problem* p, int64_t* bytes);

/* Run every chain of the batch on the device (replaces delay.py:1146-1178).
 * Host-pointer problems of at least 16 waves of baselines (2 x 8 x SM count; DGIBBS_PIPELINE_CHUNK overrides the chunk,
 * 0 = never) are processed in chunks: the upload of a chunk overlaps the kernels of the previous one and the
 * download of its results the next one (pinned host buffers make the copies asynchronous).  Results do not depend
 * on the chunking. */
int dgibbs_run(dgibbs_handle h, const dgibbs_problem* p, dgibbs_result* r);
int dgibbs_get_timing(dgibbs_handle h, dgibbs_timing* t);

/* Debug / parity: the three return values of _compute_delay_spectrum_inputs
 * (delay.py:1657-1710) for baseline `b` of the problem, written to HOST or DEVICE buffers
 * according to p->on_device:
 *   data_w : (2M, T)      windowed + whitened data, interleaved re/im rows
 *   FTNih  : (n_sys, 2M)
 *   FTNiF  : (n_sys, n_sys)
 * any of the three may be NULL. */
int dgibbs_setup_inputs(dgibbs_handle h, const dgibbs_problem* p, int32_t b,
                        double* data_w, double* FTNih, double* FTNiF);

/* Wiener filter (delay.py:2132-2201): one deterministic solve per baseline with
 * S = delay_PS (passed in p->S0); `out` is [sum_b T[b] * n_sys] doubles, baseline b at
 * offset n_sys * sum_{b'<b} T[b'], row-major (T[b], n_sys) (i.e. complex128 (T, N) in
 * complex_timedomain mode). */
int dgibbs_wiener(dgibbs_handle h, const dgibbs_problem* p, double* out, uint8_t* success);

/* Fourier matrices of delay.py:1480-1613 generated on the device (API completeness and
 * the definition of the masked forward/inverse DFT kernels); `kind`: 0 r2c (2M,N),
 * 1 c2r (N,2M), 2 c2c (2M,2N), 3 complex (M,N) as interleaved re/im.  Host output. */
int dgibbs_fourier_matrix(dgibbs_handle h, int32_t kind, int32_t N, const int64_t* fsel,
                          int32_t M, double* out);

/* ------------------------------------------------------------------------------------------
 * Cross-power-spectrum Gibbs sampler: replaces delay_spectrum_gibbs_cross
 * (draco/analysis/delay.py:1907-2099) as called per baseline by
 * DelayCrossPowerSpectrumEstimator._evaluate (delay.py:1395-1447).
 *
 * A batch of B independent problems (baselines) sharing nd datasets, N delays, T samples and
 * one channel selection fsel[M] (values in [0, N); the sampler uses the complex Fourier matrix
 * fourier_matrix, delay.py:1588-1613).  HOST pointers only.
 *   data : complex128 [B, nd, T, M]      (what _cut_data returns for the stacked datasets)
 *   Ni   : float64    [B, nd, M]
 *   S0   : complex128 [B, nd, nd, N]     initial cross spectrum (Hermitian positive definite per delay)
 * Injected draws (rng_mode == DGIBBS_RNG_INJECTED), the reference's call order per iteration
 * (standard_complex_normal x2, then per delay the Bartlett factor of complex_wishart,
 * draco/util/random.py:103-166):
 *   w1 : complex128 [B, niter, N, nd, T]     w2 : complex128 [B, niter, nd, M, T]
 *   wn : float64    [B, niter, N, 2, L]      L = nd(nd-1)/2; [..,0,:] real, [..,1,:] imaginary parts
 *                                            of the strict lower triangle in numpy.tril_indices order
 *                                            (unit-variance normals, the 1/sqrt(2) is applied here)
 *   wg : float64    [B, niter, N, nd]        gamma(T - i) variates of the diagonal
 * Output
 *   samples : complex128 [B, niter, nd, nd, N]  (unshifted delay order, every iteration)
 *   n_done  : [B] completed iterations; success[b] = 0 when a factorisation / inverse failed, which
 *             is where the reference raises RuntimeError("Exiting earlier as singular") (delay.py:2096).
 */
#define DGIBBS_CROSS_MAX_DATASETS 4

typedef struct {
  int32_t struct_size;
  int32_t B, nd, N, T, M;
  int32_t niter;
  int32_t window_id;
  int32_t rng_mode;
  int32_t reserved;
  uint64_t seed;
  const void* data;
  const double* Ni;
  const int64_t* fsel;
  const void* S0;
  const int64_t* global_id; /* [B] Philox key per problem; may be NULL */
  const void* w1;
  const void* w2;
  const double* wn;
  const double* wg;
} dgibbs_cross_problem;

int dgibbs_cross_run(dgibbs_handle h, const dgibbs_cross_problem* p, void* samples, int32_t* n_done,
                     uint8_t* success);

/* ---------------------------------------------------------------------------------------------
 * Device-side time / frequency cuts.  Replaces the per-baseline Python loop around
 * DelayTransformBase._cut_data (draco/analysis/delay.py:516-597, called from delay.py:1146-1167) for a whole
 * raw slab [B, T, nfreq]: non_zero_time / non_zero_freq masks from the fraction of positive weights
 * (time_frac, freq_frac), compaction, mean removal over the kept samples, the optional scale_freq
 * normalisation, the time-averaged weights (use_average_weights=True, the only mode the samplers accept,
 * delay.py:1697) times weight_boost, and the three "skip this baseline" exits (alive[b] = 0).
 *
 * Output layout = the ragged batch of dgibbs_problem in fixed slots: baseline b's kept block (T_out[b],
 * M_out[b]) row-major starts at data_out + b*T*nfreq; its averaged weights and channel indices
 * (channel_ind of the kept frequencies) start at Ni_out / fsel_out + b*nfreq.  With on_device = 1 the outputs
 * feed dgibbs_run in place: data_off[i] = b_i*T*nfreq, chan_off[i] = b_i*nfreq over the alive baselines b_i
 * (offsets may leave gaps, they must not overlap).
 * Masks, counts and channel lists are exact; data and weights are bit-identical to numpy's (sequential sums
 * over the time axis) except under scale_freq, whose all-element std is equal to rounding. */
typedef struct {
  int32_t struct_size;
  int32_t on_device;   /* 1: every pointer below is a device pointer; 0: host pointers */
  int32_t B, T, nfreq; /* raw slab shape */
  int32_t weight_f32;  /* element type of weight: 1 = float32, 0 = float64 */
  int32_t remove_mean; /* delay.py:570-574 */
  int32_t scale_freq;  /* delay.py:580-587 */
  int32_t data_c64;    /* element type of data: 1 = complex64 (draco's containers store visibilities as complex64,
                          core/containers.py:503; the mean is then removed in numpy's complex64 arithmetic, bit for bit),
                          0 = complex128.  scale_freq needs complex128. */
  int32_t reserved0;   /* keep 0 */
  double time_frac;    /* delay.py:552 */
  double freq_frac;    /* delay.py:560 */
  double weight_boost; /* delay.py:593 */
  const void* data;           /* [B, T, nfreq] complex128 (complex64 with data_c64) */
  const void* weight;         /* [B, T, nfreq] */
  const int64_t* channel_ind; /* [nfreq] channel of each input frequency on the padded grid (delay.py:461-512) */
  void* data_out;    /* [B, T*nfreq] complex128 */
  double* Ni_out;    /* [B, nfreq] */
  int64_t* fsel_out; /* [B, nfreq] */
  int32_t* T_out;    /* [B] kept samples (0 for a skipped baseline) */
  int32_t* M_out;    /* [B] kept frequencies */
  uint8_t* alive;    /* [B] 0 = _cut_data returned None */
  uint8_t* keep_t;   /* [B, T] non_zero_time; may be NULL */
  uint8_t* keep_f;   /* [B, nfreq] non_zero_freq; may be NULL */
} dgibbs_cut;

int dgibbs_cut_data(dgibbs_handle h, const dgibbs_cut* c);

/* ---------------------------------------------------------------------------------------------
 * Plain delay transform by inverse FFT: delay_spectrum_fft, draco/analysis/delay.py:2102-2130 (task DelaySpectrumFFT,
 * delay.py:960-980), for nrows rows of N complex128 channels at once: out[r] = ifft(data[r] * window(i / N)) in
 * numpy's convention (sign +, scale 1/N); shift = 1 adds the fftshift of DelaySpectrumFFT._estimator (delay.py:980).
 * window_id as in dgibbs_problem (-1 = none).  N = 2^m <= 4096 runs as an in-smem FFT, any other N <= 6400 as a
 * direct DFT.  data / out are host pointers (on_device = 0) or device pointers; out may not alias data.
 * ms_kernel (may be NULL) receives the CUDA-event time of the transform kernel. */
int dgibbs_delay_fft(dgibbs_handle h, int32_t on_device, int64_t nrows, int32_t N, int32_t window_id, int32_t shift,
                     const void* data, void* out, double* ms_kernel);

/* dgibbs_delay_fft for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  It has NO counterpart in the reference.
 * Same transform and arguments as dgibbs_delay_fft with on_device = 1 (data / out are DEVICE pointers on the
 * handle's device; out may not alias data), but the window and twiddle tables live in a static __device__ array of
 * the library (160 KB held by every GPU that loads it) instead of the handle's pool: the call performs no cudaMalloc /
 * cudaFree and calls no allocating routine.  Limit: N = 2^m with 8 <= N <= DGIBBS_NOALLOC_MAX_N; anything else is
 * rejected with -2 and a message (use dgibbs_delay_fft).  nrows = 0 only performs the once-per-device set-up (the
 * cudaGetSymbolAddress look-up), e.g. ahead of a capture.  sync = 1 synchronises the handle's stream before returning,
 * sync = 0 only enqueues the kernels on it.  All calls on a device share the one scratch array (rebuilt at the head
 * of every call): calls that may OVERLAP on the GPU (different streams or handles of one device) race; calls on one
 * stream are ordered and safe. */
#define DGIBBS_NOALLOC_MAX_N 4096
int dgibbs_delay_fft_noalloc(dgibbs_handle h, int64_t nrows, int32_t N, int32_t window_id, int32_t shift,
                             const void* data, void* out, int32_t sync);

/* DelaySpectrumToPowerSpectrum.process, draco/analysis/delay.py:1061-1106: var[b, n] = variance over the unmasked
 * samples t of spec[b, t, n] (complex128, [B, T, N]); mask [B, T] with 1 = masked sample (the DelayTransform's
 * spectrum_mask) or NULL.  With a mask, NaN results (no unmasked sample) become 0 and bad[b] = 1 (delay.py:1099-1104);
 * bad may be NULL. */
int dgibbs_sample_variance(dgibbs_handle h, int32_t on_device, int32_t B, int32_t T, int32_t N, const void* spec,
                           const uint8_t* mask, double* var, uint8_t* bad, double* ms_kernel);

#ifdef __cplusplus
}
#endif
#endif /* DGIBBS_H */
