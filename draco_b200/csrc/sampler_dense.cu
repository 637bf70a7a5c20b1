// Dense (delay-space) branch of the Gibbs sampler: one persistent CTA per baseline chain.
//
// Per iteration and baseline (reference: _draw_signal_sample_f + _draw_ps_sample,
// draco/analysis/delay.py:1778-1819, 1862-1880):
//   prepass   Si = 1/S ;  Y = sqrt(Si) w1 ;  u = h (z + w2)
//   phase 0   Y += u F                      masked inverse DFT as a DMMA GEMM against the shared table
//   factor    Ci = circ(c) + diag(Si) = L L^T, left-looking, 128-wide block columns, DMMA updates;
//             Ci is never stored: its tiles are generated from the circulant first column.
//             The RHS rows ride along as extra row tiles, so the forward solve is free.
//   backward  x = L^-T z  (DMMA), sample-major
//   draw      S = var_t(x) T / chi2
// W_j = inv(L_jj) is kept (packed by 8x8 blocks, bank-swizzled) in shared memory for the whole
// block column, so every triangular solve is a DMMA multiply against a smem-staged tile.
//
// Sharing a chain.  A chain is serial over its iterations, so with B chains on S SMs the last wave of a launch runs
// with (B mod S) busy SMs (4096 baselines over 8 GPUs: 512 chains = 3.46 waves executed as 4).  The kernel is therefore
// also instantiated for thread-block clusters of C = 2 and 4 CTAs that work on ONE chain: the sample rows of the
// prepass / backward solve, the row tiles below the diagonal of every block column and the delays of the
// power-spectrum draw are divided over the cluster ranks; every rank factors the diagonal tile itself (identical
// arithmetic, so no broadcast of W_j).  Row tile i belongs to rank i mod C for the whole factorisation, so a rank only
// ever reads foreign data through row panel j at step j: the split-phase cluster barrier (arrive after tile (j+1, j)
// is stored, wait at the top of step j+1) is the only synchronisation of a block column.  Every tile is computed by
// exactly the code and data of the single-CTA kernel, so the samples are bit-identical for every C.  The host
// (launch_dense_sampler) runs the full waves with C = 1 and picks C for the remainder.
//
// All hot loops are ROLLED or small: straight-line code larger than the instruction caches costs
// ~140 cycles per 128 B line on this part (measured, profiles/r01a), so register-tile code that
// touches all 64 accumulators is kept to loads/stores only.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "dense_core.cuh"
#include "tma_host.h"
#include "tma_pipe.cuh"

#ifndef DGIBBS_COST_C2
#define DGIBBS_COST_C2 0.59
#endif
#ifndef DGIBBS_COST_C4
#define DGIBBS_COST_C4 0.42
#endif

namespace dg {

// ----------------------------------------------------------------------------- the sampler
// Shared memory:  [ctab: nsys (optional)] [sWp: WP_DOUBLES] [pipe/tile: PIPE_DOUBLES] [misc]
template <int C>
__global__ void __launch_bounds__(THREADS, 1) dense_sampler_kernel(DevBatch d, int it_begin, int it_end,
                                                                   const __grid_constant__ CUtensorMap tmap) {
  extern __shared__ __align__(128) double smem[];
  const Lane ln;
  const int rank = C > 1 ? cluster_ctarank() : 0;              // rank of this CTA among the C that share a chain
  const int slot = C > 1 ? (int)blockIdx.x / C : (int)blockIdx.x;  // chain slot (workspace) of the cluster
  const int nslot = (int)gridDim.x / C;
  const int N = d.N, nsys = d.nsys, npad = d.npad, nblk = d.nblk, Kpad = d.Kpad, RT = d.RT;
  double* s_ct = smem;
  double* sWp = smem + (d.ctab_smem ? nsys : 0);
  double* pipe = sWp + WP_DOUBLES;
  double* misc = pipe + PIPE_DOUBLES;
  double* cwin = misc + 16;  // circulant window of the current tile
  TmaRing ring;
  const bool use_tma = d.use_tma != 0;  // NT updates through the TMA ring (tma_pipe.cuh) instead of cp.async
  if (use_tma) tma_ring_init(ring, reinterpret_cast<unsigned long long*>(misc + 16 + CWIN));

  const size_t wrows = (size_t)npad + (size_t)RT * TILE;
  double* Wk = d.work + (size_t)slot * d.work_stride;   // L rows then RHS rows, pitch npad
  double* Rh = Wk + (size_t)npad * npad;               // RHS rows (sample-major)
  double* Wpg = d.Wp + (size_t)blockIdx.x * nblk * WP_DOUBLES;  // packed inv(L_jj), one per block column (per CTA)
  double* uT = d.uT + (size_t)slot * RT * TILE * Kpad;
  double* Si = d.Si + (size_t)blockIdx.x * npad;                // per CTA

  long long tlast = clock64();
  const int nl = d.li_end > 0 ? d.li_end : (d.blist ? d.nlist : d.B);
  for (int li = d.li_begin + slot; li < nl; li += nslot) {
    const int b = d.blist ? d.blist[li] : li;
    if (d.failed[b]) continue;
    const bool isdiag = d.branch[b] == 3;
    const double dg_a = d.diag_ab[2 * b], dg_b = d.diag_ab[2 * b + 1];
    const int Tb = d.T[b], Mb = d.M[b];
    const double* ctg = d.ctab + (size_t)b * nsys;
    const double* ct = ctg;
    if (d.ctab_smem) {
      __syncthreads();
      for (int k = threadIdx.x; k < nsys; k += THREADS) s_ct[k] = ctg[k];
      ct = s_ct;
    }
    const double* h = d.h + (size_t)b * Kpad;
    const double* zT = d.zT + (size_t)b * d.Tmax * Kpad;
    const int* cidx = d.cidx + (size_t)b * d.nch;
    double* S = d.S + (size_t)b * N;
    const Philox ph(mix_key(d.seed, d.gid ? (unsigned long long)d.gid[b] : (unsigned long long)b));
    const int rhs_frags = (Tb + 7) >> 3;  // valid 8-row fragments of the RHS

    if (d.use_fft) {  // RHS rows the FFT projection never writes (skipped by every tile variant): keep them finite
      const long long r0 = (long long)rhs_frags * 8 * (npad / 2), r1 = (long long)RT * TILE * (npad / 2);
      for (long long e = r0 + threadIdx.x + rank * THREADS; e < r1; e += THREADS * C)
        reinterpret_cast<double2*>(Rh)[e] = make_double2(0.0, 0.0);
    }
    // fragments (8 sample rows each) of the right-hand sides this rank solves backwards
    const int frag_per = (rhs_frags + C - 1) / C;
    const int frag_lo = rank * frag_per, frag_hi = min(rhs_frags, frag_lo + frag_per);
    bool alive = true;
    for (int it = it_begin; it < it_end && alive; it++) {
      // ---------------- prepass
      __syncthreads();
      for (int k = threadIdx.x; k < npad; k += THREADS) {
        double si = 0.0;
        if (k < nsys) {
          double s = S[d.cplx ? (k >> 1) : k];
          si = (s != 0.0) ? 1.0 / s : 0.0;   // invert_no_zero, delay.py:1785
          if (d.cplx) si *= 2.0;             // delay.py:1794
        }
        Si[k] = si;
      }
      __syncthreads();
      Acc acc;
      if (d.use_fft) {
        // ---------------- fused prepass + phase 0: draws, u, masked inverse DFT (FFT), Y
        DrawCtx dc;
        dc.d = &d; dc.ph = &ph; dc.cidx = cidx; dc.h = h; dc.zT = zT;
        dc.w1 = (d.rng_mode == 0) ? d.w1 + d.w1_off[b] + (long long)it * nsys * Tb : nullptr;
        dc.w2 = (d.rng_mode == 0) ? d.w2 + d.w2_off[b] + (long long)it * 2 * Mb * Tb : nullptr;
        dc.Tb = Tb; dc.it = it; dc.half = Kpad >> 1;
        double2* fbuf = reinterpret_cast<double2*>(pipe);
        project_rows_fft(d, dc, Rh, Si, fbuf, reinterpret_cast<double2*>(sWp), rhs_frags * 8, rank, C);
        PROF(0);
      } else {
      {  // Y = sqrt(Si) * w1  (two adjacent delays per thread), zero in the padding
        const int half = npad >> 1;
        const long long tot = (long long)RT * TILE * half;
        for (long long e = threadIdx.x + rank * THREADS; e < tot; e += THREADS * C) {
          int t = (int)(e / half), k = (int)(e % half) * 2;
          double y0 = 0.0, y1 = 0.0;
          if (t < Tb && k < nsys && d.rng_mode != 2) {
            double n0, n1;
            if (d.rng_mode == 1) {
              unsigned long long blk = ((unsigned long long)t * nsys + k) >> 1;
              normal_pair(ph((unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_W1), n0, n1);
            } else {
              const double* w1 = d.w1 + d.w1_off[b] + (long long)it * nsys * Tb;
              n0 = w1[(long long)k * Tb + t];
              n1 = w1[(long long)(k + 1) * Tb + t];
            }
            y0 = sqrt(Si[k]) * n0;       // delay.py:1815
            y1 = sqrt(Si[k + 1]) * n1;
          }
          *reinterpret_cast<double2*>(Rh + (size_t)t * npad + k) = make_double2(y0, y1);
        }
      }
      {  // u = h * (z + w2)  on the full channel grid (delay.py:1814,1817), zero where masked/padded
        const int half = Kpad >> 1;
        const long long tot = (long long)RT * TILE * half;
        for (long long e = threadIdx.x + rank * THREADS; e < tot; e += THREADS * C) {
          int t = (int)(e / half), i = (int)(e % half);
          double u0 = 0.0, u1 = 0.0;
          if (t < Tb && i < d.nch) {
            int ci = cidx[i];
            if (ci >= 0) {
              double n0 = 0.0, n1 = 0.0;
              if (d.rng_mode == 1) {
                unsigned long long blk = (unsigned long long)t * half + i;
                normal_pair(ph((unsigned)blk, (unsigned)(blk >> 32), (unsigned)it, STREAM_W2), n0, n1);
              } else if (d.rng_mode == 0) {
                const double* w2 = d.w2 + d.w2_off[b] + (long long)it * 2 * Mb * Tb;
                n0 = w2[(long long)(2 * ci) * Tb + t];
                n1 = w2[(long long)(2 * ci + 1) * Tb + t];
              }
              double2 z = *reinterpret_cast<const double2*>(zT + (size_t)t * Kpad + 2 * i);
              u0 = h[2 * i] * (z.x + n0);
              u1 = h[2 * i + 1] * (z.y + n1);
            }
          }
          *reinterpret_cast<double2*>(uT + (size_t)t * Kpad + 2 * i) = make_double2(u0, u1);
        }
      }
      __syncthreads();
      if (C > 1) cluster_sync_all();  // Y and u complete before any rank's projection tiles read them
      PROF(0);

      // ---------------- phase 0: Y += u F   (column blocks divided over the ranks)
      for (int rt = 0; rt < RT; rt++) {
        const bool do0 = rt * 16 + ln.rf[0] < rhs_frags, do1 = rt * 16 + ln.rf[1] < rhs_frags;
        const int var = rhs_variant(do0, do1);
        for (int cb = rank; cb < nblk; cb += C) {
          double* ytile = Rh + (size_t)rt * TILE * npad + cb * TILE;
          tile_load_global(acc, ytile, npad, ln);
          gemm_nn<false>(pipe, acc, ln, uT + (size_t)rt * TILE * Kpad, Kpad, d.Ftab + cb * TILE, npad, 0, Kpad / KC, var);
          tile_store_global(acc, ytile, npad, ln, do0, do1);
        }
      }
      __syncthreads();
      PROF(1);
      }

      if (C > 1) cluster_sync_all();  // every sample row of y is in place
      if (C == 1 && d.project_only) {  // export y^T = (F^T h z)^T for the streaming diagonal kernel
        double* yo = d.yhat + (size_t)d.ycompact[b] * d.Tmax * d.ypitch;
        const int hp = d.ypitch / 2;
        for (long long e = threadIdx.x; e < (long long)Tb * hp; e += THREADS) {
          int t = (int)(e / hp), k = (int)(e % hp) * 2;
          double2 v = make_double2(0.0, 0.0);
          if (k < npad) v = *reinterpret_cast<const double2*>(Rh + (size_t)t * npad + k);
          *reinterpret_cast<double2*>(yo + (size_t)t * d.ypitch + k) = v;
        }
        break;
      }

      if (isdiag) alive = diag_solve_rows(Rh, Si, dg_a, dg_b, nsys, npad, Tb, misc, rank, C);

      // ---------------- factor (+ forward solve on the RHS row tiles)
      bool pending = false;  // a cluster arrive whose wait is still to come
      for (int j = 0; j < nblk && alive && !isdiag; j++) {
        if (C > 1 && pending) {  // row panel j is complete: tile (j, j-1) was stored before its owner arrived
          cluster_wait();
          pending = false;
        }
        const double* Bp = Wk + (size_t)j * TILE * npad;
        // diagonal tile
        const bool interior_j = !d.cplx && (j + 1) * TILE <= nsys;
        if (interior_j) load_circ_window(cwin, ctg, N, 0);
        acc_zero(acc);
        if (use_tma) gemm_nt_tma(&tmap, ring, pipe, acc, ln, slot, j * TILE, j * TILE, j * (TILE / TKC), ln.warp < 4 ? V_DIAG_LO : V_DIAG_HI);
        else gemm_nt(pipe, acc, ln, Bp, Bp, npad, j * (TILE / KC), ln.warp < 4 ? V_DIAG_LO : V_DIAG_HI);
        PROF(2);
        if (interior_j) {  // generated part added in registers on the way to the staged tile
          acc_add_diag_win(acc, ln, cwin, Si, j * TILE);
          stage_acc(acc, pipe, ln, true, true);
        } else {
          stage_acc(acc, pipe, ln, true, true);
          __syncthreads();
          add_generated(pipe, ct, Si, N, nsys, d.cplx, j, j, true);
        }
        __syncthreads();
        PROF(7);
        alive = factor_diag(pipe, sWp, misc, ln);
        if (!alive) break;
        PROF(3);
        for (int idx = threadIdx.x; idx < WP_DOUBLES / 2; idx += THREADS)
          reinterpret_cast<double2*>(Wpg + (size_t)j * WP_DOUBLES)[idx] = reinterpret_cast<const double2*>(sWp)[idx];
        __syncthreads();
        PROF(4);
        // tiles below the diagonal, then the RHS row tiles
        for (int i = j + 1; i < nblk + RT; i++) {
          if (C > 1 && i % C != rank) {  // another rank's row tile
            if (i == j + 1) {
              cluster_arrive();
              pending = true;
            }
            continue;
          }
          bool do0 = true, do1 = true;
          int nrows = TILE;
          double* out;
          if (i < nblk) {
            out = Wk + (size_t)i * TILE * npad + j * TILE;
          } else {
            int rt = i - nblk;
            do0 = rt * 16 + ln.rf[0] < rhs_frags;
            do1 = rt * 16 + ln.rf[1] < rhs_frags;
            nrows = min(TILE, max(0, rhs_frags * 8 - rt * TILE));
            out = Rh + (size_t)rt * TILE * npad + j * TILE;
          }
          const int var = rhs_variant(do0, do1);
          const bool interior = i < nblk && !d.cplx && (i + 1) * TILE <= nsys;
          if (interior) load_circ_window(cwin, ctg, N, i - j);
          acc_zero(acc);
          if (use_tma) gemm_nt_tma(&tmap, ring, pipe, acc, ln, slot, i * TILE, j * TILE, j * (TILE / TKC), var);
          else gemm_nt(pipe, acc, ln, Wk + (size_t)i * TILE * npad, Bp, npad, j * (TILE / KC), var);
          PROF(6);
#if DGIBBS_TRISOLVE_REG
          if (interior || i >= nblk) {
            // the tile stays in the accumulator registers: generated part (or the right-hand side) added in place,
            // solved by quad shuffles + DMMA, stored straight to the workspace
            if (interior) acc_add_circ(acc, ln, cwin);  // (gemm_nt ends with a barrier: the window is visible)
            else acc_add_global(acc, out, npad, ln, do0, do1);
            PROF(5);
            trisolve_reg<false>(acc, sWp, ln, var);
            tile_store_global(acc, out, npad, ln, do0, do1);
            __syncthreads();  // the circulant window is rewritten for the next tile
            PROF(8);
            if (C > 1 && i == j + 1) {
              cluster_arrive();
              pending = true;
            }
            continue;
          }
#endif
          if (interior) {  // interior tile below the diagonal: generated in registers (gemm_nt ends with a barrier)
            stage_acc_circ(acc, pipe, ln, cwin);
          } else {
            stage_acc(acc, pipe, ln, do0, do1);
            __syncthreads();
            if (i < nblk) add_generated(pipe, ct, Si, N, nsys, d.cplx, i, j, false);
            else add_global(pipe, out, npad, nrows);
          }
          __syncthreads();
          PROF(5);
          trisolve<false>(pipe, sWp, ln, var);
          __syncthreads();
          copy_tile_out(pipe, out, npad, nrows);
          __syncthreads();
          PROF(8);
          if (C > 1 && i == j + 1) {  // next step's row panel is stored: let the other ranks go on
            cluster_arrive();
            pending = true;
          }
        }
      }
      if (C > 1) {
        if (pending) cluster_wait();
        cluster_sync_all();  // every right-hand-side tile (forward solve) and, on the diagonal branch, every row is stored
      }
      if (!alive) {
        __syncthreads();
        if (threadIdx.x == 0 && rank == 0) {
          d.failed[b] = 1;
          d.n_done[b] = it;
        }
        break;
      }

      // ---------------- backward solve  x^T = z^T L^-1
      for (int rt = 0; rt < RT && !isdiag; rt++) {
        // the sample rows are independent: this rank solves the fragments [frag_lo, frag_hi)
        const int f0 = rt * 16 + ln.rf[0], f1 = rt * 16 + ln.rf[1];
        const bool do0 = f0 >= frag_lo && f0 < frag_hi, do1 = f1 >= frag_lo && f1 < frag_hi;
        const int var = rhs_variant(do0, do1);
        const int row_lo = min(TILE, max(0, frag_lo * 8 - rt * TILE)), nrows = min(TILE, max(0, frag_hi * 8 - rt * TILE));
        if (nrows <= row_lo) continue;  // CTA-uniform
        double* rrow = Rh + (size_t)rt * TILE * npad;
        for (int j = nblk - 1; j >= 0; j--) {
          double* xt = rrow + j * TILE;
          // W_j streams into sWp underneath the GEMM (its copy group is older than the pipeline's)
          for (int idx = threadIdx.x; idx < WP_DOUBLES / 2; idx += THREADS)
            cp_async16(sWp + idx * 2, Wpg + (size_t)j * WP_DOUBLES + idx * 2);
          cp_async_commit();
          acc_zero(acc);
          gemm_nn<true>(pipe, acc, ln, rrow, npad, Wk + j * TILE, npad, (j + 1) * TILE, (nblk - 1 - j) * (TILE / KC), var);
          PROF(9);
#if DGIBBS_TRISOLVE_REG
          acc_add_global(acc, xt, npad, ln, do0, do1);
          trisolve_reg<true>(acc, sWp, ln, var);
          tile_store_global(acc, xt, npad, ln, do0, do1);
          __syncthreads();  // x_j is read by the next update; sWp is reloaded
#else
          stage_acc(acc, pipe, ln, do0, do1);
          __syncthreads();
          add_global(pipe, xt, npad, nrows, row_lo);
          __syncthreads();
          trisolve<true>(pipe, sWp, ln, var);
          __syncthreads();
          copy_tile_out(pipe, xt, npad, nrows, row_lo);
          __syncthreads();
#endif
          PROF(10);
        }
      }

      if (C > 1) cluster_sync_all();  // x complete (rows solved by different ranks)
      if (d.wiener) {
        double* xo = d.xout + (size_t)d.xoff[b] * nsys;
        for (long long e = threadIdx.x + rank * THREADS; e < (long long)Tb * nsys; e += THREADS * C) {
          int t = (int)(e / nsys), k = (int)(e % nsys);
          xo[e] = Rh[(size_t)t * npad + k];
        }
        if (threadIdx.x == 0 && rank == 0) d.n_done[b] = 1;
        break;
      }

      // ---------------- power-spectrum draw (delay.py:1862-1880)
      for (int k = threadIdx.x + rank * THREADS; k < N; k += THREADS * C) {
        double var = 0.0;
        const int ncomp = d.cplx ? 2 : 1;
        for (int comp = 0; comp < ncomp; comp++) {
          const double* col = Rh + (d.cplx ? 2 * k + comp : k);
          double mean = 0.0;
          for (int t = 0; t < Tb; t++) mean += col[(size_t)t * npad];
          mean /= (double)Tb;
          double ss = 0.0;
          for (int t = 0; t < Tb; t++) {
            double dv = col[(size_t)t * npad] - mean;
            ss += dv * dv;
          }
          var += ss / (double)Tb;
        }
        double chi2;
        if (d.rng_mode == 1) {
          // chi^2_T = sum of T squared normals; pairs collapse to -2 ln(u)
          double acc2 = 0.0;
          const unsigned long long base = (unsigned long long)k * ((Tb + 3) / 4 + 1);
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
          chi2 = d.chi2[((size_t)b * d.niter + it) * N + k];
        }
        double snew = var * (double)Tb / chi2;
        S[k] = snew;
        d.samples[((size_t)it * d.B + b) * N + k] = snew;
      }
      if (threadIdx.x == 0 && rank == 0) d.n_done[b] = it + 1;
      __syncthreads();
      if (C > 1) cluster_sync_all();  // S of the next iteration; the right-hand sides may be overwritten
      PROF(11);
    }
  }
}

size_t dense_smem_bytes(const DevBatch& d) {
  return sizeof(double) * ((d.ctab_smem ? (size_t)d.nsys : 0) + WP_DOUBLES + PIPE_DOUBLES + MISC_DOUBLES + 2 * TNST);
}
size_t dense_wp_doubles_per_slot(const DevBatch& d) { return (size_t)d.nblk * WP_DOUBLES; }

cudaError_t dense_configure(const DevBatch& d) {
  const int bytes = (int)dense_smem_bytes(d);
  cudaError_t e = cudaFuncSetAttribute(dense_sampler_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(dense_sampler_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(dense_sampler_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

// ---- tensor map of the chain workspaces (tma_pipe.cuh / tma_host.h): (k mod 4, row, k div 4, slot)
// true: `tm` describes d.work for the TMA ring; false: this geometry keeps the cp.async loaders
static bool make_work_tensor_map(const DevBatch& d, CUtensorMap* tm) {
  if (d.project_only) {
    memset(tm, 0, sizeof *tm);
    return false;
  }
  const size_t ring_off = sizeof(double) * ((d.ctab_smem ? (size_t)d.nsys : 0) + WP_DOUBLES);
  return tma_work_map(tm, d.work, (size_t)d.npad + (size_t)d.RT * TILE, d.npad, d.work_stride, d.nslots, ring_off, TILE, TKC);
}

template <int C>
static cudaError_t launch_one(const DevBatch& d0, int it_begin, int it_end, int nclusters, cudaStream_t st) {
  DevBatch d = d0;
  CUtensorMap tm;
  d.use_tma = make_work_tensor_map(d, &tm) ? 1 : 0;
  if (C == 1) {
    dense_sampler_kernel<1><<<nclusters, THREADS, dense_smem_bytes(d), st>>>(d, it_begin, it_end, tm);
    return cudaGetLastError();
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nclusters * C);
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = dense_smem_bytes(d);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = C;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, dense_sampler_kernel<C>, d, it_begin, it_end, tm);
}

// clusters of C CTAs that the device keeps resident at once (C = 1: one CTA per SM)
template <int C>
static int max_clusters(const DevBatch& d, int sm_count) {
  if (C == 1) return sm_count;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((sm_count / C) * C);
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = dense_smem_bytes(d);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = C;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, dense_sampler_kernel<C>, &cfg) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

// Relative duration of one chain on a cluster of C CTAs (C = 1: 1.0), measured at cfg2 geometry
// (profiles/r02a_cluster_split.json.log): the diagonal tile is factored by every rank and the backward
// solve is latency-bound, so the split is worth less than C.
static double chain_cost(int C) { return C == 1 ? 1.0 : (C == 2 ? DGIBBS_COST_C2 : DGIBBS_COST_C4); }

DensePlan plan_dense(const DevBatch& d, int nchains, int sm_count, int force_cluster) {
  DensePlan p{};
  const int slots = d.nslots;  // workspace slots = concurrent chains
  int ncl[5] = {0, slots < sm_count ? slots : sm_count, 0, 0, 0};
  ncl[2] = max_clusters<2>(d, sm_count);
  ncl[4] = max_clusters<4>(d, sm_count);
  for (int C = 2; C <= 4; C *= 2) ncl[C] = ncl[C] < slots ? ncl[C] : slots;
  if (force_cluster == 1 || force_cluster == 2 || force_cluster == 4) {
    if (ncl[force_cluster] > 0) {
      p.C_tail = force_cluster;
      p.n_tail = nchains;
      p.grid_tail = nchains < ncl[force_cluster] ? nchains : ncl[force_cluster];
      return p;
    }
  }
  // full waves (one chain per SM) on single CTAs; what is left over -- all of a small batch -- on the cluster size that
  // finishes it soonest
  p.n_main = (nchains / sm_count) * sm_count;
  p.grid_main = ncl[1];
  const int rem = nchains - p.n_main;
  if (rem == 0) return p;
  double best = 1e300;
  for (int C = 1; C <= 4; C *= 2) {
    if (ncl[C] <= 0) continue;
    if (C > 1 && (d.project_only || d.nblk < 2)) continue;  // nothing to divide
    const int rounds = (rem + ncl[C] - 1) / ncl[C];
    const double cost = rounds * chain_cost(C);
    if (cost < best - 1e-9) {
      best = cost;
      p.C_tail = C;
      p.grid_tail = rem < ncl[C] ? rem : ncl[C];
    }
  }
  p.n_tail = rem;
  return p;
}

cudaError_t launch_dense_sampler(const DevBatch& d, int it_begin, int it_end, cudaStream_t st, int sm_count,
                                 int force_cluster, int* launches) {
  const int nchains = d.blist ? d.nlist : d.B;
  if (sm_count <= 0) sm_count = d.nslots;
  const DensePlan p = plan_dense(d, nchains, sm_count, force_cluster);
  static const bool verbose = getenv("DGIBBS_VERBOSE") != nullptr;
  if (verbose)
    fprintf(stderr, "[dgibbs] dense plan: %d chains = %d on %d CTAs + %d on %d clusters of %d\n", nchains, p.n_main,
            p.grid_main, p.n_tail, p.grid_tail, p.C_tail);
  cudaError_t e = cudaSuccess;
  if (p.n_main > 0) {
    DevBatch dm = d;
    dm.li_begin = 0;
    dm.li_end = p.n_main;
    e = launch_one<1>(dm, it_begin, it_end, p.grid_main, st);
    if (e != cudaSuccess) return e;
    if (launches) ++*launches;
  }
  if (p.n_tail > 0) {
    DevBatch dt = d;
    dt.li_begin = p.n_main;
    dt.li_end = p.n_main + p.n_tail;
    switch (p.C_tail) {
      case 4: e = launch_one<4>(dt, it_begin, it_end, p.grid_tail, st); break;
      case 2: e = launch_one<2>(dt, it_begin, it_end, p.grid_tail, st); break;
      default: e = launch_one<1>(dt, it_begin, it_end, p.grid_tail, st); break;
    }
    if (e != cudaSuccess) return e;
    if (launches) ++*launches;
  }
  return e;
}

}  // namespace dg
