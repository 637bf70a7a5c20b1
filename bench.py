#!/usr/bin/env python
"""Headline benchmark: baseline-Gibbs-iterations/s at nfreq=1024 (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config cfg2]

One "step" = one pass of the whole hot path (setup + all nsamp Gibbs iterations + median) over one
batch of synthetic visibilities of the named config; at N GPUs each rank runs that batch on its own
shard of the baseline axis (weak scaling; per-rank work fixed), followed by the NCCL all-reduce of
the baseline-group averages.  `value` has the inputs resident in HBM; `e2e` goes through the
host-pointer C-ABI call (H2D + D2H inside the timed region).  Rank 0 prints ONE JSON line.

The same line carries three more records (each skippable with --no-strong / --no-diag / --no-task) and, at one GPU, the
side record `extra.int8_gemm` of the int8 tensor-core GEMM behind the Wiener operator / NRML objective (--no-int8):
  strong    ONE step of BASELINE config 4 at FIXED total size: 4096 baselines x nsamp 200 divided over the ranks in
            contiguous blocks (dist.split_range, the reference's MPI layout delay.py:2297-2298), 64 groups of 64
            averaged with an NCCL all-reduce, spectra all-gathered; `result_checksum` must be identical at 1/2/4/8 GPUs.
  diag      BASELINE config 3 (diagonal branch, HBM-bound): 16384 pixels / GPUs per rank, device-resident `value`
            with its HBM roofline and `e2e` from pinned complex64 host buffers through the pipelined upload.
  e2e_task  wall time of DelayPowerSpectrumGibbs.process() on a complex64 / float32 SiderealStream shard
            (cut + pack + H2D + chains + D2H: SURVEY 8d's definition of the metric).
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

T_START = time.perf_counter()
METRIC = "baseline-Gibbs-iterations/sec @nfreq=1024"
UNIT = "baseline-iterations/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg2")
    ap.add_argument("--nbase", type=int, default=None, help="override baselines per rank (debug)")
    ap.add_argument("--nsamp", type=int, default=None, help="override Gibbs iterations (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the fixed-size cfg4 record")
    ap.add_argument("--no-diag", action="store_true", help="skip the cfg3 (diagonal branch) record")
    ap.add_argument("--no-task", action="store_true", help="skip the task-level end-to-end record")
    ap.add_argument("--no-int8", action="store_true", help="skip the side record of the int8 tensor-core GEMM (1 GPU only)")
    ap.add_argument("--strong-nbase", type=int, default=None, help="override the TOTAL baselines of the strong record (debug)")
    ap.add_argument("--strong-nsamp", type=int, default=None)
    ap.add_argument("--diag-nbase", type=int, default=None, help="override the TOTAL pixels of the diag record (debug)")
    ap.add_argument("--e2e-steps", type=int, default=5, help="e2e is timed on min(steps, this) steps")
    ap.add_argument("--budget-s", type=float, default=780.0,
                    help="wall-clock budget of the whole run: an optional record (e2e_task / strong / diag) whose estimated "
                         "cost would go past it is skipped and says so in the line")
    return ap.parse_args()


def budget_allows(elapsed_s, estimate_s, budget_s):
    """An optional record runs only if its estimated cost keeps the whole run inside the wall-clock budget."""
    return budget_s <= 0 or elapsed_s + estimate_s <= budget_s


def workload_label(cfg, nbase, nsamp):
    """Identical in both arms so that the driver's same_config check compares like with like."""
    return (f"{cfg.name}: {nbase} baselines/GPU x {cfg.nfreq} freq (N={cfg.ndelay}), T={cfg.ntime}, nsamp={nsamp}, "
            f"window={cfg.window}, weights={cfg.weights}, mask={cfg.mask}")


# --------------------------------------------------------------------------- inputs
def make_inputs(cfg, nbase, base_offset):
    """Synthetic stack -> the cut, packed batch the task would hand to the device."""
    from draco_b200 import gibbs, synth

    freq, vis, wgt = synth.make_stack(cfg, nbase=nbase, base_offset=base_offset)
    N = cfg.ndelay
    # _cut_data with defaults on synthetic data (vectorised over baselines: same mask logic,
    # delay.py:516-597): channels with any positive weight survive; mean over time removed.
    keep = (wgt > 0).any(axis=1)  # [B, nfreq]
    common = keep.all(axis=0) | ~keep.any(axis=0)
    vis = vis - vis.mean(axis=1, keepdims=True)
    wavg = wgt.mean(axis=1)
    chan = np.arange(cfg.nfreq)
    gid = np.arange(base_offset, base_offset + nbase)
    S0 = np.full((nbase, N), 10.0)
    if common.all():
        sel = keep[0]
        return gibbs.pack_uniform(vis[:, :, sel], wavg[:, sel], chan[sel], N, S0, window=cfg.window, global_id=gid)
    datas = [vis[b][:, keep[b]] for b in range(nbase)]
    return gibbs.pack_batch(datas, [wavg[b][keep[b]] for b in range(nbase)], [chan[keep[b]] for b in range(nbase)],
                            N, S0, window=cfg.window, global_id=gid)


def make_device_batch(cfg, lo, hi, device, chunk=512):
    """Device-resident packed batch of the global baselines [lo, hi) of a config with a common channel set, built
    chunk by chunk so that the host never holds more than `chunk` baselines."""
    import torch

    from draco_b200 import gibbs

    parts = [make_inputs(cfg, min(chunk, hi - b0), b0).to_device(device) for b0 in range(lo, hi, chunk)]
    if len(parts) == 1:
        return parts[0]
    T0, M0 = int(parts[0].T[0]), int(parts[0].M[0])
    for p in parts:
        assert bool((p.T == T0).all()) and bool((p.M == M0).all()), "chunks must share (T, M)"
    B = hi - lo
    dev = parts[0].data.device
    cat = lambda name: torch.cat([getattr(p, name) for p in parts])  # noqa: E731
    return gibbs.PackedBatch(parts[0].N, cat("T"), cat("M"), torch.arange(B + 1, dtype=torch.int64, device=dev) * (T0 * M0),
                             torch.arange(B + 1, dtype=torch.int64, device=dev) * M0, cat("data"), cat("Ni"), cat("fsel"),
                             cat("S0"), cat("global_id"), parts[0].complex_timedomain, parts[0].window)


def pin(pb):
    """Copy the batch's arrays into pinned host memory (numpy views of pinned torch tensors)."""
    import dataclasses

    import torch

    kw = {}
    for f in dataclasses.fields(pb):
        v = getattr(pb, f.name)
        if isinstance(v, np.ndarray):
            raw = v.view(np.float64) if v.dtype == np.complex128 else (v.view(np.float32) if v.dtype == np.complex64 else v)
            t = torch.empty(raw.shape, dtype=torch.from_numpy(raw[:0].copy()).dtype, pin_memory=True)
            t.numpy()[...] = raw
            v = t.numpy().view(v.dtype) if v.dtype in (np.complex128, np.complex64) else t.numpy()
            kw.setdefault("_keep", []).append(t)
        kw[f.name] = v
    keep = kw.pop("_keep", [])
    out = type(pb)(**kw)
    out._pinned = keep
    return out


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index, period_ms=200):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", str(period_ms), "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# --------------------------------------------------------------------------- CPU baseline
def _cpu_chain(cfg, b, niter, use_reference):
    """One baseline of the config through the reference's CPU arithmetic: (t_setup, t_per_iteration, iterations)."""
    import oracle
    from draco_b200 import synth

    freq, vis, wgt = synth.make_stack(cfg, nbase=1, base_offset=b, threads=1)
    _, chan = oracle.calculate_delays(freq)
    d, w, nzf, _ = oracle.cut_data(vis[0], wgt[0])
    N = cfg.ndelay
    if use_reference:  # the reference's own functions, executed where they lie (build container only)
        from oracle import ref_loader

        ref = ref_loader.load()
        t0 = time.perf_counter()
        ref._compute_delay_spectrum_inputs(d.copy(), N, w, chan[nzf], cfg.window, False)
        t_setup = time.perf_counter() - t0
        t0 = time.perf_counter()
        s, ok = ref.delay_power_spectrum_gibbs(d.copy(), N, w, np.full(N, 10.0), window=cfg.window, fsel=chan[nzf],
                                               niter=niter, rng=np.random.default_rng(b))
        t_chain = time.perf_counter() - t0
    else:
        t0 = time.perf_counter()
        oracle.whiten_inputs(d, N, w, chan[nzf], cfg.window, False)
        t_setup = time.perf_counter() - t0
        t0 = time.perf_counter()
        s, ok = oracle.gibbs_chain(d, N, w, np.full(N, 10.0), cfg.window, chan[nzf], niter, np.random.default_rng(b))
        t_chain = time.perf_counter() - t0
    return t_setup, (t_chain - t_setup) / max(1, len(s)), len(s)


def _cpu_worker(args):
    os.environ["OPENBLAS_NUM_THREADS"] = os.environ["OMP_NUM_THREADS"] = os.environ["MKL_NUM_THREADS"] = "1"
    cfg_name, b0, nb, niter, use_reference = args
    try:
        from threadpoolctl import threadpool_limits
        ctx = threadpool_limits(1)
    except Exception:  # pragma: no cover
        import contextlib
        ctx = contextlib.nullcontext()
    from draco_b200 import synth

    cfg = synth.CONFIGS[cfg_name]
    with ctx:
        return [_cpu_chain(cfg, b0 + i, niter, use_reference) for i in range(nb)]


def _cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.lower().startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def _blas_info():
    try:
        from threadpoolctl import threadpool_info

        return [{k: i.get(k) for k in ("internal_api", "version", "num_threads", "threading_layer", "architecture")}
                for i in threadpool_info() if i.get("user_api") == "blas"]
    except Exception:  # pragma: no cover
        return []


def cpu_baseline(cfg, nsamp, niter_sample=10, per_proc=2, procs=None, all_threads=True):
    """BASELINE.md section 3.  The reference's CPU arithmetic on this box's host cores: P single-BLAS-thread worker
    processes (P = host cores), each owning a contiguous block of `per_proc` baselines -- the reference's MPI layout
    (no mpirun / mpi4py on the box, so multiprocessing) -- for `niter_sample` iterations, set-up timed separately,
    extrapolated linearly in baselines x nsamp.  Also 1 process x all BLAS threads.  Where the reference tree is
    present (build container) the reference's own functions are timed (kind "reference"), else the oracle port."""
    import multiprocessing as mp

    from oracle import ref_loader

    use_ref = ref_loader.available()
    P = procs or os.cpu_count() or 1
    ctx = mp.get_context("spawn")
    t0 = time.perf_counter()
    with ctx.Pool(P) as pool:
        res = pool.map(_cpu_worker, [(cfg.name, p * per_proc, per_proc, niter_sample, use_ref) for p in range(P)])
    wall = time.perf_counter() - t0
    flat = [r for rr in res for r in rr]
    t_setup = float(np.mean([r[0] for r in flat]))
    t_iter = float(np.mean([r[1] for r in flat]))
    value = P * nsamp / (t_setup + nsamp * t_iter)  # P baselines advance concurrently
    out = {"value": value, "unit": UNIT, "cores": P, "kind": "reference" if use_ref else "port",
           "layout": "multiprocessing (spawn), one process per core, 1 BLAS thread each; mpirun/mpi4py absent",
           "sample": f"{P * per_proc} baselines ({per_proc} per process) x {niter_sample} iterations of {cfg.name} geometry "
                     f"(N={cfg.ndelay}, T={cfg.ntime}); t_setup={t_setup:.3f}s t_iter={t_iter:.4f}s per baseline, "
                     f"extrapolated to nsamp={nsamp}; sample wall {wall:.1f}s",
           "sample_wall_s": wall, "sample_units": P * per_proc * niter_sample,
           "t_setup_s": t_setup, "t_iter_s": t_iter, "cpu_model": _cpu_model(), "blas": _blas_info(),
           "numpy": np.__version__}
    try:
        import scipy

        out["scipy"] = scipy.__version__
    except Exception:  # pragma: no cover
        pass
    if all_threads:
        _cpu_chain(cfg, 0, 1, use_ref)  # BLAS thread pool warm-up
        t0 = time.perf_counter()
        ts, ti, n = _cpu_chain(cfg, 0, max(2, niter_sample // 2), use_ref)
        out["one_process_all_blas_threads"] = {
            "value": nsamp / (ts + nsamp * ti), "unit": UNIT, "t_setup_s": ts, "t_iter_s": ti,
            "sample": f"1 baseline x {n} iterations, BLAS threads = library default ({os.cpu_count()} cores)",
            "sample_wall_s": time.perf_counter() - t0}
    return out


# --------------------------------------------------------------------------- main
def fp64_peak_tflops():
    """cuBLAS DGEMM 8192^3 on this GPU, best of 4 (MEASURED_PEAKS.json carries no fp64 entry)."""
    import torch

    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    best = 1e30
    for i in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        c = a @ b
        e1.record()
        torch.cuda.synchronize()
        if i:
            best = min(best, e0.elapsed_time(e1))
    del a, b, c
    torch.cuda.empty_cache()
    return 2.0 * n**3 / best / 1e9


# dram__bytes_read.sum + dram__bytes_write.sum of dense_sampler_kernel<1> (TMA operand ring, register tri-solves) from the
# committed `ncu --set full` capture profiles/r03n_dense_sampler_ncu_raw.csv: 66.73 GB read + 6.22 GB written for a launch
# of 296 baseline-iterations of cfg2 geometry = 246.5 MB per baseline-iteration (r03c / r02p, earlier kernels of the
# round: 66.74 + 6.23 / 66.78 + 6.22; the r01 captures of the cp.async kernel: 72.8-73.0 GB).  Twice the A-panel model: the B row panel is re-read per row tile and
# does not stay in L2; 25 % of the DRAM peak, not the bound.
# Reported per launch of THIS run by scaling with the launch's baseline-iterations (cfg2 geometry only).
DENSE_TRAFFIC_PER_UNIT = (66.727257e9 + 6.224364e9) / 296
DENSE_TRAFFIC_SOURCE = "ncu --set full capture of 296 baseline-iterations scaled to this launch (profiles/r03n_dense_sampler_ncu_raw.csv)"


# dram__bytes_read.sum + dram__bytes_write.sum of diag_resident_kernel<384,0> from the committed `ncu --set full` capture
# profiles/r02p_diag_path_ncu_raw.csv: 257.75 MB read + 62.32 MB written for a launch of 240 baselines x 20 iterations of
# cfg3 geometry (the projected data is read once per baseline and then stays in distributed shared memory).
DIAG_TRAFFIC_PER_UNIT = (257.753088e6 + 62.3232e6) / (240 * 20)


def hbm_roofline(nbytes, ms, units=None):
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "6650 GB/s (of fallback)"
    ach = nbytes / (ms / 1e3) / 1e9 if ms > 0 else 0.0
    return {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "traffic": DIAG_TRAFFIC_PER_UNIT * units if units else None,
            "traffic_source": "ncu --set full capture of 4800 baseline-iterations scaled to this launch (profiles/r02p_diag_path_ncu_raw.csv)",
            "peak_source": src, "kernel": "diag_resident_kernel (cluster-resident; diag_sampler_kernel when a slice does not fit)",
            "byte_model": "Q_d = 8 n T + 16 N bytes per baseline-iteration (read projected data once, read+write S)"}


def run_reference(args, cfg, nsamp):
    """Reference arm: the reference's CPU implementation of the path on the box's host cores (oracle port; the
    reference's own functions where /root/reference exists).  Each step is a bounded sample (2 baselines per core x 10
    iterations); `value` is that sample extrapolated linearly to the workload's nsamp, `ms_per_step` the wall time of
    one sample step as actually run."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals, walls, last = [], [], None
    for i in range(args.warmup + args.steps):
        last = cpu_baseline(cfg, nsamp, all_threads=(i == args.warmup + args.steps - 1))
        if i >= args.warmup:
            vals.append(last["value"])
            walls.append(last["sample_wall_s"])
    v = float(np.mean(vals)) if vals else last["value"]
    last["value"] = v
    nb = args.nbase or cfg.nbase
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(walls)) if walls else None,
        "ms_per_full_step_extrapolated": 1e3 * nb * nsamp / v, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_label(cfg, nb, nsamp),
                   "note": "reference CPU path on all host cores; every step is a BOUNDED SAMPLE of the workload "
                           "(cpu_baseline.sample), value = sample extrapolated linearly in baselines x nsamp"},
        "cpu_baseline": last,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def sha16(t):
    """First 16 hex digits of the SHA-256 of a tensor's / array's bytes."""
    a = t.detach().cpu().numpy() if hasattr(t, "detach") else np.asarray(t)
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def strong_record(args, handle, rank, world, local, sync):
    """ONE step of BASELINE config 4 at fixed total size, divided over the ranks like the reference's baseline axis."""
    import torch
    import torch.distributed as dist

    from draco_b200 import dist as ddist
    from draco_b200 import gibbs, synth

    cfg = synth.CONFIGS["cfg4"]
    total = args.strong_nbase or cfg.nbase
    nsamp = args.strong_nsamp or cfg.nsamp
    group_size = 64
    ngroup = -(-total // group_size)
    lo, hi = ddist.split_range(total, world, rank)
    t0 = time.perf_counter()
    pb = make_device_batch(cfg, lo, hi, local)
    t_gen = time.perf_counter() - t0
    gids = np.arange(lo, hi)
    seed = 4242  # one seed for every N: Philox is keyed on the global baseline id

    def step():
        res = gibbs.run_batch(pb, nsamp, handle=handle, seed=seed)
        avg, cnt = ddist.group_average(res.spectrum, gids, ngroup, group_size, weights=res.success)
        full = ddist.gather_spectra(res.spectrum, total)
        return res, avg, cnt, full

    sync()
    clocks = ClockSampler(local, 500)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res, avg, cnt, full = step()
    e1.record()
    sync()
    clk = clocks.stop()
    ms = torch.tensor([e0.elapsed_time(e1), float(res.timing["ms_sampler"])], dtype=torch.float64, device="cuda")
    nok = torch.tensor([float(res.success.sum().item())], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(nok)
    if rank != 0:
        return None
    # group averages recomputed from the gathered spectra in a fixed order (identical at every N); the all-reduced
    # averages (summation order depends on N) must agree with them to rounding
    fixed = full[: ngroup * group_size].view(ngroup, group_size, -1).mean(dim=1) if total % group_size == 0 else None
    nsm = torch.cuda.get_device_properties(local).multi_processor_count
    nloc = hi - lo
    out = {
        "workload": f"cfg4: {total} baselines TOTAL x {cfg.nfreq} freq (N={cfg.ndelay}), T={cfg.ntime}, nsamp={nsamp}, "
                    f"{ngroup} groups of {group_size}, window={cfg.window}, weights={cfg.weights}",
        "scaling": "strong", "n_gpus": world, "steps": 1,
        "partition": f"contiguous blocks of the baseline axis (dist.split_range), rank 0 owns [{lo}, {hi})",
        "baseline_iter_per_s": total * nsamp / (float(ms[0].item()) / 1e3), "unit": UNIT,
        "ms_step": float(ms[0].item()), "ms_sampler_max": float(ms[1].item()),
        "chains_per_gpu": nloc, "waves_per_gpu_ideal": nloc / nsm,
        "dense_plan": "full waves one CTA per chain + remainder on thread-block clusters (dgibbs_problem.cluster_size = 0)",
        "sampler_launches": int(res.timing["sampler_launches"]),
        "collectives": "NCCL all-reduce of [ngroup, N+1] group sums + all-gather of [B, N] spectra" if world > 1
                       else "none (1 GPU)",
        "chains_ok": int(nok.item()), "inputs": f"device-resident (generated in {t_gen:.1f} s, outside the timed region)",
        "result_checksum": sha16(full), "group_average_checksum": sha16(fixed) if fixed is not None else None,
        "allreduce_vs_fixed_order_max_rel": float(((avg - fixed).abs().max() / fixed.abs().max()).item()) if fixed is not None else None,
        "clocks": clk,
    }
    return out


def int8_gemm_record(device=0, batch=37):
    """The fp64-quality GEMM on the INT8 tensor cores (draco_b200/csrc/ozaki_gemm.cuh: Ozaki split, tcgen05.mma kind::i8, TMEM
    accumulators) that the Wiener delay-transform operator and the NRML objective use for their contractions, through its C-ABI
    entry point with HOST operands: `batch` x [512 x 1024 x 1024], the operator's A = T1 P^H shape at 512 channels.  The timed
    region is the device time of split + GEMM reported by the library (CUDA events around its kernels); the copies are outside.
    A side record: it never raises (an error is returned as text) and never touches the headline numbers."""
    try:
        from draco_b200 import powerspec

        rng = np.random.default_rng(0)
        M, N, K = 512, 1024, 1024
        A, B = rng.standard_normal((batch, M, K)), rng.standard_normal((batch, N, K))
        powerspec.gemm_nt_int8(A[:1], B[:1], device=device)  # library load, kernel attributes
        best = None
        for _ in range(2):
            t = {}
            C = powerspec.gemm_nt_int8(A, B, device=device, timing=t)
            best = t["ms_device"] if best is None else min(best, t["ms_device"])
        ref = A[0, :128] @ B[0, :64].T
        bound = np.abs(A[0, :128]) @ np.abs(B[0, :64]).T
        flop = 2.0 * M * N * K * batch
        return {"what": "C[b] = A[b] B[b]^T, fp64 in / fp64 out, 8 signed 7-bit slices per row x 256 columns on tcgen05.mma kind::i8 "
                        "(36 int8 GEMMs, int32 accumulators in TMEM); dgwiener_gemm_ozaki, split + GEMM device time",
                "shape": [batch, M, N, K], "ms_device": best, "tflops_equivalent_incl_split": flop / best / 1e9,
                "max_diff_vs_numpy_dgemm_over_absA_absB": float((np.abs(C[0, :128, :64] - ref) / bound).max()),
                "reference_points": "GEMM alone 61.3 TFLOP/s-equivalent, cuBLAS DgemmStridedBatched 35.7, fp64 DMMA pipe 37.0 "
                                    "(profiles/r04j_ozaki_gemm.log)"}
    except Exception as exc:  # noqa: BLE001 -- a side record must not cost the bench line
        return {"error": f"{type(exc).__name__}: {exc}"[:300]}


def task_record(args, cfg, nsamp, nbase, rank, world, local, sync, handle):
    """Wall time of DelayPowerSpectrumGibbs.process() on a complex64 / float32 SiderealStream holding this rank's shard:
    device cut + pack + H2D + all chains + D2H ('wall time of _evaluate')."""
    import torch
    import torch.distributed as dist

    from draco_b200 import _compat, delay, synth

    freq, vis, wgt = synth.make_stack(cfg, nbase=nbase, base_offset=rank * nbase)
    vis32, wgt32 = vis.astype(np.complex64), wgt.astype(np.float32)
    del vis, wgt
    # container layout of draco's SiderealStream: vis [freq, stack, ra], weight likewise
    ss = _compat.SiderealStream(freq=freq, stack=nbase, ra=cfg.ntime, native_dtypes=True)
    ss.vis[:] = np.ascontiguousarray(vis32.transpose(2, 0, 1))
    ss.weight[:] = np.ascontiguousarray(wgt32.transpose(2, 0, 1))
    task = delay.DelayPowerSpectrumGibbs()
    task.read_config({"nsamp": nsamp, "sample_axis": "ra", "window": cfg.window or "nuttall",
                      "apply_window": cfg.window is not None, "seed": 31, "device": local})
    task.setup()
    task.process(ss)  # warm-up
    sync()
    n = 1
    t0 = time.perf_counter()
    for _ in range(n):
        out = task.process(ss)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    if rank != 0:
        return None
    spec = np.asarray(getattr(out.spectrum[:], "local_array", out.spectrum[:]))
    return {"value": world * nbase * nsamp * n / float(dt.item()), "unit": UNIT, "steps": n,
            "what": "DelayPowerSpectrumGibbs.process(SiderealStream complex64/float32): flatten_axes + device _cut_data + "
                    "pack + H2D + chains + D2H into the DelaySpectrum container, host wall clock",
            "h2d_bytes_per_step": int(vis32.nbytes + wgt32.nbytes), "d2h_bytes_per_step": int(spec.nbytes),
            "spectrum_finite": bool(np.isfinite(spec).all())}


def main():
    args = parse()
    from draco_b200 import synth

    cfg = synth.CONFIGS[args.config]
    nsamp = args.nsamp or cfg.nsamp
    if args.impl == "reference":
        return run_reference(args, cfg, nsamp)

    import torch
    import torch.distributed as dist

    from draco_b200 import dist as ddist
    from draco_b200 import gibbs

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbase = args.nbase or cfg.nbase  # per rank: weak scaling
    group_size = 64
    ngroup = max(1, -(-world * nbase // group_size))

    pb_host = make_inputs(cfg, nbase, base_offset=rank * nbase)
    pb_dev = pb_host.to_device(local)
    gids = np.arange(rank * nbase, (rank + 1) * nbase)
    stream = torch.cuda.current_stream()
    handle = gibbs.Handle(local, stream=stream.cuda_stream)

    def step_device(i):
        res = gibbs.run_batch(pb_dev, nsamp, handle=handle, seed=1000 + i)
        avg, cnt = ddist.group_average(res.spectrum, gids, ngroup, group_size, weights=res.success)
        return res, avg

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    peak = fp64_peak_tflops()
    for i in range(args.warmup):
        step_device(i)
    sync()
    clocks = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flop = ms_sampler = nbytes = ms_diag = 0.0
    launches = n_diag = 0
    e0.record()
    for i in range(args.steps):
        res, avg = step_device(args.warmup + i)
        flop += res.timing["flop_sampler"]
        ms_sampler += res.timing["ms_sampler"]
        nbytes += res.timing["bytes_sampler"]
        ms_diag += res.timing["ms_sampler_diag"]
        n_diag = int(res.timing["n_diag"])
        launches += int(res.timing["launches"])
    e1.record()
    sync()
    clk = clocks.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    n_ok = int(res.success.sum().item())
    value = world * nbase * nsamp * args.steps / (ms / 1e3)
    weak_checksum = sha16(res.spectrum[: min(nbase, 64)])

    # ---- end to end through the host-pointer C-ABI call (pinned host buffers, H2D + D2H timed)
    pb_pin = pin(pb_host)
    h2d = sum(getattr(pb_pin, f).nbytes for f in ("T", "M", "data_off", "chan_off", "data", "Ni", "fsel", "S0", "global_id"))
    d2h = nbase * cfg.ndelay * 8 + nbase * 5
    e2e_steps = max(1, min(args.steps, args.e2e_steps))

    def step_e2e(i):
        r = gibbs.run_batch(pb_pin, nsamp, handle=handle, seed=2000 + i)
        spec = torch.from_numpy(r.spectrum).to("cuda", non_blocking=True)
        ok = torch.from_numpy(r.success).to("cuda", non_blocking=True)
        avg, _ = ddist.group_average(spec, gids, ngroup, group_size, weights=ok)
        return float(avg[0, 0].item())  # device -> host read of the step's result

    step_e2e(0)
    sync()
    e0.record()
    for i in range(e2e_steps):
        step_e2e(1 + i)
    e1.record()
    sync()
    ms2 = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    e2e_value = world * nbase * nsamp * e2e_steps / (float(ms2.item()) / 1e3)
    del pb_pin

    extra = {}
    skipped = {}
    rate = value / world  # baseline-iterations/s of one GPU, for the cost estimates

    def fits(name, est_s):
        """True if the optional record `name` (estimated `est_s` seconds) still fits the run's wall-clock budget;
        decided on the slowest rank's clock so that every rank takes the same branch (the records hold collectives)."""
        el = torch.tensor([time.perf_counter() - T_START], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(el, op=dist.ReduceOp.MAX)
        ok = budget_allows(float(el.item()), est_s, args.budget_s)
        if not ok:
            skipped[name] = f"skipped: {float(el.item()):.0f} s elapsed + ~{est_s:.0f} s estimated > --budget-s {args.budget_s:.0f}"
        return ok

    if not args.no_task and fits("e2e_task", 25.0 + 2.2 * nbase * nsamp / rate):
        extra["e2e_task"] = task_record(args, cfg, nsamp, nbase, rank, world, local, sync, handle)
    del pb_dev
    torch.cuda.empty_cache()
    c4 = synth.CONFIGS["cfg4"]
    strong_units = (args.strong_nbase or c4.nbase) * (args.strong_nsamp or c4.nsamp) / world
    if not args.no_strong and fits("strong", 20.0 + 0.004 * (args.strong_nbase or c4.nbase) / world + 1.08 * strong_units / rate):
        extra["strong"] = strong_record(args, handle, rank, world, local, sync)
        torch.cuda.empty_cache()
    if not args.no_diag and fits("diag", 15.0 + 45.0 * (args.diag_nbase or synth.CONFIGS["cfg3"].nbase) / 16384 / world):
        import bench_diag

        extra["diag"] = bench_diag.diag_record(args, handle, rank, world, local, sync)
        torch.cuda.empty_cache()

    if world == 1 and not args.no_int8 and fits("int8_gemm", 15.0):
        extra["int8_gemm"] = int8_gemm_record(local)

    if rank == 0:
        ach = flop / (max(ms_sampler, 1e-9) / 1e3) / 1e12
        if n_diag > 0:
            # diagonal(+rank-1) branch: cluster-resident kernel, judged against the HBM stream of its algorithmic bytes (SURVEY 8d Q_d = 8 n T + 16 N per baseline-iteration)
            roof = hbm_roofline(nbytes, ms_diag, units=n_diag * nsamp * args.steps if cfg.ndelay == 2048 else None)
            branch = f"diagonal(+rank-1) cluster-resident sampler ({n_diag} of {nbase} baselines)"
        else:
            roof = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                    "traffic": (DENSE_TRAFFIC_PER_UNIT * nbase * nsamp) if cfg.ndelay == 2048 else None,
                    "traffic_source": DENSE_TRAFFIC_SOURCE,
                    "peak_source": "fp64: builder-measured cuBLAS DGEMM 8192^3 in this run (MEASURED_PEAKS.json has no fp64 entry; "
                                   "tcgen05 has no fp64 kind, DMMA pipe probe = 37.0 TFLOP/s, profiles/r01_fp64_pipe_probe.log)",
                    "kernel": "dense_sampler_kernel", "flop_model": "W_f = n^3/3 + 2 n^2 T + 4 n M T per baseline-iteration"}
            branch = "dense (_f), fp64 DMMA Cholesky"
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {
                "workload": workload_label(cfg, nbase, nsamp),
                "branch": branch, "rng": gibbs.rng_description(),
                "parallelism": f"baseline-sharded x{world}, group-average allreduce ({ngroup} groups of {group_size})",
                "l2": "inputs (1 GiB visibilities + 5 GiB factor workspace) larger than L2; no flush needed",
                "chains_ok": n_ok, "spectrum_checksum_first64": weak_checksum,
            },
            "clocks": clk,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps},
            "gpu_launches": launches,
            "timing_last_step_ms": {k: round(float(v), 3) for k, v in res.timing.items() if k.startswith("ms_")},
            "roofline": roof,
        }
        if extra.get("e2e_task"):
            out["e2e_task"] = extra.pop("e2e_task")
        if extra.get("strong"):
            out["strong"] = extra.pop("strong")
        out["extra"] = {k: v for k, v in extra.items() if v is not None}
        if skipped:
            out["extra"]["skipped_for_time"] = skipped
        if not args.no_cpu_baseline and world == 1:
            out["cpu_baseline"] = cpu_baseline(cfg, nsamp)
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
