#!/usr/bin/env python
"""Headline benchmark: baseline-Gibbs-iterations/s at nfreq=1024 (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config cfg2]

One "step" = one pass of the whole hot path (setup + all nsamp Gibbs iterations + median) over one
batch of synthetic visibilities of the named config; at N GPUs each rank runs that batch on its own
shard of the baseline axis (weak scaling; per-rank work fixed), followed by the NCCL all-reduce of
the baseline-group averages.  `value` has the inputs resident in HBM; `e2e` goes through the
host-pointer C-ABI call (H2D + D2H inside the timed region).  Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

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
    return ap.parse_args()


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


def pin(pb):
    """Copy the batch's arrays into pinned host memory (numpy views of pinned torch tensors)."""
    import dataclasses

    import torch

    kw = {}
    for f in dataclasses.fields(pb):
        v = getattr(pb, f.name)
        if isinstance(v, np.ndarray):
            raw = v.view(np.float64) if v.dtype == np.complex128 else v
            t = torch.empty(raw.shape, dtype=torch.from_numpy(raw[:0].copy()).dtype, pin_memory=True)
            t.numpy()[...] = raw
            v = t.numpy().view(np.complex128) if v.dtype == np.complex128 else t.numpy()
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

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(index)], stdout=subprocess.PIPE, text=True)
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
def _cpu_worker(args):
    os.environ["OPENBLAS_NUM_THREADS"] = os.environ["OMP_NUM_THREADS"] = os.environ["MKL_NUM_THREADS"] = "1"
    cfg_name, b, niter = args
    try:
        from threadpoolctl import threadpool_limits
        ctx = threadpool_limits(1)
    except Exception:  # pragma: no cover
        import contextlib
        ctx = contextlib.nullcontext()
    import oracle
    from draco_b200 import synth

    cfg = synth.CONFIGS[cfg_name]
    with ctx:
        freq, vis, wgt = synth.make_stack(cfg, nbase=1, base_offset=b)
        _, chan = oracle.calculate_delays(freq)
        d, w, nzf, _ = oracle.cut_data(vis[0], wgt[0])
        N = cfg.ndelay
        t0 = time.perf_counter()
        oracle.whiten_inputs(d, N, w, chan[nzf], cfg.window, False)
        t_setup = time.perf_counter() - t0
        t0 = time.perf_counter()
        s, ok = oracle.gibbs_chain(d, N, w, np.full(N, 10.0), cfg.window, chan[nzf], niter, np.random.default_rng(b))
        t_chain = time.perf_counter() - t0
    return t_setup, (t_chain - t_setup) / max(1, len(s)), len(s)


def cpu_baseline(cfg, nsamp, niter_sample=6, procs=None):
    """The oracle (numpy + LAPACK, the reference's own arithmetic) on this box's host cores:
    P single-threaded worker processes, one baseline block each -- the reference's MPI layout
    (no mpirun/mpi4py on the box, so multiprocessing).  Times a bounded sample and extrapolates
    linearly in baselines x nsamp."""
    import multiprocessing as mp

    P = procs or os.cpu_count() or 1
    ctx = mp.get_context("spawn")
    t0 = time.perf_counter()
    with ctx.Pool(P) as pool:
        res = pool.map(_cpu_worker, [(cfg.name, b, niter_sample) for b in range(P)])
    wall = time.perf_counter() - t0
    t_setup = float(np.mean([r[0] for r in res]))
    t_iter = float(np.mean([r[1] for r in res]))
    value = P * nsamp / (t_setup + nsamp * t_iter)  # P baselines advance concurrently
    return {"value": value, "unit": UNIT, "cores": P, "kind": "port",
            "sample": f"{P} baselines x {niter_sample} iterations of {cfg.name} geometry (N={cfg.ndelay}, T={cfg.ntime}), "
                      f"1 BLAS thread per process; t_setup={t_setup:.3f}s t_iter={t_iter:.4f}s, extrapolated to nsamp={nsamp}; "
                      f"sample wall {wall:.1f}s",
            "t_setup_s": t_setup, "t_iter_s": t_iter}


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


# dram__bytes_read.sum + dram__bytes_write.sum of dense_sampler_kernel from the committed `ncu --set full`
# captures profiles/r01e (72.76 GB), r01h (72.89 GB) and r01i_dense_sampler_ncu_raw.csv (66.73 GB read + 6.22 GB written =
# 72.96 GB) for a launch of 296 baseline-iterations of cfg2
# geometry = 245.8 MB per baseline-iteration (L panels re-read from DRAM; 14 % of DRAM peak, not the bound).
# Reported per launch of THIS run by scaling with the launch's baseline-iterations (cfg2 geometry only).
DENSE_TRAFFIC_PER_UNIT = 72.985e9 / 296  # r01o capture: 66.76 GB read + 6.22 GB written


# dram__bytes_read.sum + dram__bytes_write.sum of diag_resident_kernel from the committed `ncu --set full` capture
# profiles/r01n_diag_path_ncu_raw.csv: 257.65 MB read + 61.77 MB written for a launch of 240 baselines x 20 iterations of
# cfg3 geometry (the projected data is read once per baseline and then stays in distributed shared memory).
DIAG_TRAFFIC_PER_UNIT = (257.651712e6 + 61.768704e6) / (240 * 20)


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
            "traffic_source": "ncu --set full capture of 4800 baseline-iterations scaled to this launch (profiles/r01n_diag_path_ncu_raw.csv)",
            "peak_source": src, "kernel": "diag_resident_kernel (cluster-resident; diag_sampler_kernel when a slice does not fit)",
            "byte_model": "Q_d = 8 n T + 16 N bytes per baseline-iteration (read projected data once, read+write S)"}


def run_reference(args, cfg, nsamp):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals, last = [], None
    for i in range(args.warmup + args.steps):
        last = cpu_baseline(cfg, nsamp)
        if i >= args.warmup:
            vals.append(last["value"])
    v = float(np.mean(vals)) if vals else last["value"]
    last["value"] = v
    nb = args.nbase or cfg.nbase
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * nb * nsamp / v, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{cfg.name}: {nb} baselines x {cfg.nfreq} freq (N={cfg.ndelay}), T={cfg.ntime}, nsamp={nsamp}",
                   "note": "reference CPU path (numpy + scipy LAPACK restatement; the Python reference cannot travel) on all host cores"},
        "cpu_baseline": last,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


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

    # ---- end to end through the host-pointer C-ABI call (pinned host buffers, H2D + D2H timed)
    pb_pin = pin(pb_host)
    h2d = sum(getattr(pb_pin, f).nbytes for f in ("T", "M", "data_off", "chan_off", "data", "Ni", "fsel", "S0", "global_id"))
    d2h = nbase * cfg.ndelay * 8 + nbase * 5

    def step_e2e(i):
        r = gibbs.run_batch(pb_pin, nsamp, handle=handle, seed=2000 + i)
        spec = torch.from_numpy(r.spectrum).to("cuda", non_blocking=True)
        ok = torch.from_numpy(r.success).to("cuda", non_blocking=True)
        avg, _ = ddist.group_average(spec, gids, ngroup, group_size, weights=ok)
        return float(avg[0, 0].item())  # device -> host read of the step's result

    step_e2e(0)
    sync()
    e0.record()
    for i in range(args.steps):
        step_e2e(1 + i)
    e1.record()
    sync()
    ms2 = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    e2e_value = world * nbase * nsamp * args.steps / (float(ms2.item()) / 1e3)

    if rank == 0:
        ach = flop / (max(ms_sampler, 1e-9) / 1e3) / 1e12
        if n_diag > 0:
            # diagonal(+rank-1) branch: cluster-resident kernel, judged against the HBM stream of its algorithmic bytes (SURVEY 8d Q_d = 8 n T + 16 N per baseline-iteration)
            roof = hbm_roofline(nbytes, ms_diag, units=n_diag * nsamp if cfg.ndelay == 2048 else None)
            branch = f"diagonal(+rank-1) cluster-resident sampler ({n_diag} of {nbase} baselines)"
        else:
            roof = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                    "traffic": (DENSE_TRAFFIC_PER_UNIT * nbase * nsamp) if cfg.ndelay == 2048 else None,
                    "traffic_source": "ncu --set full capture of 296 baseline-iterations scaled to this launch (profiles/r01o_dense_sampler_ncu_raw.csv)",
                    "peak_source": "fp64: cuBLAS DGEMM 8192^3 measured in this run (MEASURED_PEAKS.json has no fp64 entry; "
                                   "tcgen05 has no fp64 kind, DMMA pipe probe = 37.0 TFLOP/s, profiles/r01_fp64_pipe_probe.log)",
                    "kernel": "dense_sampler_kernel", "flop_model": "W_f = n^3/3 + 2 n^2 T + 4 n M T per baseline-iteration"}
            branch = "dense (_f), fp64 DMMA Cholesky"
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {
                "workload": f"{cfg.name}: {nbase} baselines/GPU x {cfg.nfreq} freq (N={cfg.ndelay}), T={cfg.ntime}, "
                            f"nsamp={nsamp}, window={cfg.window}, weights={cfg.weights}, mask={cfg.mask}",
                "branch": branch, "rng": "device Philox4x32-10",
                "parallelism": f"baseline-sharded x{world}, group-average allreduce ({ngroup} groups of {group_size})",
                "l2": "inputs (1 GiB visibilities + 5 GiB factor workspace) larger than L2; no flush needed",
                "chains_ok": n_ok,
            },
            "clocks": clk,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
            "gpu_launches": launches,
            "timing_last_step_ms": {k: round(float(v), 3) for k, v in res.timing.items() if k.startswith("ms_")},
            "roofline": roof,
        }
        if not args.no_cpu_baseline and world == 1:
            out["cpu_baseline"] = cpu_baseline(cfg, nsamp)
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
