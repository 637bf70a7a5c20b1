"""BASELINE config 5 at its stated size on ONE GPU: 40 000 baselines x 2048 freq (N = 4096, T = 16), 40 % of the channels
flagged in contiguous bands redrawn per block of 500 baselines, non-uniform weights, nsamp = 10 -- everything resident
in HBM at once (inputs, whitened copy, every sample, the 148 factor workspaces), one dgibbs_run call.

    python scratch/perf_cfg5_full.py [--nbase 40000] > profiles/r02o_cfg5_full.json.log
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from draco_b200 import gibbs, synth  # noqa: E402


def concat(parts):
    import torch

    off_d, off_c = 0, 0
    doff, coff = [], []
    for p in parts:
        doff.append(p.data_off[:-1] + off_d)
        coff.append(p.chan_off[:-1] + off_c)
        off_d += int(p.data_off[-1])
        off_c += int(p.chan_off[-1])
    dev = parts[0].data.device
    end = lambda v: torch.tensor([v], dtype=torch.int64, device=dev)  # noqa: E731
    cat = lambda name: torch.cat([getattr(p, name) for p in parts])  # noqa: E731
    return gibbs.PackedBatch(parts[0].N, cat("T"), cat("M"), torch.cat(doff + [end(off_d)]), torch.cat(coff + [end(off_c)]),
                             cat("data"), cat("Ni"), cat("fsel"), cat("S0"), cat("global_id"), False, parts[0].window)


def main():
    import torch

    ap = argparse.ArgumentParser()
    ap.add_argument("--nbase", type=int, default=None)
    ap.add_argument("--chunk", type=int, default=2000)
    a = ap.parse_args()
    cfg = synth.CONFIGS["cfg5"]
    nb, nsamp = a.nbase or cfg.nbase, cfg.nsamp
    h = gibbs.Handle(0)
    gibbs.run_batch(bench.make_inputs(cfg, 8, 0).to_device(0), 1, handle=h, seed=0)  # kernel loading
    t0 = time.perf_counter()
    parts = [bench.make_inputs(cfg, min(a.chunk, nb - b0), b0).to_device(0) for b0 in range(0, nb, a.chunk)]
    pb = concat(parts)
    del parts
    torch.cuda.empty_cache()
    t_gen = time.perf_counter() - t0
    M = pb.M.cpu().numpy()
    free0, total = torch.cuda.mem_get_info(0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res = gibbs.run_batch(pb, nsamp, handle=h, seed=5)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    free1, _ = torch.cuda.mem_get_info(0)
    t = res.timing
    print(json.dumps({
        "workload": f"cfg5 at its stated size: {nb} baselines x {cfg.nfreq} freq (N={cfg.ndelay}), T={cfg.ntime}, nsamp={nsamp}, "
                    f"40 % flagged in bands per 500-baseline block (M = {int(M.min())} .. {int(M.max())}), weights perfreq, ONE call, 1 GPU",
        "baseline_iter_per_s": nb * nsamp / (ms / 1e3), "ms_call": ms, "ms_sampler": t["ms_sampler"], "ms_setup": t["ms_setup"],
        "ms_finish": t["ms_finish"], "tflops_sampler": t["flop_sampler"] / t["ms_sampler"] / 1e9,
        "chains_ok": int(res.success.sum().item()), "branches": {int(k): int(v) for k, v in zip(*np.unique(res.branch.cpu().numpy(), return_counts=True))},
        "sampler_launches": int(t["sampler_launches"]),
        "hbm_in_use_gb_during_call": (total - free1) / 1e9, "hbm_in_use_gb_inputs_only": (total - free0) / 1e9, "hbm_total_gb": total / 1e9,
        "input_generation_s": t_gen, "spectrum_checksum_first64": bench.sha16(res.spectrum[:64]),
    }), flush=True)


if __name__ == "__main__":
    main()
