"""Chain shared by a cluster of C CTAs: duration of one chain and of a launch with a partial last wave.

    python scratch/perf_cluster.py [--nsamp 4] > profiles/r02a_cluster_split.json.log

cfg2 geometry (N=2048, T=64, M=1024).  Prints one JSON line per measurement (CUDA-event time of the sampler
launches, `ms_sampler` of dgibbs_timing)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from draco_b200 import gibbs, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nsamp", type=int, default=4)
    ap.add_argument("--config", default="cfg2")
    ap.add_argument("--tail", type=int, default=512, help="chains of the partial-wave measurement")
    a = ap.parse_args()
    cfg = synth.CONFIGS[a.config]
    import torch

    nsm = torch.cuda.get_device_properties(0).multi_processor_count
    h = gibbs.Handle(0)
    pb_all = bench.make_inputs(cfg, a.tail, 0).to_device(0)

    def sub(n):
        import dataclasses

        kw = {f.name: getattr(pb_all, f.name) for f in dataclasses.fields(pb_all)}
        for k in ("T", "M", "global_id"):
            kw[k] = kw[k][:n]
        kw["data_off"], kw["chan_off"], kw["S0"] = kw["data_off"][: n + 1], kw["chan_off"][: n + 1], kw["S0"][:n]
        return gibbs.PackedBatch(**kw)

    base = None
    for C, n in ((1, nsm), (2, nsm // 2), (4, nsm // 4), (1, nsm // 2), (1, nsm // 4)):
        pb = sub(n)
        gibbs.run_batch(pb, 1, handle=h, seed=1, cluster_size=C)
        r = gibbs.run_batch(pb, a.nsamp, handle=h, seed=2, cluster_size=C)
        ms = r.timing["ms_sampler"] / a.nsamp
        if C == 1 and base is None:
            base = ms
        print(json.dumps({"what": "one wave", "config": a.config, "cluster_size": C, "chains": n, "ms_per_iteration": ms,
                          "relative_chain_time": ms / base, "ok": int(r.success.sum())}), flush=True)
    for C in (1, 0):
        gibbs.run_batch(pb_all, 1, handle=h, seed=1, cluster_size=C)
        r = gibbs.run_batch(pb_all, a.nsamp, handle=h, seed=2, cluster_size=C)
        ms = r.timing["ms_sampler"] / a.nsamp
        print(json.dumps({"what": "partial last wave", "config": a.config, "cluster_size": C or "planned", "chains": a.tail,
                          "ms_per_iteration": ms, "waves_ideal": a.tail / nsm, "waves_measured": ms / base,
                          "sampler_launches": int(r.timing["sampler_launches"]),
                          "baseline_iter_per_s": a.tail / (ms / 1e3)}), flush=True)


if __name__ == "__main__":
    main()
