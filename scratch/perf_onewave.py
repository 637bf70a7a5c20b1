"""One full wave of cfg2 chains (one CTA per chain on every SM): ms per Gibbs iteration of the library named by
DGIBBS_LIB_PATH (default: the in-tree build).  Prints one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from draco_b200 import gibbs, synth  # noqa: E402

cfg = synth.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "cfg2"]
nsamp = int(sys.argv[2]) if len(sys.argv) > 2 else 4
import torch  # noqa: E402

nsm = torch.cuda.get_device_properties(0).multi_processor_count
pb = bench.make_inputs(cfg, nsm, 0).to_device(0)
h = gibbs.Handle(0)
gibbs.run_batch(pb, 1, handle=h, seed=1, cluster_size=1)
best = 1e30
for rep in range(3):
    r = gibbs.run_batch(pb, nsamp, handle=h, seed=2 + rep, cluster_size=1)
    best = min(best, r.timing["ms_sampler"] / nsamp)
print(json.dumps({"lib": os.environ.get("DGIBBS_LIB_PATH", "in-tree"), "no_tma": os.environ.get("DGIBBS_NO_TMA"), "config": cfg.name,
                  "chains": nsm, "ms_per_iteration": best, "baseline_iter_per_s": nsm / best * 1e3,
                  "tflops": r.timing["flop_sampler"] / nsamp / best / 1e9}))
