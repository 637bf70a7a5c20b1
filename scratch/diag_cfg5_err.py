"""Per-iteration error of the cfg5-geometry parity chain against the oracle (injected draws): how fast the rounding
differences of two correct fp64 implementations grow along an ill-conditioned chain."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402
from _cases import recorded_chain  # noqa: E402
from draco_b200 import gibbs, synth  # noqa: E402

cfg = synth.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "cfg5"]
niter = int(sys.argv[2]) if len(sys.argv) > 2 else 4
datas, Nis, fsels, refs, draws = [], [], [], [], []
for b, off in enumerate((0, 500)):
    freq, vis, wgt = synth.make_stack(cfg, nbase=1, base_offset=off)
    delays, chan = oracle.calculate_delays(freq)
    N = len(delays)
    d, w, nzf, _ = oracle.cut_data(vis[0], wgt[0])
    s, ok, dr = recorded_chain(d, N, w, np.full(N, 10.0), cfg.window, chan[nzf], niter, False, seed=500 + b)
    datas.append(d); Nis.append(w); fsels.append(chan[nzf]); refs.append(s); draws.append(dr)
pb = gibbs.pack_batch(datas, Nis, fsels, N, np.full((2, N), 10.0), window=cfg.window)
res = gibbs.run_batch(pb, niter, draws=gibbs.pack_draws(pb, niter, *zip(*draws)), save_samples=True)
for b in range(2):
    rel = np.abs(res.samples[:, b] - refs[b]) / np.abs(refs[b])
    print(os.environ.get("DGIBBS_LIB_PATH", "in-tree"), cfg.name, "baseline", b, "max rel err per iteration:",
          " ".join(f"{x:.2e}" for x in rel.max(axis=1)), " median:", " ".join(f"{x:.1e}" for x in np.median(rel, axis=1)))
