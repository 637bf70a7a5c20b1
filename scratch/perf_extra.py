"""Extra device timings (not the headline): the _t branch against the delay-space system on heavily masked
data, and the cross-spectrum sampler at a realistic size."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from draco_b200 import gibbs

out = {}
# ---- _t branch: N=2048, 80 % of the channels flagged (M=205 -> 2M=410), T=64
N, T, nb, niter = 2048, 64, 296, 10
r = np.random.default_rng(0)
for M in (205, 410):
    fsel = np.sort(r.choice(N // 2, size=M, replace=False))
    data = r.standard_normal((nb, T, M)) + 1j * r.standard_normal((nb, T, M))
    Ni = r.uniform(0.5, 2.0, (nb, M)) * 100
    pb = gibbs.pack_uniform(data, Ni, fsel, N, np.full((nb, N), 10.0), window="nuttall", global_id=np.arange(nb)).to_device(0)
    for name, fb in (("freq", 0), ("delay", 1)):
        gibbs.run_batch(pb, 2, seed=1, force_branch=fb)
        res = gibbs.run_batch(pb, niter, seed=2, force_branch=fb)
        ms = res.timing["ms_sampler"]
        out[f"t_branch_M{M}_{name}"] = {"branch": int(res.branch[0]), "ms": ms, "baseline_iter_per_s": nb * niter / ms * 1e3,
                                        "ok": int(res.success.sum())}
# ---- cross sampler: nd=2, N=1024 (real embedding 4096), T=32, full band
nd, N, T, M, B, niter = 2, 1024, 32, 1024, 148, 3
data = r.standard_normal((B, nd, T, M)) + 1j * r.standard_normal((B, nd, T, M))
Ni = r.uniform(0.5, 2.0, (B, nd, M)) * 10
S0 = np.zeros((nd, nd, N), dtype=complex)
S0[0, 0] = S0[1, 1] = 10.0
gibbs.cross_batch(data[:8], N, Ni[:8], S0, "nuttall", np.arange(M), 1, seed=1)
t0 = time.perf_counter()
s, nd_, ok = gibbs.cross_batch(data, N, Ni, S0, "nuttall", np.arange(M), niter, seed=2)
dt = time.perf_counter() - t0
n = 2 * nd * N
flop = B * niter * (n**3 / 3 + 2 * n * n * T)
out["cross_nd2_N1024"] = {"wall_s_incl_copies": dt, "problem_iter_per_s": B * niter / dt, "tflops_lower_bound": flop / dt / 1e12,
                          "ok": int(ok.sum())}
print(json.dumps(out, indent=1))
