"""Small runs of every sampler kernel for compute-sanitizer (memcheck / racecheck / synccheck)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _cases import make_case  # noqa: E402
from draco_b200 import gibbs  # noqa: E402

which = sys.argv[1:] or ["dense256", "dense256c2", "dense512", "freq", "cross", "diag"]
r = np.random.default_rng(0)
if "dense256" in which or "dense256c2" in which:
    N, M, T = 256, 102, 32
    data, Ni, fsel, S0 = make_case(N, M, T, False, False, seed=1)
    pb = gibbs.pack_batch([data, data], [Ni, Ni * 1.1], [fsel, fsel], N, np.full((2, N), 10.0), window="nuttall")
    for cs in ([1] if "dense256" in which else []) + ([2] if "dense256c2" in which else []):
        res = gibbs.run_batch(pb, 2, seed=3, cluster_size=cs, force_branch=1)
        print("dense N=256 cluster", cs, "ok", res.success.tolist(), flush=True)
if "dense512" in which:
    N, M, T = 512, 200, 16
    data, Ni, fsel, S0 = make_case(N, M, T, False, False, seed=2)
    pb = gibbs.pack_batch([data], [Ni], [fsel], N, np.full((1, N), 10.0), window="nuttall")
    res = gibbs.run_batch(pb, 1, seed=4, cluster_size=1, force_branch=1)
    print("dense N=512 ok", res.success.tolist(), flush=True)
if "freq" in which:
    N, M, T = 256, 40, 8
    data, Ni, fsel, S0 = make_case(N, M, T, False, False, seed=5)
    pb = gibbs.pack_batch([data], [Ni], [fsel], N, np.full((1, N), 10.0), window="nuttall")
    res = gibbs.run_batch(pb, 2, seed=6, force_branch=2)
    print("freq branch", res.branch.tolist(), "ok", res.success.tolist(), flush=True)
if "cross" in which:
    nd, N, T, M, B = 2, 64, 8, 64, 2
    data = r.standard_normal((B, nd, T, M)) + 1j * r.standard_normal((B, nd, T, M))
    Ni = r.uniform(0.5, 2.0, (B, nd, M)) * 10
    S0 = np.zeros((nd, nd, N), dtype=complex)
    S0[0, 0] = S0[1, 1] = 10.0
    s, ndone, ok = gibbs.cross_batch(data, N, Ni, S0, "nuttall", np.arange(M), 2, seed=7)
    print("cross ok", ok.tolist(), flush=True)
if "diag" in which:
    N, M, T = 256, 129, 16
    data = r.standard_normal((4, T, M)) + 1j * r.standard_normal((4, T, M))
    pb = gibbs.pack_uniform(data, np.ones((4, M)), np.arange(M), N, np.full((4, N), 10.0), window="nuttall")
    res = gibbs.run_batch(pb, 3, seed=8)
    print("diag branch", res.branch.tolist(), "ok", res.success.tolist(), flush=True)
