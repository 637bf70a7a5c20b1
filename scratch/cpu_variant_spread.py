"""How far two correct LAPACK solutions of the cfg5-geometry system are from each other (first Gibbs iteration, flat S0):
upper Cholesky (the reference's call) vs lower Cholesky vs LU, compared on the quantity the samples are made of (the
variance of the signal draws over time).  Pure CPU."""
import os
import sys

import numpy as np
import scipy.linalg as la

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from oracle import gibbs_oracle as go  # noqa: E402
from draco_b200 import synth  # noqa: E402

cfg = synth.CONFIGS["cfg5"]
for b, off in enumerate((0, 500)):
    freq, vis, wgt = synth.make_stack(cfg, nbase=1, base_offset=off)
    delays, chan = oracle.calculate_delays(freq)
    N = len(delays)
    d, w, nzf, _ = oracle.cut_data(vis[0], wgt[0])
    z, FTNih, FTNiF = go.whiten_inputs(d, N, w, chan[nzf], cfg.window, False)
    S = np.full(N, 10.0)
    Si = go.invert_no_zero(S)
    Ci = FTNiF.copy()
    Ci[np.diag_indices_from(Ci)] += Si
    rng = np.random.default_rng(500 + b)
    w1 = rng.standard_normal((N, z.shape[1]))
    w2 = rng.standard_normal(z.shape)
    rhs = w1 * (Si**0.5)[:, None] + FTNih @ (z + w2)
    xs = {
        "potrf upper (reference)": la.cho_solve(la.cho_factor(Ci, lower=False), rhs),
        "potrf lower": la.cho_solve(la.cho_factor(Ci, lower=True), rhs),
        "getrf (LU)": la.solve(Ci, rhs),
    }
    ref = xs["potrf upper (reference)"].var(axis=1)
    ev = np.linalg.eigvalsh(Ci)
    print(f"baseline {b}: N={N} M={d.shape[1]} T={z.shape[1]}  cond(Ci)={ev[-1] / ev[0]:.2e}")
    for k, x in xs.items():
        rel = np.abs(x.var(axis=1) - ref) / ref
        print(f"   {k:26s} max rel diff of var_t(x): {rel.max():.2e}   median {np.median(rel):.1e}   > 1e-6: {(rel > 1e-6).sum()} of {N}")
