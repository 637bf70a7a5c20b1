"""Distribution check of the device RNG paths after a change of generator: dense kernel vs diagonal kernel vs the CPU
oracle on one uniform-weight geometry, many chains, KS p-values at every delay (expected: uniform p-values)."""
import os
import sys

import numpy as np
from scipy import stats

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402
from draco_b200 import gibbs  # noqa: E402
from test_gpu_diag import uniform_case  # noqa: E402

N, T, nb, niter = 128, 64, 512, 60
data, Ni, fsel, S0 = uniform_case(N, T, False, False, seed=N + T)
pb = gibbs.pack_batch([data] * nb, [Ni] * nb, [fsel] * nb, N, np.full((nb, N), 10.0), window=None, global_id=np.arange(nb)).to_device(0)
a = gibbs.run_batch(pb, niter, seed=15, save_samples=True)
b = gibbs.run_batch(pb, niter, seed=16, save_samples=True, force_branch=gibbs.BRANCH_DELAY_SPACE)
sa = np.log(a.samples[-1].cpu().numpy())
sb = np.log(b.samples[-1].cpu().numpy())
so = []
for c in range(nb):
    s, ok = oracle.gibbs_chain(data, N, Ni, S0, None, fsel, niter, np.random.default_rng(9000 + c))
    so.append(np.log(s[-1]))
so = np.array(so)
for name, x, y in (("diag vs dense", sa, sb), ("dense vs oracle", sb, so), ("diag vs oracle", sa, so)):
    p = np.array([stats.ks_2samp(x[:, k], y[:, k]).pvalue for k in range(N)])
    z = (x.mean(0) - y.mean(0)) / np.sqrt(x.var(0, ddof=1) / nb + y.var(0, ddof=1) / nb)
    print(f"{name}: KS p-values over {N} delays: min {p.min():.2e}, frac<0.01 {np.mean(p < 0.01):.3f}, frac<0.1 {np.mean(p < 0.1):.3f}, "
          f"median {np.median(p):.2f};  mean-diff z: max|z| {np.abs(z).max():.2f}, frac|z|<2 {np.mean(np.abs(z) < 2):.3f}; "
          f"KS of p against U(0,1): {stats.kstest(p, 'uniform').pvalue:.3f}")
