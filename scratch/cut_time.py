import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from draco_b200 import gibbs
r = np.random.default_rng(0)
B, T, F = 1024, 64, 1024
vis = (r.standard_normal((B, T, F)) + 1j * r.standard_normal((B, T, F)))
wgt = r.uniform(0.5, 2.0, (B, T, F)).astype(np.float32)
wgt[:, :, r.choice(F, 100, replace=False)] = 0.0
wgt[:, 5, :] = 0.0
d_vis = torch.from_numpy(vis.view(np.float64).reshape(B, T, F, 2)).cuda()
d_w = torch.from_numpy(wgt).cuda()
for i in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    cr = gibbs.cut_batch(d_vis, d_w, np.arange(F), 2 * F, 10.0, time_frac=0.0, freq_frac=0.0, remove_mean=True)
    print("cut_batch wall ms", (time.perf_counter() - t0) * 1e3)
