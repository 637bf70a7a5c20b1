"""Device timings of the f3 / f4 rows: the device-side cut at cfg2 size against the host loop, and one NRML
likelihood evaluation at N = 2048, M = 1024 against the numpy oracle."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tests/golden")
import oracle
from oracle import nrml_oracle
from draco_b200 import delayopt, gibbs
from make_golden_nrml import make_nrml_case

out = {}
# ---- f3: cut 1024 baselines x 64 samples x 1024 channels
r = np.random.default_rng(0)
B, T, F = 1024, 64, 1024
vis = (r.standard_normal((B, T, F)) + 1j * r.standard_normal((B, T, F)))
wgt = r.uniform(0.5, 2.0, (B, T, F)).astype(np.float32)
wgt[:, :, r.choice(F, 100, replace=False)] = 0.0
wgt[:, 5, :] = 0.0
cind = np.arange(F)
gibbs.cut_batch(vis[:8], wgt[:8], cind, 2 * F, 10.0, time_frac=0.0, freq_frac=0.0, remove_mean=True)
t0 = time.perf_counter()
cr = gibbs.cut_batch(vis, wgt, cind, 2 * F, 10.0, time_frac=0.0, freq_frac=0.0, remove_mean=True)
t_dev = time.perf_counter() - t0
import torch
d_vis = torch.from_numpy(vis.view(np.float64).reshape(B, T, F, 2)).cuda()
d_w = torch.from_numpy(wgt).cuda()
torch.cuda.synchronize()
t0 = time.perf_counter()
cr2 = gibbs.cut_batch(d_vis, d_w, cind, 2 * F, 10.0, time_frac=0.0, freq_frac=0.0, remove_mean=True)
t_res = time.perf_counter() - t0
t0 = time.perf_counter()
nhost = 64
for b in range(nhost):
    oracle.cut_data(vis[b], wgt[b])
t_host = (time.perf_counter() - t0) * B / nhost
out["cut_B1024_T64_F1024"] = {"device_incl_h2d_s": t_dev, "device_resident_inputs_s": t_res, "host_numpy_loop_s_extrapolated": t_host,
                              "alive": int(cr.alive.sum())}
# ---- f4: one likelihood evaluation (value + gradient + Hessian) at N = 2048, M = 1024
N, M, T = 2048, 1024, 64
data, Ni, fsel = make_nrml_case(N, M, T, 20, 3)
_, X, Fm, Nm = nrml_oracle.maxpost_inputs(data, N, Ni, "nuttall", fsel)
like = delayopt.LogLikePS(X, Fm, Nm, T, bounds=(1e-15, 1e10))
x = np.random.default_rng(1).normal(0.0, 1.0, N)
like.value(x); like.gradient(x); like.hessian(x)
like.eval_ms = like.eval_flop = 0.0
t0 = time.perf_counter()
reps = 5
for i in range(reps):
    xi = x + 1e-3 * i
    v = like.value(xi); g = like.gradient(xi); H = like.hessian(xi)
wall = (time.perf_counter() - t0) / reps
t0 = time.perf_counter()
ora = nrml_oracle.LogLike(X, Fm, Nm, T, bounds=(1e-15, 1e10))
vo = ora.value(xi); go = ora.gradient(xi); Ho = ora.hessian(xi)
t_cpu = time.perf_counter() - t0
out["nrml_eval_N2048_M1004"] = {"device_ms_events": like.eval_ms / reps, "wall_ms_incl_hessian_d2h": wall * 1e3,
                                "tflops_device": like.eval_flop / (like.eval_ms * 1e-3) / 1e12, "oracle_cpu_s": t_cpu,
                                "value_rel_err": abs(v - vo) / abs(vo), "grad_rel_err": float(np.abs(g - go).max() / np.abs(go).max()),
                                "hess_rel_err": float(np.abs(H - Ho).max() / np.abs(Ho).max())}
print(json.dumps(out))
