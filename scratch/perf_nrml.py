"""End-to-end NRML (delay_power_spectrum_maxpost) for one baseline at N = 2048, M = 1024 (20 flagged), T = 64:
device objective + device pseudo-inverse start against the numpy/scipy oracle (= the reference's algorithm)."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tests/golden")
from oracle import nrml_oracle
from draco_b200 import delayopt
from make_golden_nrml import make_nrml_case

N, M, T, maxiter = 2048, 1024, 64, 8
data, Ni, fsel = make_nrml_case(N, M, T, 20, 3)
delayopt.delay_power_spectrum_maxpost(data[:, :64], 128, Ni[:64], None, window="nuttall", fsel=np.arange(64), maxiter=2)  # warm-up
out = {}
import os
# PERF_NRML_ONLY=host_pinv: only the objective's own path (the device starting point uses cuSOLVER's eigensolver), no oracle
only = os.environ.get("PERF_NRML_ONLY")
for tag, kw in (("device_start", {}), ("host_pinv", dict(host_pinv=True))):
    if only and tag != only:
        continue
    st = {}
    t0 = time.perf_counter()
    s, ok = delayopt.delay_power_spectrum_maxpost(data, N, Ni, None, window="nuttall", fsel=fsel, maxiter=maxiter, tol=1e-3, stats=st, **kw)
    out[tag] = {"wall_s": time.perf_counter() - t0, "iterates": len(s), "success": bool(ok), "device_ms": st["ms"], "evals": st["evals"],
                "tflops_device": st["flop"] / (st["ms"] * 1e-3) / 1e12}
    last = s[-1]
if only:
    print(json.dumps(out))
    sys.exit(0)
t0 = time.perf_counter()
so, oko = nrml_oracle.maxpost_chain(data, N, Ni, None, window="nuttall", fsel=fsel, maxiter=maxiter, tol=1e-3)
out["oracle_cpu"] = {"wall_s": time.perf_counter() - t0, "iterates": len(so), "success": bool(oko)}
out["max_rel_diff_last_iterate"] = float(np.max(np.abs(last - so[-1]) / np.abs(so[-1]))) if len(so) == len(s) else None
print(json.dumps(out))
