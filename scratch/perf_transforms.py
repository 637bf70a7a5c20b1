"""Timing of the plain FFT delay transform and of the sample-variance reduction (device-resident data, CUDA events
recorded by the library around the kernel), with the numpy oracle on the host beside them."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import oracle  # noqa: E402
from draco_b200 import gibbs  # noqa: E402
from draco_b200 import _lib  # noqa: E402
import ctypes as C  # noqa: E402

out = {}
h = gibbs.get_handle(0)
for N, nrows in ((1024, 65536), (2048, 32768), (4096, 16384), (256, 262144), (1000, 4096)):
    x = torch.randn(nrows, N, 2, dtype=torch.float64, device="cuda")
    t = {}
    for _ in range(3):
        gibbs.fft_rows(x, "nuttall", shift=True, handle=h, timing=t)
    ms = min(gibbs.fft_rows(x, "nuttall", shift=True, handle=h, timing=t) is None or t["ms_kernel"] for _ in range(5))
    nbytes = 2.0 * nrows * N * 16
    xs = x[:2048].cpu().numpy().reshape(2048, 2 * N).view(np.complex128)
    t0 = time.perf_counter()
    oracle.delay_spectrum_fft(xs, N, "nuttall")
    cpu = (time.perf_counter() - t0) / 2048 * nrows
    out[f"fft_N{N}"] = {"rows": nrows, "ms_kernel": ms, "GB/s": nbytes / ms / 1e6, "numpy_1thread_ms_extrapolated": cpu * 1e3}
    del x
B, T, N = 1024, 64, 1024
spec = torch.randn(B, T, N, 2, dtype=torch.float64, device="cuda")
mask = (torch.rand(B, T, device="cuda") < 0.2).to(torch.uint8)
var = torch.empty(B, N, dtype=torch.float64, device="cuda")
bad = torch.zeros(B, dtype=torch.uint8, device="cuda")
ms = C.c_double(0.0)
best = 1e9
for _ in range(6):
    torch.cuda.synchronize()
    h.check(h.lib.dgibbs_sample_variance(h._h, 1, B, T, N, spec.data_ptr(), mask.data_ptr(), var.data_ptr(), bad.data_ptr(),
                                         C.byref(ms)), "var")
    best = min(best, ms.value)
hs = spec[:64].cpu().numpy().reshape(64, T, 2 * N).view(np.complex128)
hm = mask[:64].cpu().numpy().astype(bool)
t0 = time.perf_counter()
oracle.delay_spectrum_to_power_spectrum(hs, hm)
cpu = (time.perf_counter() - t0) / 64 * B
out["sample_variance_1024x64x1024"] = {"ms_kernel": best, "GB/s": B * T * N * 16 / best / 1e6, "numpy_ms_extrapolated": cpu * 1e3}
print(json.dumps(out, indent=1))
