"""Wiener delay-transform operator: device time per pixel against the oracle on the host cores.

    python scratch/perf_wiener_op.py [F nra nel]      (prints one JSON line per configuration)
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden"))
from make_golden_wiener_op import make_wiener_case  # noqa: E402

from draco_b200 import powerspec  # noqa: E402
from oracle import wiener_op_oracle as wo  # noqa: E402


def problem(F, nra, nel, cplx, seed=1):
    case = make_wiener_case(F, nra, nel, 1, cplx, seed)
    freq = case["freq"]
    win = wo.get_window(freq, "blackman_harris")
    tau, ntau = wo.delay_axis(freq, win)
    S = wo.get_prior(tau, 2.8e-5, 0.02)
    filt = np.ascontiguousarray(case["filter"][0].transpose(2, 0, 1))
    cov = np.ascontiguousarray(case["freq_cov"][0].transpose(2, 0, 1))
    weight = np.ascontiguousarray(case["weight"][0].transpose(2, 1, 0))
    beam = np.ascontiguousarray(case["dirty_beam_power"][0, 0].T)
    return freq, weight, beam, filt, cov, win, S, tau, ntau


def run(F, nra, nel, cplx, cpu_pixels):
    freq, weight, beam, filt, cov, win, S, tau, ntau = problem(F, nra, nel, cplx)
    Fm = wo.fourier_operator(freq, tau, ntau)
    res = {"F": F, "ndelay": int(tau.size), "nra": nra, "nel": nel, "complex_inputs": cplx}
    solvers = ((1, "batched"), (2, "streams"), (3, "own_kernels"), (4, "own_kernels_int8"))
    if os.environ.get("PERF_WIENER_ONLY"):  # e.g. PERF_WIENER_ONLY=3 for a launch list of the own-kernel path alone
        solvers = tuple(x for x in solvers if str(x[0]) == os.environ["PERF_WIENER_ONLY"])
    for solver, label in solvers:
        best = None
        for rep in range(3):
            t = {}
            t0 = time.perf_counter()
            got = powerspec.wiener_delay_operator(Fm, S, win, filt, cov, weight, beam, solver=solver, timing=t)
            t["wall_ms"] = (time.perf_counter() - t0) * 1e3
            if best is None or t["ms_compute"] < best["ms_compute"]:
                best = t
        res[label] = {"ms_compute": round(best["ms_compute"], 3), "ms_device": round(best["ms_device"], 3),
                      "wall_ms": round(best["wall_ms"], 2),
                      "us_per_pixel": round(best["ms_compute"] * 1e3 / (nra * nel), 2),
                      "tflops": round(best["flop"] / best["ms_compute"] / 1e9, 2), "launches": best["launches"]}
    if os.environ.get("PERF_WIENER_ONLY"):
        print(json.dumps(res), flush=True)
        return
    # oracle on the host (numpy + LAPACK, all BLAS threads), a bounded number of pixels
    nr = max(1, min(nra, cpu_pixels))
    t0 = time.perf_counter()
    ref = wo.construct_operator(freq, weight[:1, :nr], beam[:1], filt[:nr], cov[:nr], win, S, tau, ntau)
    dt = time.perf_counter() - t0
    res["oracle_cpu"] = {"pixels": nr, "ms_per_pixel": round(dt * 1e3 / nr, 3), "cores": os.cpu_count()}
    err = np.abs(got[:nr, 0] - ref[:, 0]).max() / np.abs(ref).max()
    res["max_rel_err_vs_oracle"] = float(err)
    k = "streams" if "streams" in res else "batched"
    res["speedup_compute"] = round(res["oracle_cpu"]["ms_per_pixel"] * 1e3 / min(res[x]["us_per_pixel"] for x in ("batched", "streams", "own_kernels", "own_kernels_int8") if x in res), 1)
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    if len(sys.argv) >= 4:
        run(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), False, 16)
    else:
        run(64, 1024, 4, False, 64)
        run(128, 512, 4, False, 32)
        run(256, 256, 2, False, 16)
        run(256, 256, 2, True, 16)
        run(512, 96, 2, False, 8)
        run(1024, 32, 1, False, 4)
