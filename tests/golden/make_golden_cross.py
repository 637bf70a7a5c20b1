"""Generate tests/golden/reference_cross.npz by running the REFERENCE's own cross-spectrum sampler.

Build-container only (needs /root/reference): executes ``delay_spectrum_gibbs_cross``
(draco/analysis/delay.py:1907-2099) through ``oracle/ref_loader.py`` with the complex normal / Wishart
helpers of draco/util/random.py, and stores every sample of short seeded chains.  The inputs are
re-created from the same seeds by the tests (``make_cross_case``); the draws are re-created by seeding
numpy identically and drawing in the reference's order.

    python tests/golden/make_golden_cross.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

# name, nd, N, M, T, window, niter
CROSS_CASES = [
    ("x_nd2_n32", 2, 32, 32, 8, "nuttall", 4),
    ("x_nd2_n64_masked", 2, 64, 40, 6, "blackman_harris", 3),
    ("x_nd3_n24_nowin", 3, 24, 20, 7, None, 3),
    ("x_nd1_n16", 1, 16, 12, 5, "hann", 3),
    ("x_nd4_n16", 4, 16, 16, 9, "nuttall", 2),
]
DRAW_SEED = 7


def make_cross_case(nd, N, M, T, seed):
    """Seeded inputs of one cross-spectrum problem: correlated datasets, per-dataset weights."""
    r = np.random.default_rng(seed)
    fsel = np.sort(r.choice(N, size=M, replace=False))
    common = r.standard_normal((T, M)) + 1j * r.standard_normal((T, M))
    data = np.stack([common * (1.0 + 0.3 * d) + 0.5 * (r.standard_normal((T, M)) + 1j * r.standard_normal((T, M)))
                     for d in range(nd)])
    Ni = r.uniform(0.5, 2.0, (nd, M)) * 10
    S0 = np.zeros((nd, nd, N), dtype=np.complex128)
    for d in range(nd):
        S0[d, d] = 10.0
    return data, Ni, fsel, S0


def main():
    from oracle import ref_loader

    ref = ref_loader.load()
    out = {}
    for name, nd, N, M, T, window, niter in CROSS_CASES:
        data, Ni, fsel, S0 = make_cross_case(nd, N, M, T, seed=31)
        spec = ref.delay_spectrum_gibbs_cross(data.copy(), N, Ni, S0.copy(), window=window, fsel=fsel, niter=niter,
                                              rng=np.random.default_rng(DRAW_SEED))
        out[f"{name}_samples"] = np.array(spec)
    # a non-positive-definite prior makes the reference raise (delay.py:2096-2097)
    data, Ni, fsel, S0 = make_cross_case(2, 32, 32, 8, seed=31)
    try:
        ref.delay_spectrum_gibbs_cross(data.copy(), 32, Ni, -S0, window="nuttall", fsel=fsel, niter=2,
                                       rng=np.random.default_rng(DRAW_SEED))
        raised = ""
    except RuntimeError as e:
        raised = str(e)
    out["x_fail_message"] = np.array(raised)
    path = os.path.join(HERE, "reference_cross.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
