"""Golden vectors of the REFERENCE's ``_cut_data`` (draco/analysis/delay.py:516-597) on the dtypes draco's containers
really hold: complex64 visibilities and float32 weights (draco/core/containers.py:503-520).  numpy then removes the mean
in complex64 arithmetic (float32 pairwise column sums, complex128 division by the count, float32 subtraction) and
averages the weights in float32; the device-side cut has to reproduce that bit for bit.

Run in the build container only (needs ``/root/reference``):

    python tests/golden/make_golden_cut_c64.py
"""

from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_loader  # noqa: E402

CASES = [("default", {}), ("timefrac", dict(time_frac=0.3)), ("freqfrac", dict(freq_frac=0.85)),
         ("nomean", dict(remove_mean=False)), ("boost", dict(weight_boost=2.5, time_frac=0.1))]


def make_inputs():
    """Two raw baselines: a short one (10 samples: sequential / single-block sums) and a long one (150 samples: the
    recursive branch of numpy's pairwise summation)."""
    out = {}
    for tag, (T, F), seed in (("short", (10, 24), 31), ("long", (150, 20), 32)):
        r = np.random.default_rng(seed)
        vis = ((r.standard_normal((T, F)) + 1j * r.standard_normal((T, F))) * 37.0 + (5.0 - 3.0j)).astype(np.complex64)
        wgt = r.uniform(0.5, 2.0, (T, F)).astype(np.float32)
        wgt[:, [3, 4, 17]] = 0.0
        wgt[2, :] = 0.0
        wgt[5, : F - 4] = 0.0
        vis[wgt == 0] = 0.0
        out[tag] = (vis, wgt)
    return out


def main():
    ref = ref_loader.load()
    out = {}
    base = dict(time_frac=0.0, freq_frac=0.0, remove_mean=True, scale_freq=False, use_average_weights=True,
                weight_boost=1.0)
    for tag, (vis, wgt) in make_inputs().items():
        out[f"{tag}_vis"], out[f"{tag}_wgt"] = vis, wgt
        for name, kw in CASES:
            res = ref.cut_data({**base, **kw}, vis.copy(), wgt.copy())
            for nm, a in zip(("data", "weight", "nzf", "nzt"), res):
                out[f"{tag}_{name}_{nm}"] = a
        assert out[f"{tag}_default_data"].dtype == np.complex64 and out[f"{tag}_default_weight"].dtype == np.float32
    path = os.path.join(HERE, "reference_cut_c64.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
