"""Generate the golden vectors in this directory by running the REFERENCE's own functions.

Run in the build container only (needs ``/root/reference``):

    python tests/golden/make_golden.py

The reference has no tests or fixtures for the delay code, so these
vectors are produced by executing ``draco/analysis/delay.py:1475-2201`` and
``draco/util/tools.py:547-601`` in place through ``oracle/ref_loader.py``.  Inputs are
seeded; random draws are regenerated in the tests from the recorded seeds (numpy
``default_rng`` / PCG64 streams), so only small outputs are stored.
"""

from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_loader  # noqa: E402

# (name, N, M, T, complex_timedomain, window, niter, include_nyquist)
CHAIN_CASES = [
    ("f_real_n64", 64, 29, 8, False, "nuttall", 6, False),
    ("f_real_nowin_nyq", 64, 33, 8, False, None, 4, True),
    ("f_real_cfg1", 256, 102, 32, False, "nuttall", 20, False),
    ("f_cplx_n64", 64, 50, 8, True, "nuttall", 5, False),
    ("f_real_n100_hann", 100, 40, 5, False, "hann", 4, False),
    ("t_real_n64", 64, 12, 8, False, "nuttall", 6, False),
    ("t_cplx_n64", 64, 14, 6, True, "blackman", 4, False),
    ("t_real_n256", 256, 40, 16, False, "blackman_harris", 5, False),
]


def make_case(N, M, T, cplx, nyq, seed):
    """Seeded single-baseline inputs (shared with tests/_cases.py)."""
    r = np.random.default_rng(seed)
    nch = N if cplx else N // 2 + 1
    fsel = np.sort(r.choice(nch - (0 if (nyq or cplx) else 1), size=M, replace=False))
    if nyq and not cplx:
        fsel[-1] = N // 2  # make sure the Nyquist channel is exercised
        fsel[0] = 0
    data = r.standard_normal((T, M)) + 1j * r.standard_normal((T, M))
    Ni = r.uniform(0.5, 2.0, M) * 100
    S0 = np.full(N, 10.0)
    return data, Ni, fsel, S0


class _Rec:
    def __init__(self, rng):
        self.rng = rng
        self.calls = []

    def standard_normal(self, shape):
        self.calls.append(("normal", tuple(shape)))
        return self.rng.standard_normal(shape)

    def chisquare(self, df, size=None):
        self.calls.append(("chi2", int(df), int(size)))
        return self.rng.chisquare(df, size=size)


def main():
    ref = ref_loader.load()
    out = {}

    # --- Fourier matrices and windows
    N, fsel = 32, np.array([0, 1, 3, 7, 16])
    out["fm_N"] = N
    out["fm_fsel"] = fsel
    out["fm_r2c"] = ref.fourier_matrix_r2c(N, fsel)
    out["fm_c2r"] = ref.fourier_matrix_c2r(N, fsel)
    out["fm_c2c"] = ref.fourier_matrix_c2c(N, fsel)
    out["fm_complex"] = ref.fourier_matrix(N, fsel)
    out["fm_r2c_full"] = ref.fourier_matrix_r2c(16)
    x = np.linspace(-0.1, 1.1, 49)
    out["win_x"] = x
    for w in ["uniform", "hann", "hanning", "hamming", "blackman", "nuttall", "blackman_nuttall", "blackman_harris",
              "triangular", "tukey-0.5"]:
        out["win_" + w] = ref.window_generalised(x, window=w)

    # --- setup triple, Wiener filter and chains
    for name, N, M, T, cplx, window, niter, nyq in CHAIN_CASES:
        data, Ni, fsel, S0 = make_case(N, M, T, cplx, nyq, seed=11)
        z, fh, ff = ref._compute_delay_spectrum_inputs(data.copy(), N, Ni, fsel, window, cplx)
        if N <= 64:
            out[f"{name}_z"], out[f"{name}_FTNih"], out[f"{name}_FTNiF"] = z, fh, ff
        else:  # keep fixtures small: checksums + a corner
            out[f"{name}_z"] = z
            out[f"{name}_FTNih_sum"] = np.array([fh.sum(), np.abs(fh).sum(), (fh**2).sum()])
            out[f"{name}_FTNiF_col0"] = ff[:, 0].copy()
            out[f"{name}_FTNiF_diag"] = np.diag(ff).copy()
        out[f"{name}_wiener"] = ref.delay_spectrum_wiener_filter(S0, data.copy(), N, Ni, window, fsel, cplx)
        rec = _Rec(np.random.default_rng(7))
        spec, ok = ref.delay_power_spectrum_gibbs(data.copy(), N, Ni, S0, window=window, fsel=fsel, niter=niter,
                                                  rng=rec, complex_timedomain=cplx)
        out[f"{name}_samples"] = np.array(spec)
        out[f"{name}_success"] = np.array(ok)
        if name in ("f_real_n64", "t_real_n64", "f_cplx_n64"):
            out[f"{name}_draw_order"] = np.array([repr(c) for c in rec.calls[:3]])

    # --- failure semantics: a negative chi2 at iteration 2 makes S negative, the next
    # factorisation fails and the chain is truncated (delay.py:1891-1896)
    data, Ni, fsel, S0 = make_case(64, 29, 8, False, False, seed=11)

    class _Bad(_Rec):
        n = 0

        def chisquare(self, df, size=None):
            c = self.rng.chisquare(df, size=size)
            if self.n == 2:
                c = -c
            self.n += 1
            return c

    spec, ok = ref.delay_power_spectrum_gibbs(data.copy(), 64, Ni, S0, window="nuttall", fsel=fsel, niter=6,
                                              rng=_Bad(np.random.default_rng(7)))
    out["fail_samples"] = np.array(spec)
    out["fail_success"] = np.array(ok)
    spec, ok = ref.delay_power_spectrum_gibbs(data.copy(), 64, Ni, -S0, window="nuttall", fsel=fsel, niter=3,
                                              rng=np.random.default_rng(7))
    out["fail0_len"] = np.array(len(spec))
    out["fail0_success"] = np.array(ok)

    # --- host-side task methods
    freq = np.linspace(800.0, 400.0, 64, endpoint=False)
    base = dict(freq_zero=None, freq_spacing=None, nfreq=None, skip_nyquist=True, complex_timedomain=False)
    for tag, kw in [("default", {}), ("nonyq", dict(skip_nyquist=False)), ("cplx", dict(complex_timedomain=True)),
                    ("nfreq", dict(nfreq=80)), ("zero", dict(freq_zero=806.25, freq_spacing=3.125))]:
        d, ci = ref.calculate_delays({**base, **kw}, freq)
        out[f"delays_{tag}"], out[f"chanind_{tag}"] = d, ci
    r = np.random.default_rng(21)
    vis = r.standard_normal((10, 24)) + 1j * r.standard_normal((10, 24))
    wgt = r.uniform(0.5, 2, (10, 24))
    wgt[:, [3, 4, 17]] = 0.0
    wgt[2, :] = 0.0
    wgt[5, :20] = 0.0
    vis[wgt == 0] = 0.0
    out["cut_vis"], out["cut_wgt"] = vis, wgt
    cutbase = dict(time_frac=0.0, freq_frac=0.0, remove_mean=True, scale_freq=False, use_average_weights=True,
                   weight_boost=1.0)
    for tag, kw in [("default", {}), ("timefrac", dict(time_frac=0.3)), ("freqfrac", dict(freq_frac=0.85)),
                    ("nomean", dict(remove_mean=False)), ("scale", dict(scale_freq=True, weight_boost=2.5))]:
        res = ref.cut_data({**cutbase, **kw}, vis.copy(), wgt.copy())
        for nm, a in zip(("data", "weight", "nzf", "nzt"), res):
            out[f"cut_{tag}_{nm}"] = a
    out["cut_dead_is_none"] = np.array(ref.cut_data(cutbase, vis.copy(), np.zeros_like(wgt)) is None)
    out["cut_zero_is_none"] = np.array(ref.cut_data({**cutbase, "remove_mean": False}, np.zeros_like(vis), wgt.copy()) is None)

    path = os.path.join(HERE, "reference_delay.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
