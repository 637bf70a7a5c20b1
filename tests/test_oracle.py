"""The CPU oracle against the golden vectors produced by the reference itself
(tests/golden/make_golden.py), and against the live reference when its tree is present."""
import numpy as np
import pytest

import oracle
from oracle import ref_loader
from _cases import CASES, make_case

TIGHT = dict(rtol=1e-13, atol=1e-13)


def test_fourier_matrices(golden):
    N, fsel = int(golden["fm_N"]), golden["fm_fsel"]
    np.testing.assert_allclose(oracle.dft_r2c(N, fsel), golden["fm_r2c"], **TIGHT)
    np.testing.assert_allclose(oracle.dft_c2r(N, fsel), golden["fm_c2r"], **TIGHT)
    np.testing.assert_allclose(oracle.dft_c2c(N, fsel), golden["fm_c2c"], **TIGHT)
    np.testing.assert_allclose(oracle.dft_complex(N, fsel), golden["fm_complex"], **TIGHT)
    np.testing.assert_allclose(oracle.dft_r2c(16), golden["fm_r2c_full"], **TIGHT)
    # c2r is the inverse of r2c on the full channel set
    np.testing.assert_allclose(oracle.dft_c2r(16) @ oracle.dft_r2c(16), np.eye(16), atol=1e-12)


@pytest.mark.parametrize("name", ["uniform", "hann", "hanning", "hamming", "blackman", "nuttall", "blackman_nuttall",
                                  "blackman_harris", "triangular", "tukey-0.5"])
def test_windows(golden, name):
    np.testing.assert_allclose(oracle.cosine_sum_window(golden["win_x"], name), golden["win_" + name], **TIGHT)


@pytest.mark.parametrize("name", list(CASES))
def test_setup_wiener_chain(golden, name):
    _, N, M, T, cplx, window, niter, nyq = CASES[name]
    data, Ni, fsel, S0 = make_case(N, M, T, cplx, nyq, seed=11)
    z, fh, ff = oracle.whiten_inputs(data, N, Ni, fsel, window, cplx)
    np.testing.assert_allclose(z, golden[f"{name}_z"], **TIGHT)
    if f"{name}_FTNih" in golden:
        np.testing.assert_allclose(fh, golden[f"{name}_FTNih"], **TIGHT)
        np.testing.assert_allclose(ff, golden[f"{name}_FTNiF"], rtol=1e-12, atol=1e-9)
    else:
        np.testing.assert_allclose([fh.sum(), np.abs(fh).sum(), (fh**2).sum()], golden[f"{name}_FTNih_sum"], rtol=1e-11)
        np.testing.assert_allclose(ff[:, 0], golden[f"{name}_FTNiF_col0"], rtol=1e-12, atol=1e-9)
        np.testing.assert_allclose(np.diag(ff), golden[f"{name}_FTNiF_diag"], rtol=1e-12)
    np.testing.assert_allclose(oracle.wiener_filter(S0, data, N, Ni, window, fsel, cplx), golden[f"{name}_wiener"],
                               rtol=1e-9, atol=1e-12)
    samples, ok = oracle.gibbs_chain(data, N, Ni, S0, window, fsel, niter, np.random.default_rng(7), cplx)
    assert ok == bool(golden[f"{name}_success"])
    np.testing.assert_allclose(np.array(samples), golden[f"{name}_samples"], rtol=1e-9)


def test_draw_order(golden):
    # per iteration: normal(n_sys, T) -> normal(2M, T) -> chisquare(T, N)   
    rec = oracle.RecordingDraws(np.random.default_rng(0))
    data, Ni, fsel, S0 = make_case(64, 29, 8, False, False, seed=11)
    oracle.gibbs_chain(data, 64, Ni, S0, "nuttall", fsel, 2, rec, False)
    assert [a.shape for a in rec.normals] == [(64, 8), (58, 8), (64, 8), (58, 8)]
    assert [a.shape for a in rec.chisq] == [(64,), (64,)]
    assert list(golden["f_real_n64_draw_order"]) == ["('normal', (64, 8))", "('normal', (58, 8))", "('chi2', 8, 64)"]
    assert list(golden["f_cplx_n64_draw_order"])[0] == "('normal', (128, 8))"


def test_failure_semantics(golden):
    data, Ni, fsel, S0 = make_case(64, 29, 8, False, False, seed=11)

    class Bad(oracle.RecordingDraws):
        n = 0

        def chisquare(self, df, size=None):
            c = super().chisquare(df, size)
            if self.n == 2:
                c = -c
            self.n += 1
            return c

    samples, ok = oracle.gibbs_chain(data, 64, Ni, S0, "nuttall", fsel, 6, Bad(np.random.default_rng(7)), False)
    assert not ok and not bool(golden["fail_success"])
    np.testing.assert_allclose(np.array(samples), golden["fail_samples"], rtol=1e-9)
    samples, ok = oracle.gibbs_chain(data, 64, Ni, -S0, "nuttall", fsel, 3, np.random.default_rng(7), False)
    assert len(samples) == int(golden["fail0_len"]) == 0 and not ok
    spec, samples, ok = oracle.gibbs_estimate(data, Ni, -S0, 64, fsel, 3, np.random.default_rng(7))
    assert not ok and np.array_equal(spec, -S0)


def test_injected_draws_reproduce_chain():
    data, Ni, fsel, S0 = make_case(64, 29, 8, False, False, seed=3)
    rec = oracle.RecordingDraws(np.random.default_rng(5))
    a, _ = oracle.gibbs_chain(data, 64, Ni, S0, "nuttall", fsel, 4, rec, False)
    b, _ = oracle.gibbs_chain(data, 64, Ni, S0, "nuttall", fsel, 4, oracle.InjectedDraws(*rec.packed()), False)
    assert np.array_equal(np.array(a), np.array(b))


def test_calculate_delays(golden):
    freq = np.linspace(800.0, 400.0, 64, endpoint=False)
    for tag, kw in [("default", {}), ("nonyq", dict(skip_nyquist=False)), ("cplx", dict(complex_timedomain=True)),
                    ("nfreq", dict(nfreq=80)), ("zero", dict(freq_zero=806.25, freq_spacing=3.125))]:
        d, ci = oracle.calculate_delays(freq, **kw)
        np.testing.assert_allclose(d, golden[f"delays_{tag}"], **TIGHT)
        assert np.array_equal(ci, golden[f"chanind_{tag}"])


def test_cut_data(golden):
    vis, wgt = golden["cut_vis"], golden["cut_wgt"]
    for tag, kw in [("default", {}), ("timefrac", dict(time_frac=0.3)), ("freqfrac", dict(freq_frac=0.85)),
                    ("nomean", dict(remove_mean=False)), ("scale", dict(scale_freq=True, weight_boost=2.5))]:
        res = oracle.cut_data(vis.copy(), wgt.copy(), **kw)
        for nm, a in zip(("data", "weight", "nzf", "nzt"), res):
            np.testing.assert_allclose(a, golden[f"cut_{tag}_{nm}"], **TIGHT)
    assert oracle.cut_data(vis, np.zeros_like(wgt)) is None and bool(golden["cut_dead_is_none"])
    assert oracle.cut_data(np.zeros_like(vis), wgt, remove_mean=False) is None and bool(golden["cut_zero_is_none"])


def test_cut_data_complex64():
    """The container dtypes (complex64 visibilities, float32 weights): the oracle keeps the reference's dtypes and
    arithmetic, bit for bit (tests/golden/make_golden_cut_c64.py)."""
    import os

    from _cases import C64_CASES

    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_cut_c64.npz"))
    for tag in ("short", "long"):
        vis, wgt = g[f"{tag}_vis"], g[f"{tag}_wgt"]
        assert vis.dtype == np.complex64 and wgt.dtype == np.float32
        for name, kw in C64_CASES:
            res = oracle.cut_data(vis.copy(), wgt.copy(), **kw)
            for nm, a in zip(("data", "weight", "nzf", "nzt"), res):
                ref = g[f"{tag}_{name}_{nm}"]
                assert a.dtype == ref.dtype and np.array_equal(a, ref), (tag, name, nm)


def test_ftnif_is_circulant():
    # structure the CUDA path relies on: FTNiF[t,t'] = c[(t-t') mod N] for any mask/weights/window
    data, Ni, fsel, _ = make_case(64, 29, 8, False, False, seed=2)
    _, _, ff = oracle.whiten_inputs(data, 64, Ni, fsel, "nuttall", False)
    idx = (np.arange(64)[:, None] - np.arange(64)[None, :]) % 64
    np.testing.assert_allclose(ff, ff[:, 0][idx], atol=1e-9 * np.abs(ff).max())


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present (GPU box)")
def test_against_live_reference():
    ref = ref_loader.load()
    for name in ("f_real_n64", "t_real_n64", "f_cplx_n64", "t_cplx_n64"):
        _, N, M, T, cplx, window, niter, nyq = CASES[name]
        data, Ni, fsel, S0 = make_case(N, M, T, cplx, nyq, seed=99)
        a, oka = ref.delay_power_spectrum_gibbs(data.copy(), N, Ni, S0, window=window, fsel=fsel, niter=niter,
                                                rng=np.random.default_rng(1), complex_timedomain=cplx)
        b, okb = oracle.gibbs_chain(data.copy(), N, Ni, S0, window, fsel, niter, np.random.default_rng(1), cplx)
        assert oka == okb
        np.testing.assert_allclose(np.array(b), np.array(a), rtol=1e-12)
