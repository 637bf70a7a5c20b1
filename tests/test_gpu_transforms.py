"""Device path of the plain FFT delay transform (``dgibbs_delay_fft``; delay.py:960-980, 2102-2130) and of
DelaySpectrumToPowerSpectrum (``dgibbs_sample_variance``; delay.py:1061-1106) through the C ABI, against vectors
produced by the reference's own code and against the numpy oracle."""
import os
import sys

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_golden_transforms import FFT_CASES, VAR_CASES, fft_input, var_input  # noqa: E402

FFT_ATOL = 1e-13  # of the largest output; fp64 FFT rounding (numpy / FFTW / this kernel differ at that level)


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "reference_transforms.npz"))


def test_fft_matches_reference_golden(golden):
    from draco_b200 import gibbs

    for i, (name, ns, N, window) in enumerate(FFT_CASES):
        x = fft_input(ns, N, 900 + i)
        keep = x.copy()
        got = gibbs.delay_spectrum_fft(x, N, window)
        assert np.array_equal(x, keep)
        ref = golden[name]
        np.testing.assert_allclose(got, ref, rtol=0, atol=FFT_ATOL * np.abs(ref).max(), err_msg=name)
        sh = gibbs.fft_rows(x, window, shift=True)
        np.testing.assert_allclose(sh, np.fft.fftshift(ref, axes=-1), rtol=0, atol=FFT_ATOL * np.abs(ref).max())
    with pytest.raises(ValueError):
        gibbs.delay_spectrum_fft(fft_input(2, 60, 1), 64, "nuttall")  # the reference's broadcast error


@pytest.mark.parametrize("N,window", [(1024, "nuttall"), (4096, "blackman"), (16, None), (1000, "nuttall"), (3, "hann"),
                                      (6400, None)])
def test_fft_matches_oracle_sizes(N, window):
    """Both transform kernels (in-smem FFT for N = 2^m <= 4096, direct DFT otherwise), host and device pointers."""
    import torch

    from draco_b200 import gibbs

    nrow = 300 if N <= 1024 else 7
    x = fft_input(nrow, N, N)
    ref = oracle.delay_spectrum_fft(x, N, window)
    t = {}
    got = gibbs.fft_rows(x, window, timing=t)
    scale = np.abs(ref).max(axis=-1, keepdims=True)
    assert np.max(np.abs(got - ref) / scale) < FFT_ATOL * max(1.0, N / 1024)
    assert t["ms_kernel"] > 0
    d = torch.from_numpy(x.view(np.float64).reshape(nrow, N, 2)).cuda()
    dev = gibbs.fft_rows(d, window, shift=True).cpu().numpy().reshape(nrow, 2 * N).view(np.complex128)
    assert np.array_equal(dev, np.fft.fftshift(got, axes=-1))
    assert gibbs.fft_rows(np.zeros((0, N), dtype=np.complex128), window).shape == (0, N)


def test_fft_rejects_unserved_sizes():
    from draco_b200 import gibbs

    with pytest.raises(gibbs.DGibbsError):
        gibbs.fft_rows(np.zeros((1, 8192), dtype=np.complex128), None)


def test_sample_variance_matches_reference_golden(golden):
    from draco_b200 import gibbs

    for i, (name, B, T, N, masked) in enumerate(VAR_CASES):
        spec, mask = var_input(B, T, N, masked, 950 + i)
        var, bad = gibbs.sample_variance(spec, mask)
        ref = golden[f"{name}_spectrum"]
        np.testing.assert_allclose(var, ref, rtol=1e-13, atol=0, err_msg=name)
        assert np.array_equal(bad, golden[f"{name}_mask"])
        assert np.array_equal(var == 0.0, ref == 0.0)
    spec, _ = var_input(5, 33, 700, False, 3)  # no mask: plain variance, NaN stays NaN
    spec[2, 4, 10] = np.nan
    var, bad = gibbs.sample_variance(spec)
    ref = np.var(spec, axis=1)
    assert np.isnan(var[2, 10]) and not bad.any()
    ok = ~np.isnan(ref)
    np.testing.assert_allclose(var[ok], ref[ok], rtol=1e-13)


def _complex_stream(nfreq, nstack, nra, seed):
    from draco_b200 import _compat

    r = np.random.default_rng(seed)
    freq = np.linspace(800.0, 400.0, nfreq, endpoint=False)
    ss = _compat.SiderealStream(freq=freq, stack=nstack, ra=nra)
    vis = r.standard_normal((nfreq, nstack, nra)) + 1j * r.standard_normal((nfreq, nstack, nra))
    wgt = r.uniform(0.5, 2.0, (nfreq, nstack, nra))
    wgt[:, 2, :] = 0.0  # a dead baseline
    wgt[:, 1, 3] = 0.0  # a dead sample on baseline 1
    vis[wgt == 0] = 0.0
    ss.vis[:] = vis
    ss.weight[:] = wgt
    return ss, vis, wgt


@pytest.mark.parametrize("nfreq", [32, 24])
def test_fft_task_then_power_spectrum_task(nfreq):
    """DelaySpectrumFFT -> DelaySpectrumToPowerSpectrum against the reference's loops restated with the oracle."""
    from draco_b200 import delay

    nstack, nra = 5, 9
    ss, vis, wgt = _complex_stream(nfreq, nstack, nra, 11)
    cfg = dict(sample_axis="ra", complex_timedomain=True, save_spectrum_mask=True, window="blackman_harris")
    outs = [delay.DelaySpectrumFFT.from_config({**cfg, **kw}).process(ss)
            for kw in (dict(device_cut=True), dict(device_cut=False), dict(device_cut=True, batch_baselines=2))]
    delays, chan = oracle.calculate_delays(ss.freq, complex_timedomain=True)
    N = len(delays)
    assert N == nfreq
    ref = np.zeros((nstack, nra, N), dtype=np.complex128)
    rmask = np.zeros((nstack, nra), dtype=bool)
    for b in range(nstack):  # delay.py:899-930
        t = oracle.cut_data(vis[:, b, :].T.copy(), wgt[:, b, :].T.copy())
        if t is None:
            rmask[b] = True
            continue
        d, w, nzf, nzt = t
        ref[b, nzt] = np.fft.fftshift(oracle.delay_spectrum_fft(d, N, "blackman_harris"), axes=-1)
        rmask[b][~nzt] = True
    for o in outs:
        np.testing.assert_allclose(o.spectrum[:].local_array, ref, rtol=0, atol=1e-13 * np.abs(ref).max())
        assert np.array_equal(o.datasets["spectrum_mask"][:].local_array, rmask)
    assert np.array_equal(outs[0].spectrum[:].local_array, outs[1].spectrum[:].local_array)
    ps = delay.DelaySpectrumToPowerSpectrum.from_config({}).process(outs[0])
    rps, rflag = oracle.delay_spectrum_to_power_spectrum(ref, rmask)
    np.testing.assert_allclose(ps.spectrum[:].local_array, rps, rtol=1e-12, atol=1e-13 * np.abs(rps).max())
    assert np.array_equal(ps.datasets["spectrum_mask"][:].local_array, rflag) and rflag[2] and not rflag[0]
    # real time-domain mode: N = 2 nfreq channels are expected, the window cannot broadcast (as in the reference)
    with pytest.raises(ValueError):
        delay.DelaySpectrumFFT.from_config(dict(sample_axis="ra")).process(ss)


def test_deprecated_estimators_raise():
    from draco_b200 import delay

    for cls in (delay.DelayPowerSpectrumStokesIEstimator, delay.DelayPowerSpectrumGeneralEstimator):
        with pytest.raises(DeprecationWarning):
            cls.from_config(dict(sample_axis="ra")).setup()
