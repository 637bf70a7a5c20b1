"""DelayCrossPowerSpectrumEstimator: host wiring (CPU, device call replaced by the oracle) and, on the GPU,
the whole task against the reference's per-baseline loop (delay.py:1395-1447) restated with the oracle."""
import numpy as np
import pytest

import oracle
from draco_b200 import _compat, delay, gibbs


def make_streams(nd, nfreq=8, nstack=3, nra=6, seed=0):
    freq = np.linspace(800.0, 400.0, nfreq, endpoint=False)
    r = np.random.default_rng(seed)
    common = r.standard_normal((nfreq, nstack, nra)) + 1j * r.standard_normal((nfreq, nstack, nra))
    out = []
    for d in range(nd):
        ss = _compat.SiderealStream(freq=freq, stack=nstack, ra=nra)
        ss.vis[:] = common * (1 + 0.2 * d) + 0.3 * (r.standard_normal(common.shape) + 1j * r.standard_normal(common.shape))
        ss.weight[:] = r.uniform(1, 2, common.shape)
        out.append(ss)
    return out


def reference_loop(task, sslist):
    """delay.py:1395-1447 with the oracle sampler and the two wrapper defects fixed (see the class docstring)."""
    delays, channel_ind = task._calculate_delays(sslist)
    dvs, wvs, axes = task._prepare_inputs(sslist)
    nd, N = len(sslist), len(delays)
    nbase = dvs[0].local_array.shape[0]
    spec_out = np.zeros((nd, nd, nbase, N))
    samp_out = np.zeros((task.nsamp, nd, nd, nbase, N))
    rng = np.random.default_rng(task.seed)
    for bi in range(nbase):
        data = np.array([d.local_array[bi] for d in dvs])
        weight = np.array([w.local_array[bi] for w in wvs])
        t = task._cut_data(data, weight)
        if t is None:
            continue
        data, weight, nzf, _ = t
        S0 = np.identity(nd)[..., None] * np.full(N, task.initial_amplitude)
        spec = oracle.cross_gibbs_chain(data.copy(), N, weight, S0.astype(complex), task.window if task.apply_window else None,
                                        channel_ind[nzf], task.nsamp, rng)
        spec_av = np.median(spec[-(task.nsamp // 2):], axis=0)
        spec_out[..., bi, :] = np.fft.fftshift(spec_av).real
        samp_out[..., bi, :] = np.array(spec).real
    return spec_out, samp_out


def test_host_wiring_with_oracle_backend(monkeypatch):
    def fake_cross_batch(data, N, Ni, S0, window="nuttall", fsel=None, niter=20, *, draws=None, **kw):
        B = data.shape[0]
        out = np.empty((B, niter, data.shape[1], data.shape[1], N), dtype=complex)
        for b in range(B):
            inj = oracle.CrossInjectedDraws(*(d[b] for d in draws))
            out[b] = np.array(oracle.cross_gibbs_chain(data[b].copy(), N, Ni[b], S0[b].copy(), window, fsel, niter, inj))
        return out, np.full(B, niter, dtype=np.int32), np.ones(B, dtype=bool)

    monkeypatch.setattr(gibbs, "cross_batch", fake_cross_batch)
    monkeypatch.setattr(gibbs, "get_handle", lambda dev=0: None)
    sslist = make_streams(2)
    cfg = dict(sample_axis="ra", nsamp=4, save_samples=True, seed=3, rng_mode="injected", device=0)
    task = delay.DelayCrossPowerSpectrumEstimator.from_config(cfg)
    out = task.process(sslist)
    assert out.spectrum.global_shape == (2, 2, 3, 16)
    assert out.datasets["spectrum_samples"].global_shape == (4, 2, 2, 3, 16)
    assert list(out.attrs["baseline_axes"]) == ["stack"] and out.attrs["window_los"] == "nuttall"
    spec, samp = reference_loop(delay.DelayCrossPowerSpectrumEstimator.from_config(cfg), sslist)
    np.testing.assert_allclose(out.spectrum.local_array, spec, rtol=1e-12)
    np.testing.assert_allclose(out.datasets["spectrum_samples"].local_array, samp, rtol=1e-12)


def test_input_validation():
    task = delay.DelayCrossPowerSpectrumEstimator.from_config(dict(sample_axis="ra", nsamp=2))
    with pytest.raises(ValueError, match="No datasets"):
        task._prepare_inputs([])
    a, b = make_streams(2)
    b.index_map["freq"] = b.index_map["freq"] + 1.0
    with pytest.raises(ValueError, match="same frequencies"):
        task._prepare_inputs([a, b])


@pytest.mark.gpu
def test_task_on_gpu_matches_reference_loop():
    gibbs.get_handle(0)
    sslist = make_streams(2, nfreq=16, nstack=4, nra=8, seed=4)
    cfg = dict(sample_axis="ra", nsamp=4, save_samples=True, seed=9, rng_mode="injected", device=0, window="blackman_harris")
    out = delay.DelayCrossPowerSpectrumEstimator.from_config(cfg).process(sslist)
    spec, samp = reference_loop(delay.DelayCrossPowerSpectrumEstimator.from_config(cfg), sslist)
    scale = np.abs(samp).max(axis=(1, 2), keepdims=True)
    assert np.max(np.abs(out.datasets["spectrum_samples"].local_array - samp) / scale) < 1e-6
    assert np.max(np.abs(out.spectrum.local_array - spec)) < 1e-6 * np.abs(spec).max()
    # native RNG runs too and gives finite spectra
    out2 = delay.DelayCrossPowerSpectrumEstimator.from_config(dict(sample_axis="ra", nsamp=6, seed=1, device=0)).process(sslist)
    assert np.isfinite(out2.spectrum.local_array).all() and np.abs(out2.spectrum.local_array).max() > 0
