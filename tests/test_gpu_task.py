"""The task classes end to end on the GPU against the reference's loop (delay.py:1117-1185)
restated with the oracle: same container in, same DelaySpectrum layout out."""
import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


def _stream(nfreq=32, nstack=6, nra=8, seed=0, dead=(4,)):
    from draco_b200 import _compat, synth

    cfg = synth.SynthConfig("t", nstack, nfreq, nra, 6, "nuttall", seed, mask="none")
    freq, vis, wgt = synth.make_stack(cfg)  # [stack, ra, freq]
    bad = [3, 7, 8, 20]
    vis[:, :, bad] = 0.0
    wgt[:, :, bad] = 0.0
    for b in dead:  # a fully masked baseline must be skipped and flagged
        vis[b] = 0.0
        wgt[b] = 0.0
    wgt[1, 2, :] = 0.0  # one dead time sample on baseline 1
    vis[1, 2, :] = 0.0
    ss = _compat.SiderealStream(freq=freq, stack=nstack, ra=nra)
    ss.vis[:] = np.transpose(vis, (2, 0, 1))
    ss.weight[:] = np.transpose(wgt, (2, 0, 1))
    return ss, vis, wgt


def _reference_loop(vis, wgt, freq, nsamp, rng, median_frac=0.5, time_frac=0.0):
    """DelayPowerSpectrumBase._evaluate + DelayPowerSpectrumGibbs._estimator with the oracle."""
    delays, chan = oracle.calculate_delays(freq)
    N = len(delays)
    nb = vis.shape[0]
    spectrum = np.zeros((nb, N))
    samples = np.zeros((nsamp, nb, N))
    mask = np.zeros(nb, dtype=bool)
    for b in range(nb):
        t = oracle.cut_data(vis[b], wgt[b], time_frac=time_frac)
        if t is None:
            mask[b] = True
            continue
        d, w, nzf, _ = t
        spec, samp, ok = oracle.gibbs_estimate(d, w, np.full(N, 10.0), N, chan[nzf], nsamp, rng, median_frac=median_frac)
        spectrum[b] = spec
        mask[b] |= not ok
        if samp:
            samples[-len(samp):, b] = np.array(samp)
    return delays, spectrum, samples, mask


def test_gibbs_task_injected_matches_reference_loop():
    from draco_b200 import delay

    ss, vis, wgt = _stream()
    task = delay.DelayPowerSpectrumGibbs.from_config(dict(
        sample_axis="ra", nsamp=6, save_samples=True, save_spectrum_mask=True, rng_mode="injected", seed=11, time_frac=0.01))
    out = task.process(ss)
    delays, spectrum, samples, mask = _reference_loop(vis, wgt, ss.freq, 6, np.random.default_rng(np.random.SeedSequence(11)),
                                                      time_frac=0.01)
    np.testing.assert_allclose(out.index_map["delay"], delays, rtol=1e-13)
    np.testing.assert_allclose(out.spectrum[:].local_array, spectrum, rtol=1e-6)
    np.testing.assert_allclose(out.datasets["spectrum_samples"][:].local_array, samples, rtol=1e-6)
    assert np.array_equal(out.datasets["spectrum_mask"][:].local_array, mask) and mask[4] and mask.sum() == 1
    assert out.attrs["window_los"] == "nuttall" and list(out.attrs["baseline_axes"]) == ["stack"]
    assert out.spectrum.global_shape == (6, 64)


def test_gibbs_task_philox_reproducible_and_batched():
    from draco_b200 import delay

    ss, _, _ = _stream(dead=())
    cfg = dict(sample_axis="ra", nsamp=8, seed=5, apply_window=False)
    a = delay.DelayPowerSpectrumGibbs.from_config(cfg).process(ss).spectrum[:].local_array.copy()
    b = delay.DelayPowerSpectrumGibbs.from_config({**cfg, "batch_baselines": 4}).process(ss).spectrum[:].local_array
    c = delay.DelayPowerSpectrumStochasticEstimator.from_config({**cfg, "sample_axis": None, "average_axis": "ra"}).process(ss)
    assert np.array_equal(a, b)  # Philox keyed on the global baseline id: batching changes nothing
    assert np.array_equal(a, c.spectrum[:].local_array)
    assert c.attrs["window_los"] == "None"
    assert np.all(a > 0)


def test_single_baseline_estimator_api():
    from draco_b200 import delay

    ss, vis, wgt = _stream(dead=())
    task = delay.DelayPowerSpectrumGibbs.from_config(dict(sample_axis="ra", nsamp=5, rng_mode="injected", seed=2))
    delays, chan = task._calculate_delays(ss)
    d, w, nzf, _ = task._cut_data(vis[0], wgt[0])
    spec, samples, ok = task._estimator(d, w, np.full(len(delays), 10.0), len(delays), chan[nzf])
    rspec, rsamp, rok = oracle.gibbs_estimate(d, w, np.full(len(delays), 10.0), len(delays), chan[nzf], 5,
                                              np.random.default_rng(np.random.SeedSequence(2)))
    assert ok == rok
    np.testing.assert_allclose(spec, rspec, rtol=1e-6)
    np.testing.assert_allclose(np.array(samples), np.array(rsamp), rtol=1e-6)


def test_wiener_filter_task():
    from draco_b200 import _compat, delay

    ss, vis, wgt = _stream(dead=(4,))
    gib = delay.DelayPowerSpectrumGibbs.from_config(dict(sample_axis="ra", nsamp=6, seed=3))
    dps = gib.process(ss)
    task = delay.DelaySpectrumWienerFilter.from_config(dict(sample_axis="ra", save_spectrum_mask=True, time_frac=0.01))
    task.setup(dps)
    out = task.process(ss)
    assert out.spectrum.global_shape == (6, 8, 64)
    delays, chan = oracle.calculate_delays(ss.freq)
    N = len(delays)
    prior = dps.spectrum[:].local_array
    for b in range(6):
        t = oracle.cut_data(vis[b], wgt[b], time_frac=0.01)
        if t is None:
            assert np.all(out.spectrum[:].local_array[b] == 0) and out.datasets["spectrum_mask"][:].local_array[b].all()
            continue
        d, w, nzf, nzt = t
        y = oracle.wiener_filter(np.fft.fftshift(prior[b]), d, N, w, "nuttall", chan[nzf], False)
        ref = np.zeros((8, N), dtype=np.complex128)
        ref[nzt] = np.fft.fftshift(y, axes=-1)
        got = out.spectrum[:].local_array[b]
        assert np.max(np.abs(got - ref)) < 1e-8 * np.max(np.abs(ref))
        assert np.array_equal(out.datasets["spectrum_mask"][:].local_array[b], ~nzt)
    it = delay.DelaySpectrumWienerFilterIteratePS.from_config(dict(sample_axis="ra"))
    out2 = it.process(ss, dps)
    np.testing.assert_allclose(out2.spectrum[:].local_array[0], out.spectrum[:].local_array[0], rtol=1e-12)


def test_gibbs_task_device_cut_matches_host_cut():
    """``device_cut`` (dgibbs_cut_data) changes where ``_cut_data`` runs, not one bit of the result."""
    from draco_b200 import delay

    ss, _, _ = _stream()
    cfg = dict(sample_axis="ra", nsamp=8, seed=5, save_samples=True, save_spectrum_mask=True, time_frac=0.01,
               weight_boost=3.0)
    a = delay.DelayPowerSpectrumGibbs.from_config({**cfg, "device_cut": True}).process(ss)
    b = delay.DelayPowerSpectrumGibbs.from_config({**cfg, "device_cut": False}).process(ss)
    c = delay.DelayPowerSpectrumGibbs.from_config({**cfg, "device_cut": True, "batch_baselines": 4}).process(ss)
    for x in (b, c):
        assert np.array_equal(a.spectrum[:].local_array, x.spectrum[:].local_array)
        assert np.array_equal(a.datasets["spectrum_samples"][:].local_array, x.datasets["spectrum_samples"][:].local_array)
        assert np.array_equal(a.datasets["spectrum_mask"][:].local_array, x.datasets["spectrum_mask"][:].local_array)
    assert a.datasets["spectrum_mask"][:].local_array[4] and np.all(a.spectrum[:].local_array[0] > 0)


def test_tasks_on_container_dtypes_complex64_float32():
    """draco's SiderealStream stores complex64 visibilities and float32 weights: the raw slab is uploaded as it is and
    cut on the device in numpy's complex64 / float32 arithmetic, so both tasks return what the host ``_cut_data``
    route returns, bit for bit; under ``scale_freq`` the task falls back to the host cut by itself."""
    from draco_b200 import _compat, delay

    ss64, vis, wgt = _stream()
    ss = _compat.SiderealStream(freq=ss64.freq, stack=vis.shape[0], ra=vis.shape[1], native_dtypes=True)
    ss.vis[:] = np.transpose(vis, (2, 0, 1))
    ss.weight[:] = np.transpose(wgt, (2, 0, 1))
    assert ss.vis[:].dtype == np.complex64 and ss.weight[:].dtype == np.float32
    cfg = dict(sample_axis="ra", nsamp=8, seed=5, save_samples=True, save_spectrum_mask=True, time_frac=0.01,
               weight_boost=3.0)
    for extra in ({}, {"scale_freq": True}):
        a = delay.DelayPowerSpectrumGibbs.from_config({**cfg, **extra, "device_cut": True}).process(ss)
        b = delay.DelayPowerSpectrumGibbs.from_config({**cfg, **extra, "device_cut": False}).process(ss)
        assert np.array_equal(a.spectrum[:].local_array, b.spectrum[:].local_array)
        assert np.array_equal(a.datasets["spectrum_samples"][:].local_array, b.datasets["spectrum_samples"][:].local_array)
        assert np.array_equal(a.datasets["spectrum_mask"][:].local_array, b.datasets["spectrum_mask"][:].local_array)
        assert np.all(a.spectrum[:].local_array[0] > 0)
    # not the complex128 result: the float32 rounding of the inputs is visible
    c = delay.DelayPowerSpectrumGibbs.from_config({**cfg, "device_cut": True}).process(ss64)
    a = delay.DelayPowerSpectrumGibbs.from_config({**cfg, "device_cut": True}).process(ss)
    assert not np.array_equal(a.spectrum[:].local_array, c.spectrum[:].local_array)
    np.testing.assert_allclose(a.spectrum[:].local_array, c.spectrum[:].local_array, rtol=1e-2)
    dps = delay.DelayPowerSpectrumGibbs.from_config(dict(sample_axis="ra", nsamp=6, seed=3)).process(ss)
    outs = []
    for kw in (dict(device_cut=True), dict(device_cut=False)):
        task = delay.DelaySpectrumWienerFilter.from_config(dict(sample_axis="ra", time_frac=0.01, **kw))
        task.setup(dps)
        outs.append(task.process(ss))
    assert np.array_equal(outs[0].spectrum[:].local_array, outs[1].spectrum[:].local_array)


def test_wiener_task_device_cut_matches_host_cut():
    from draco_b200 import delay

    ss, _, _ = _stream(dead=(4,))
    dps = delay.DelayPowerSpectrumGibbs.from_config(dict(sample_axis="ra", nsamp=6, seed=3)).process(ss)
    outs = []
    for kw in (dict(device_cut=True), dict(device_cut=False), dict(device_cut=True, batch_baselines=4)):
        task = delay.DelaySpectrumWienerFilter.from_config(dict(sample_axis="ra", save_spectrum_mask=True, time_frac=0.01, **kw))
        task.setup(dps)
        outs.append(task.process(ss))
    for x in outs[1:]:
        assert np.array_equal(outs[0].spectrum[:].local_array, x.spectrum[:].local_array)
        assert np.array_equal(outs[0].datasets["spectrum_mask"][:].local_array, x.datasets["spectrum_mask"][:].local_array)
    assert np.abs(outs[0].spectrum[:].local_array[0]).max() > 0
