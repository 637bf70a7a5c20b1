"""``draco_b200.install()``: an unmodified pipeline config (``type: draco.analysis.delay.DelayPowerSpectrumGibbs``)
must resolve to the GPU evaluation; only the plug-in surface of SURVEY 8b is rebound and ``uninstall`` restores it."""
import importlib

import numpy as np
import pytest

import _fake_draco


def test_install_rebinds_only_the_plugin_surface(monkeypatch):
    import draco_b200
    from draco_b200 import delay as ours, hooks

    fake = _fake_draco.inject(monkeypatch)
    ref_cls = importlib.import_module("draco.analysis.delay").DelayPowerSpectrumGibbs  # what the YAML `type:` resolves to
    before = dict(vars(ref_cls))
    bound = draco_b200.install()
    try:
        assert "DelayPowerSpectrumGibbs._evaluate" in bound and "DelaySpectrumWienerFilter._evaluate" in bound
        assert ref_cls._evaluate is ours.DelayPowerSpectrumGibbs._evaluate
        assert ref_cls._estimator is ours.DelayPowerSpectrumGibbs._estimator
        assert ref_cls._device is ours.DelayTransformBase._device
        # the reference keeps what SURVEY 8b says it keeps
        for keep in ("process",):
            assert vars(ref_cls)[keep] is before[keep]
        for keep in ("_calculate_delays", "_cut_data", "_prepare_inputs", "_create_output"):
            assert keep not in vars(ref_cls) or vars(ref_cls)[keep] is before.get(keep)
        # optional parameters exist with defaults that change nothing; YAML parameters keep working
        t = ref_cls.from_config(dict(nsamp=3, sample_axis="ra", window="hann", median_frac=0.4))
        assert t.device is None and t.batch_baselines == 0 and t.rng_mode == "philox" and t.device_cut is True
        assert t.nsamp == 3 and t.window == "hann" and t.median_frac == 0.4
        t2 = ref_cls.from_config(dict(device=1, rng_mode="injected"))
        assert t2.device == 1 and t2.rng_mode == "injected"
        # historical alias inherits the rebinding
        assert fake.DelayPowerSpectrumStochasticEstimator._evaluate is ours.DelayPowerSpectrumGibbs._evaluate
    finally:
        draco_b200.uninstall()
    assert vars(ref_cls)["_evaluate"] is before["_evaluate"] and vars(ref_cls)["_estimator"] is before["_estimator"]
    assert vars(ref_cls)["_device"] is None and not hooks._saved


def test_stand_alone_classes_refuse_a_multi_rank_mpiarray():
    """Advisor finding r01: under real caput + MPI the stand-alone classes would treat one rank's frequency slab as the
    global array.  They must refuse; install() is the supported route there."""
    from draco_b200 import _compat, delay

    class FakeComm:
        size, rank = 2, 0

    class FakeMPIArray:
        comm = FakeComm()
        local_array = np.zeros((4, 3, 5), dtype=np.complex128)

    ds = _compat.Dataset(FakeMPIArray(), ["freq", "stack", "ra"])
    with pytest.raises(RuntimeError, match="install"):
        delay.flatten_axes(ds, ["ra", "freq"])


@pytest.mark.gpu
def test_unmodified_config_runs_on_the_gpu(monkeypatch):
    import draco_b200
    from draco_b200 import delay as ours
    from test_gpu_task import _stream

    _fake_draco.inject(monkeypatch)
    ss, _, _ = _stream(dead=())
    params = dict(sample_axis="ra", nsamp=6, seed=5, save_samples=True)  # an existing config: no new keys
    ref_cls = importlib.import_module("draco.analysis.delay").DelayPowerSpectrumGibbs
    with pytest.raises(_fake_draco.CpuLoopRan):
        ref_cls.from_config(params).process(ss)
    draco_b200.install()
    try:
        task = ref_cls.from_config(params)
        out = task.process(ss)
        assert task.reference_process_calls == 1  # the reference's own process() drove it
    finally:
        draco_b200.uninstall()
    want = ours.DelayPowerSpectrumGibbs.from_config(params).process(ss)
    assert np.array_equal(out.spectrum[:].local_array, want.spectrum[:].local_array)
    assert np.array_equal(out.datasets["spectrum_samples"][:].local_array, want.datasets["spectrum_samples"][:].local_array)
    assert np.all(out.spectrum[:].local_array > 0)
