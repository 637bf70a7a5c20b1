"""Host side of the task (no GPU): parameters, delay grid, cuts, axis flattening, sharding and
the two collectives under a world_size-2 gloo group."""
import os
import sys

import numpy as np
import pytest

from draco_b200 import _compat, delay
from draco_b200 import dist as ddist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# name -> default, as declared by the reference (delay.py:403-429, 693-694, 761-763, 1235-1236)
REFERENCE_DEFAULTS = dict(
    freq_zero=None, freq_spacing=None, nfreq=None, skip_nyquist=True, apply_window=True, window="nuttall",
    complex_timedomain=False, use_average_weights=True, weight_boost=1.0, freq_frac=0.0, time_frac=0.0,
    remove_mean=True, scale_freq=False, dataset=None, nsamp=1, save_samples=False, save_spectrum_mask=False,
    initial_amplitude=10.0, median_frac=0.5,
)


def test_parameter_names_and_defaults():
    t = delay.DelayPowerSpectrumGibbs()
    for k, v in REFERENCE_DEFAULTS.items():
        assert getattr(t, k) == v, k
    assert t.sample_axis is None  # no default in the reference either (delay.py:694)
    t = delay.DelayPowerSpectrumGibbs.from_config(dict(nsamp=20, sample_axis="ra", window="blackman_harris", seed=5))
    assert (t.nsamp, t.sample_axis, t.window, t.seed) == (20, "ra", "blackman_harris", 5)
    with pytest.raises(ValueError):
        delay.DelayPowerSpectrumGibbs.from_config(dict(window="boxcar"))
    with pytest.raises(ValueError):
        delay.DelayPowerSpectrumGibbs.from_config(dict(no_such_parameter=1))


def test_nrml_task_parameters():
    t = delay.DelayPowerSpectrumNRML()
    assert t.maxpost_tol == 1e-3 and t._get_prior(3, 8, float) == [None, None, None]  # delay.py:1279-1283
    for k, v in REFERENCE_DEFAULTS.items():
        if hasattr(t, k):
            assert getattr(t, k) == v, k


def test_legacy_names():
    assert delay.DelayPowerSpectrumStochasticEstimator is delay.DelayPowerSpectrumGibbs
    assert delay.DelaySpectrumEstimator is delay.DelayPowerSpectrumGibbs
    t = delay.DelayPowerSpectrumGibbs.from_config(dict(average_axis="ra"))
    assert t._sample_axis() == "ra"


def test_calculate_delays_matches_reference(golden):
    freq = np.linspace(800.0, 400.0, 64, endpoint=False)
    ss = _compat.SiderealStream(freq=freq, stack=1, ra=1)
    for tag, kw in [("default", {}), ("nonyq", dict(skip_nyquist=False)), ("cplx", dict(complex_timedomain=True)),
                    ("nfreq", dict(nfreq=80)), ("zero", dict(freq_zero=806.25, freq_spacing=3.125))]:
        d, ci = delay.DelayPowerSpectrumGibbs.from_config(kw)._calculate_delays(ss)
        np.testing.assert_allclose(d, golden[f"delays_{tag}"], rtol=1e-13, atol=1e-13)
        assert np.array_equal(ci, golden[f"chanind_{tag}"])


def test_cut_data_matches_reference(golden):
    vis, wgt = golden["cut_vis"], golden["cut_wgt"]
    for tag, kw in [("default", {}), ("timefrac", dict(time_frac=0.3)), ("freqfrac", dict(freq_frac=0.85)),
                    ("nomean", dict(remove_mean=False)), ("scale", dict(scale_freq=True, weight_boost=2.5))]:
        res = delay.DelayPowerSpectrumGibbs.from_config(kw)._cut_data(vis.copy(), wgt.copy())
        for nm, a in zip(("data", "weight", "nzf", "nzt"), res):
            np.testing.assert_allclose(a, golden[f"cut_{tag}_{nm}"], rtol=1e-13, atol=1e-13)
    t = delay.DelayPowerSpectrumGibbs()
    assert t._cut_data(vis, np.zeros_like(wgt)) is None
    assert delay.DelayPowerSpectrumGibbs.from_config(dict(remove_mean=False))._cut_data(np.zeros_like(vis), wgt) is None


def test_flatten_axes_and_output_layout():
    freq = np.linspace(800.0, 400.0, 8, endpoint=False)
    ss = _compat.SiderealStream(freq=freq, stack=3, ra=5)
    r = np.random.default_rng(0)
    ss.vis[:] = r.standard_normal((8, 3, 5)) + 1j * r.standard_normal((8, 3, 5))
    ss.weight[:] = r.uniform(1, 2, (8, 3, 5))
    t = delay.DelayPowerSpectrumGibbs.from_config(dict(sample_axis="ra", nsamp=4, save_samples=True, save_spectrum_mask=True))
    dv, wv, axes = t._prepare_inputs(ss)
    assert list(axes) == ["stack"] and dv.local_array.shape == (3, 5, 8)
    np.testing.assert_array_equal(dv.local_array[1], ss.vis.local_array[:, 1, :].T)
    np.testing.assert_array_equal(wv.local_array[2], ss.weight.local_array[:, 2, :].T)
    delays, _ = t._calculate_delays(ss)
    out = t._create_output(ss, delays, axes)
    assert out.spectrum.global_shape == (3, 16)
    assert out.datasets["spectrum_samples"].global_shape == (4, 3, 16)
    assert out.datasets["spectrum_mask"].global_shape == (3,)
    assert list(out.attrs["baseline_axes"]) == ["stack"] and "stack" in out.index_map
    np.testing.assert_array_equal(out.attrs["freq"], freq)
    # sample over stack instead: baselines are the ra axis
    t2 = delay.DelayPowerSpectrumGibbs.from_config(dict(sample_axis="stack"))
    dv2, _, axes2 = t2._prepare_inputs(ss)
    assert list(axes2) == ["ra"] and dv2.local_array.shape == (5, 3, 8)
    with pytest.raises(ValueError):
        delay.DelayPowerSpectrumGibbs.from_config(dict(sample_axis="el"))._prepare_inputs(ss)


def test_weights_missing_axes_are_broadcast():
    data = _compat.Dataset(np.zeros((4, 3, 5)), ["freq", "stack", "ra"])
    wgt = _compat.Dataset(np.arange(20.0).reshape(4, 5), ["freq", "ra"])
    wv, _ = delay.flatten_axes(wgt, ["ra", "freq"], match_dset=data)
    assert wv.local_array.shape == (3, 5, 4)
    np.testing.assert_array_equal(wv.local_array[2], np.arange(20.0).reshape(4, 5).T)


def test_split_range_is_caput_block_layout():
    for n, size in [(10, 3), (7, 7), (5, 8), (4096, 8), (1, 2)]:
        blocks = [ddist.split_range(n, size, r) for r in range(size)]
        assert blocks[0][0] == 0 and blocks[-1][1] == n
        assert all(blocks[i][1] == blocks[i + 1][0] for i in range(size - 1))
        lens = [b - a for a, b in blocks]
        assert max(lens) - min(lens) <= 1 and lens == sorted(lens, reverse=True)


def _gloo_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nb, N = 11, 6
        full = np.arange(nb * N, dtype=np.float64).reshape(nb, N)
        arr = _compat.DistArray.from_global(full)
        a, b = ddist.split_range(nb, world, rank)
        assert arr.local_offset == a and arr.local_array.shape[0] == b - a
        assert [g for _, g in arr.enumerate(0)] == list(range(a, b))
        local = torch.from_numpy(arr.local_array.copy())
        gathered = ddist.gather_spectra(local, nb)
        assert np.array_equal(gathered.numpy(), full)
        gids = np.arange(a, b)
        w = torch.ones(b - a, dtype=torch.float64)
        if rank == 0:
            w[0] = 0.0  # one failed chain is left out of its group
        avg, cnt = ddist.group_average(local, gids, ngroup=3, group_size=4, weights=w)
        ref = np.stack([full[1:4].mean(0), full[4:8].mean(0), full[8:11].mean(0)])
        assert np.allclose(avg.numpy(), ref) and cnt.tolist() == [3.0, 4.0, 3.0]
        assert ddist.allreduce_count(rank + 1) == sum(range(1, world + 1))
        assert ddist.broadcast_int(100 + rank) == 100
        # RandomTask stand-in: ranks get different, reproducible streams
        t = delay.DelayPowerSpectrumGibbs.from_config(dict(seed=3))
        q.put((rank, float(t.rng.standard_normal())))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharding_and_collectives():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    vals = dict(q.get() for _ in range(2))
    assert vals[0] != vals[1]
