"""Device-side ``_cut_data`` (``dgibbs_cut_data``) against the reference's golden vectors and
the numpy oracle (``oracle.cut_data`` = delay.py:516-597, itself pinned on the same golden vectors)."""

import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(__file__)
TAGS = [("default", {}), ("timefrac", dict(time_frac=0.3)), ("freqfrac", dict(freq_frac=0.85)),
        ("nomean", dict(remove_mean=False)), ("scale", dict(scale_freq=True, weight_boost=2.5))]


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "reference_delay.npz"))


def _unpack(cr, n, T, F):
    """Host copies of the n-th surviving baseline of a CutResult: (data[Tb, Mb], Ni[Mb], fsel[Mb])."""
    pb = cr.batch
    Tb, Mb = int(pb.T[n]), int(pb.M[n])
    do, co = int(pb.data_off[n]), int(pb.chan_off[n])
    d = pb.data.view(-1, 2)[do : do + Tb * Mb].cpu().numpy().view(np.complex128).reshape(Tb, Mb)
    return d, pb.Ni[co : co + Mb].cpu().numpy(), pb.fsel[co : co + Mb].cpu().numpy()


def test_cut_matches_reference_golden(golden):
    from draco_b200 import gibbs

    vis, wgt = golden["cut_vis"], golden["cut_wgt"]
    T, F = vis.shape
    cind = np.arange(F) * 2 + 1  # any injective channel map
    for tag, kw in TAGS:
        full = dict(time_frac=0.0, freq_frac=0.0, remove_mean=True, scale_freq=False, weight_boost=1.0)
        full.update(kw)
        cr = gibbs.cut_batch(vis[None], wgt[None], cind, 2 * F, 1.0, **full)
        assert cr.alive.tolist() == [True]
        assert np.array_equal(cr.keep_t[0], golden[f"cut_{tag}_nzt"])
        assert np.array_equal(cr.keep_f[0], golden[f"cut_{tag}_nzf"])
        d, ni, fs = _unpack(cr, 0, T, F)
        assert np.array_equal(fs, cind[golden[f"cut_{tag}_nzf"]])
        if tag == "scale":  # numpy's all-element std is a pairwise sum: equal to rounding
            np.testing.assert_allclose(d, golden[f"cut_{tag}_data"], rtol=1e-13, atol=0)
        else:
            assert np.array_equal(d, golden[f"cut_{tag}_data"])
        assert np.array_equal(ni, golden[f"cut_{tag}_weight"])
    # the three "return None" exits (delay.py:545, 564, 577)
    assert not gibbs.cut_batch(vis[None], np.zeros_like(wgt)[None], cind, 2 * F, 1.0, time_frac=0.0, freq_frac=0.0).alive[0]
    assert not gibbs.cut_batch(np.zeros_like(vis)[None], wgt[None], cind, 2 * F, 1.0, time_frac=0.0, freq_frac=0.0).alive[0]
    assert not gibbs.cut_batch(vis[None], wgt[None], cind, 2 * F, 1.0, time_frac=0.0, freq_frac=1.0).alive[0]


@pytest.mark.parametrize("tag", ["short", "long"])
def test_cut_complex64_matches_reference_golden(tag):
    """complex64 visibilities + float32 weights (draco's container dtypes) go up as they are and come out as the
    reference's own complex64 ``_cut_data`` result, widened -- bit for bit."""
    import os

    from draco_b200 import gibbs
    from _cases import C64_CASES

    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_cut_c64.npz"))
    vis, wgt = g[f"{tag}_vis"], g[f"{tag}_wgt"]
    T, F = vis.shape
    cind = np.arange(F) * 2 + 1
    for name, kw in C64_CASES:
        full = dict(time_frac=0.0, freq_frac=0.0, remove_mean=True, scale_freq=False, weight_boost=1.0)
        full.update(kw)
        cr = gibbs.cut_batch(vis[None], wgt[None], cind, 2 * F, 1.0, **full)
        assert cr.alive.tolist() == [True]
        assert np.array_equal(cr.keep_t[0], g[f"{tag}_{name}_nzt"])
        assert np.array_equal(cr.keep_f[0], g[f"{tag}_{name}_nzf"])
        d, ni, fs = _unpack(cr, 0, T, F)
        assert np.array_equal(fs, cind[g[f"{tag}_{name}_nzf"]])
        assert np.array_equal(d, g[f"{tag}_{name}_data"].astype(np.complex128)), name
        assert np.array_equal(ni, g[f"{tag}_{name}_weight"].astype(np.float64)), name
    with pytest.raises(ValueError):  # numpy's float32 std is not reproduced on the device: the task cuts on the host
        gibbs.cut_batch(vis[None], wgt[None], cind, 2 * F, 1.0, scale_freq=True)


def test_cut_complex64_ragged_matches_oracle():
    import oracle
    from draco_b200 import gibbs

    rng = np.random.default_rng(77)
    B, T, F = 9, 200, 70
    vis = ((rng.standard_normal((B, T, F)) + 1j * rng.standard_normal((B, T, F))) * 1e3 + 250.0).astype(np.complex64)
    wgt = rng.uniform(0.2, 3.0, (B, T, F)).astype(np.float32)
    for b in range(B):
        wgt[b][rng.random(T) < 0.1 * b] = 0.0
        wgt[b][:, rng.random(F) < 0.07 * b] = 0.0
    wgt[4] = 0.0  # dead baseline
    vis[wgt == 0] = 0.0
    cind = np.arange(F)
    kw = dict(time_frac=0.05, freq_frac=0.05, remove_mean=True, weight_boost=3.0)
    cr = gibbs.cut_batch(vis, wgt, cind, 2 * F, 1.0, **kw)
    n = 0
    for b in range(B):
        t = oracle.cut_data(vis[b].copy(), wgt[b].copy(), use_average_weights=True, **kw)
        assert (t is not None) == bool(cr.alive[b])
        if t is None:
            continue
        d, ni, fs = _unpack(cr, n, T, F)
        n += 1
        assert t[0].dtype == np.complex64
        assert np.array_equal(d, t[0].astype(np.complex128))
        assert np.array_equal(ni, t[1].astype(np.float64))
        assert np.array_equal(fs, cind[t[2]])


@pytest.mark.parametrize("wdtype", [np.float64, np.float32])
@pytest.mark.parametrize("kw", [dict(remove_mean=True), dict(remove_mean=False, time_frac=0.4, freq_frac=0.3),
                                dict(remove_mean=True, scale_freq=True, weight_boost=1e3, time_frac=0.2)])
def test_cut_matches_oracle_ragged(wdtype, kw):
    import oracle
    from draco_b200 import gibbs

    rng = np.random.default_rng(5)
    B, T, F = 41, 19, 300
    vis = rng.standard_normal((B, T, F)) + 1j * rng.standard_normal((B, T, F))
    wgt = rng.uniform(0.5, 2.0, (B, T, F)).astype(wdtype)
    wgt[rng.random((B, T, F)) < 0.15] = 0.0
    for b in range(B):  # whole channels / samples flagged, differently per baseline
        wgt[b, :, rng.choice(F, size=rng.integers(0, F // 2), replace=False)] = 0.0
        wgt[b, rng.choice(T, size=rng.integers(0, T // 2), replace=False), :] = 0.0
    wgt[3] = 0.0  # dead: no weight
    vis[7] = 0.0  # dead with remove_mean=False: all-zero data (constant data dies with remove_mean=True)
    vis[8] = 2.5 - 1j
    wgt[9] = -1.0  # negative weights count as flagged
    vis[wgt <= 0] = 0.0
    cind = np.sort(rng.choice(2 * F, size=F, replace=False))
    full = dict(time_frac=0.0, freq_frac=0.0, scale_freq=False, weight_boost=1.0)
    full.update(kw)
    cr = gibbs.cut_batch(vis, wgt, cind, 4 * F, 1.0, **full)
    n = 0
    for b in range(B):
        t = oracle.cut_data(vis[b].copy(), wgt[b].copy(), use_average_weights=True, **full)
        assert (t is not None) == bool(cr.alive[b]), b
        if t is None:
            continue
        d, w, nzf, nzt = t
        assert cr.index[n] == b
        assert np.array_equal(cr.keep_t[b], nzt) and np.array_equal(cr.keep_f[b], nzf)
        dd, ni, fs = _unpack(cr, n, T, F)
        assert dd.shape == d.shape
        assert np.array_equal(fs, cind[nzf])
        if full["scale_freq"]:
            np.testing.assert_allclose(dd, d, rtol=1e-12, atol=0)
        else:  # numpy reduces the (time-contiguous) columns pairwise; the kernel implements the same algorithm
            assert np.array_equal(dd, d), b
        assert np.array_equal(ni, np.asarray(w, dtype=np.float64)), b
        n += 1
    assert n == len(cr.index) and n >= B - 6


def test_cut_feeds_sampler_in_place():
    """dgibbs_cut_data -> dgibbs_run without a host round trip gives the chains of the host-cut batch bit for bit."""
    import oracle
    from draco_b200 import gibbs

    rng = np.random.default_rng(11)
    B, T, F, N = 9, 12, 33, 64
    vis = rng.standard_normal((B, T, F)) + 1j * rng.standard_normal((B, T, F))
    wgt = rng.uniform(0.5, 2.0, (B, T, F))
    wgt[:, :, [4, 5, 20]] = 0.0
    wgt[2] = 0.0
    wgt[5, 3:7] = 0.0
    vis[wgt == 0] = 0.0
    cind = np.arange(F)
    cr = gibbs.cut_batch(vis, wgt, cind, N, 10.0, time_frac=0.0, freq_frac=0.0, remove_mean=True)
    assert cr.alive.sum() == B - 1
    res = gibbs.run_batch(cr.batch, 12, seed=77, save_samples=True)
    datas, Nis, fsels = [], [], []
    for b in cr.index:
        d, w, nzf, _ = oracle.cut_data(vis[b].copy(), wgt[b].copy())
        datas.append(d)
        Nis.append(np.asarray(w, dtype=np.float64))
        fsels.append(cind[nzf])
    pb = gibbs.pack_batch(datas, Nis, fsels, N, 10.0, global_id=cr.index)
    ref = gibbs.run_batch(pb, 12, seed=77, save_samples=True)
    assert np.array_equal(res.samples.cpu().numpy(), ref.samples)
    assert np.array_equal(res.spectrum.cpu().numpy(), ref.spectrum)
    assert bool(res.success.all())


def test_cut_host_pointers_through_the_c_abi():
    """``on_device = 0``: the library stages the raw slab itself and copies every output back (plain C callers)."""
    import ctypes as C

    import oracle
    from draco_b200 import _lib, gibbs

    rng = np.random.default_rng(2)
    B, T, F = 5, 9, 40
    vis = np.ascontiguousarray(rng.standard_normal((B, T, F)) + 1j * rng.standard_normal((B, T, F)))
    wgt = rng.uniform(0.5, 2.0, (B, T, F)).astype(np.float32)
    wgt[:, :, 7] = 0.0
    wgt[1, 4] = 0.0
    wgt[3] = 0.0
    cind = np.arange(F, dtype=np.int64) + 3
    out = np.zeros((B, T * F), dtype=np.complex128)
    Ni, fsel = np.zeros((B, F)), np.zeros((B, F), dtype=np.int64)
    Tk, Mk = np.zeros(B, dtype=np.int32), np.zeros(B, dtype=np.int32)
    alive, kt, kf = np.zeros(B, dtype=np.uint8), np.zeros((B, T), dtype=np.uint8), np.zeros((B, F), dtype=np.uint8)
    c = _lib.Cut()
    c.struct_size = C.sizeof(_lib.Cut)
    c.on_device, c.B, c.T, c.nfreq, c.weight_f32, c.remove_mean, c.scale_freq = 0, B, T, F, 1, 1, 0
    c.time_frac, c.freq_frac, c.weight_boost = 0.0, 0.0, 2.0
    for name, a in (("data", vis), ("weight", wgt), ("channel_ind", cind), ("data_out", out), ("Ni_out", Ni), ("fsel_out", fsel),
                    ("T_out", Tk), ("M_out", Mk), ("alive", alive), ("keep_t", kt), ("keep_f", kf)):
        setattr(c, name, a.ctypes.data)
    h = gibbs.get_handle(0)
    h.check(h.lib.dgibbs_cut_data(h._h, C.byref(c)), "dgibbs_cut_data")
    assert alive.tolist() == [1, 1, 1, 0, 1]
    for b in (0, 1, 2, 4):
        d, w, nzf, nzt = oracle.cut_data(vis[b].copy(), wgt[b].copy(), weight_boost=2.0)
        assert (Tk[b], Mk[b]) == d.shape
        assert np.array_equal(out[b, : d.size].reshape(d.shape), d)
        assert np.array_equal(Ni[b, : Mk[b]], w.astype(np.float64))
        assert np.array_equal(fsel[b, : Mk[b]], cind[nzf])
        assert np.array_equal(kt[b].astype(bool), nzt) and np.array_equal(kf[b].astype(bool), nzf)
    c.struct_size = 8
    assert h.lib.dgibbs_cut_data(h._h, C.byref(c)) < 0  # bad struct size is refused
