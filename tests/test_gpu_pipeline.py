"""Host-pointer batches large enough to be split go through the pipelined path of dgibbs_run (chunked upload on
its own stream, double-buffered, results drained on a third stream) and may ship complex64 visibilities
(`data_c64`).  Neither may change a bit of the result: every chain is keyed on its global baseline id."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gibbs():
    from draco_b200 import gibbs as g

    g.get_handle(0)
    return g


def _ragged(gibbs, nb, N=64, seed=1, uniform=False, c64=False):
    rng = np.random.default_rng(seed)
    nch = N // 2 + 1
    datas, Nis, fsels = [], [], []
    for b in range(nb):
        if uniform:  # diagonal(+rank-1) branch: every channel but Nyquist, uniform weights, no window
            M, T = nch - 1, 6
            fsel = np.arange(M)
            Ni = np.full(M, 25.0)
        else:
            M, T = nch - 3 - (b % 5), 4 + (b % 3)
            fsel = np.sort(rng.choice(nch - 1, size=M, replace=False))
            Ni = rng.uniform(0.5, 2.0, M) * 30.0
        d = (rng.standard_normal((T, M)) + 1j * rng.standard_normal((T, M))) * 0.4
        if c64:
            d = d.astype(np.complex64)
        datas.append(d); Nis.append(Ni); fsels.append(fsel)
    return datas, Nis, fsels


@pytest.mark.parametrize("uniform", [False, True])
def test_pipelined_host_path_bit_identical(gibbs, monkeypatch, uniform):
    import torch

    nsm = torch.cuda.get_device_properties(0).multi_processor_count
    nb, N, niter = 2 * nsm + 37, 64, 5
    datas, Nis, fsels = _ragged(gibbs, nb, N, uniform=uniform)
    window = None if uniform else "nuttall"
    gid = np.arange(1000, 1000 + nb)
    pb = gibbs.pack_batch(datas, Nis, fsels, N, np.full((nb, N), 10.0), window=window, global_id=gid)
    monkeypatch.setenv("DGIBBS_PIPELINE_CHUNK", "0")
    ref = gibbs.run_batch(pb, niter, seed=9, save_samples=True, want_final_S=True)
    monkeypatch.setenv("DGIBBS_PIPELINE_CHUNK", str(nsm))  # three chunks: nsm, nsm, 37
    res = gibbs.run_batch(pb, niter, seed=9, save_samples=True, want_final_S=True)
    assert res.success.all() and ref.success.all()
    assert np.array_equal(res.spectrum, ref.spectrum)
    assert np.array_equal(res.samples, ref.samples)
    assert np.array_equal(res.final_S, ref.final_S)
    assert np.array_equal(res.n_done, ref.n_done) and np.array_equal(res.branch, ref.branch)
    assert (res.branch == (gibbs.BRANCH_DIAGONAL if uniform else gibbs.BRANCH_DELAY_SPACE)).all()
    assert int(res.timing["n_dense"] + res.timing["n_diag"]) == nb
    # and the device-resident call agrees with both
    dev = gibbs.run_batch(pb.to_device(0), niter, seed=9, save_samples=True)
    assert np.array_equal(dev.spectrum.cpu().numpy(), ref.spectrum)


def test_pipelined_default_global_ids(gibbs, monkeypatch):
    """Without global_id the Philox key is the position in the WHOLE batch, not in the chunk."""
    import torch

    nsm = torch.cuda.get_device_properties(0).multi_processor_count
    nb, N = 2 * nsm + 5, 32
    datas, Nis, fsels = _ragged(gibbs, nb, N, seed=4)
    pb = gibbs.pack_batch(datas, Nis, fsels, N, np.full((nb, N), 10.0))
    monkeypatch.setenv("DGIBBS_PIPELINE_CHUNK", "0")
    ref = gibbs.run_batch(pb, 3, seed=2)
    monkeypatch.setenv("DGIBBS_PIPELINE_CHUNK", str(nsm))
    res = gibbs.run_batch(pb, 3, seed=2)
    assert np.array_equal(res.spectrum, ref.spectrum)


@pytest.mark.parametrize("uniform", [False, True])
def test_complex64_visibilities(gibbs, uniform):
    """complex64 input (draco's container dtype) is widened exactly on the device: same bits as the complex128 copy,
    from host buffers and from device-resident ones."""
    nb, N = 9, 128
    datas, Nis, fsels = _ragged(gibbs, nb, N, seed=7, uniform=uniform, c64=True)
    window = None if uniform else "hann"
    M, T = datas[0].shape[1], datas[0].shape[0]
    if uniform:
        vis = np.stack(datas)
        pb32 = gibbs.pack_uniform(vis, np.stack(Nis), fsels[0], N, np.full((nb, N), 10.0), window=window)
        assert pb32.data.dtype == np.complex64
        pb64 = gibbs.pack_uniform(vis.astype(np.complex128), np.stack(Nis), fsels[0], N, np.full((nb, N), 10.0), window=window)
    else:
        pb64 = gibbs.pack_batch(datas, Nis, fsels, N, np.full((nb, N), 10.0), window=window)
        assert pb64.data.dtype == np.complex128
        import dataclasses

        pb32 = dataclasses.replace(pb64, data=pb64.data.astype(np.complex64))
    ref = gibbs.run_batch(pb64, 4, seed=3, save_samples=True)
    res = gibbs.run_batch(pb32, 4, seed=3, save_samples=True)
    assert ref.success.all()
    assert np.array_equal(res.samples, ref.samples)
    dev = gibbs.run_batch(pb32.to_device(0), 4, seed=3, save_samples=True)
    assert np.array_equal(dev.samples.cpu().numpy(), ref.samples)
