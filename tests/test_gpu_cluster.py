"""One chain shared by a thread-block cluster of C CTAs (dense branch, `cluster_size` of dgibbs_problem).

The cluster ranks divide the sample rows, the row tiles of every block column and the delays of the
power-spectrum draw, but every tile is computed by the code and data of the single-CTA kernel: the
samples must be BIT-identical for C = 1, 2, 4, and therefore hold the 1e-6 parity with the oracle
under injected draws for every C.  The automatic plan (full waves on single CTAs, the remainder on
clusters) must give the same bits as well -- that is what makes the 1/2/4/8-GPU results identical.
"""
import numpy as np
import pytest

import oracle
from _cases import CASES, make_case, recorded_chain

pytestmark = pytest.mark.gpu

RTOL_SAMPLES = 1e-6


@pytest.fixture(scope="module")
def gibbs():
    from draco_b200 import gibbs as g

    g.get_handle(0)
    return g


@pytest.mark.parametrize("C", [1, 2, 4])
@pytest.mark.parametrize("name", ["f_real_cfg1", "f_cplx_n64", "f_real_n100_hann", "f_real_nowin_nyq"])
def test_golden_chain_on_clusters(gibbs, golden, name, C):
    """Reference-generated golden chains through clusters of C CTAs (injected draws, rtol 1e-6)."""
    if name not in CASES:
        pytest.skip("case not in the golden set")
    _, N, M, T, cplx, window, niter, nyq = CASES[name]
    data, Ni, fsel, S0 = make_case(N, M, T, cplx, nyq, seed=11)
    n_sys = 2 * N if cplx else N
    pb = gibbs.pack_batch([data], [Ni], [fsel], N, S0[None], window, cplx)
    w1, w2, c2 = gibbs._host_draws(np.random.default_rng(7), niter, n_sys, M, T, N)
    res = gibbs.run_batch(pb, niter, draws=gibbs.pack_draws(pb, niter, [w1], [w2], [c2]), save_samples=True,
                          cluster_size=C)
    ref = golden[f"{name}_samples"]
    assert bool(res.success[0]) == bool(golden[f"{name}_success"])
    np.testing.assert_allclose(res.samples[niter - len(ref):, 0], ref, rtol=RTOL_SAMPLES)


@pytest.mark.parametrize("C", [2, 4])
def test_cfg1_geometry_full_chain_on_clusters(gibbs, C):
    """BASELINE config 1 geometry (N=256: two block columns), full nsamp=20 chain against the oracle."""
    from draco_b200 import synth

    cfg = synth.CONFIGS["cfg1"]
    nb = 5
    freq, vis, wgt = synth.make_stack(cfg, nbase=nb)
    delays, chan = oracle.calculate_delays(freq)
    N = len(delays)
    datas, Nis, fsels, refs, draws = [], [], [], [], []
    for b in range(nb):
        d, w, nzf, _ = oracle.cut_data(vis[b], wgt[b])
        s, ok, dr = recorded_chain(d, N, w, np.full(N, 10.0), cfg.window, chan[nzf], cfg.nsamp, False, seed=300 + b)
        assert ok
        datas.append(d); Nis.append(w); fsels.append(chan[nzf]); refs.append(s); draws.append(dr)
    pb = gibbs.pack_batch(datas, Nis, fsels, N, np.full((nb, N), 10.0), window=cfg.window)
    res = gibbs.run_batch(pb, cfg.nsamp, draws=gibbs.pack_draws(pb, cfg.nsamp, *zip(*draws)), save_samples=True,
                          cluster_size=C)
    assert res.success.all()
    for b in range(nb):
        np.testing.assert_allclose(res.samples[:, b], refs[b], rtol=RTOL_SAMPLES)


def _philox_batch(gibbs, N, nb, T, window="nuttall", cplx=False, seed=3, frac=0.9):
    nch = N if cplx else N // 2 + 1
    rng = np.random.default_rng(seed)
    datas, Nis, fsels = [], [], []
    for b in range(nb):
        M = int(frac * nch) - (b % 3)
        fsel = np.sort(rng.choice(nch - 1, size=M, replace=False))
        Tb = T - (b % 2) * 3
        datas.append((rng.standard_normal((Tb, M)) + 1j * rng.standard_normal((Tb, M))) * 0.3)
        Nis.append(rng.uniform(0.5, 2.0, M) * 50.0)
        fsels.append(fsel)
    return gibbs.pack_batch(datas, Nis, fsels, N, np.full((nb, N), 10.0), window, cplx)


@pytest.mark.parametrize("N,T,cplx,nofft", [(512, 24, False, False), (640, 40, False, False), (256, 150, False, False),
                                            (128, 16, True, False), (512, 24, False, True)])
def test_cluster_sizes_bit_identical(gibbs, monkeypatch, N, T, cplx, nofft):
    """Philox chains: C = 2 and 4 reproduce the single-CTA samples bit for bit (power-of-two and other N, more than
    one right-hand-side tile, complex time domain, the DMMA projection instead of the FFT)."""
    if nofft:
        monkeypatch.setenv("DGIBBS_NO_FFT", "1")
    pb = _philox_batch(gibbs, N, 7, T, cplx=cplx).to_device(0)
    ref = gibbs.run_batch(pb, 6, seed=17, save_samples=True, cluster_size=1, force_branch=gibbs.BRANCH_DELAY_SPACE)
    assert bool(ref.success.all())
    for C in (2, 4):
        res = gibbs.run_batch(pb, 6, seed=17, save_samples=True, cluster_size=C, force_branch=gibbs.BRANCH_DELAY_SPACE)
        assert bool((res.samples == ref.samples).all()), f"C={C} differs"
        assert bool((res.spectrum == ref.spectrum).all())


def test_planned_tail_bit_identical(gibbs):
    """More chains than SMs: the planner runs the full waves on single CTAs and the remainder on clusters; same bits
    as the all-single-CTA launch, and every chain is done exactly once."""
    import torch

    nsm = torch.cuda.get_device_properties(0).multi_processor_count
    nb = nsm + 5
    pb = _philox_batch(gibbs, 256, nb, 8, frac=0.8).to_device(0)
    ref = gibbs.run_batch(pb, 3, seed=5, save_samples=True, cluster_size=1, force_branch=gibbs.BRANCH_DELAY_SPACE)
    res = gibbs.run_batch(pb, 3, seed=5, save_samples=True, cluster_size=0, force_branch=gibbs.BRANCH_DELAY_SPACE)
    assert int(res.timing["sampler_launches"]) == 2  # full wave + clustered remainder
    assert bool((res.n_done == 3).all()) and bool(res.success.all())
    assert bool((res.samples == ref.samples).all())


@pytest.mark.parametrize("C", [2, 4])
def test_failure_semantics_on_clusters(gibbs, golden, C):
    """A non-positive pivot stops the chain at the same iteration on every rank (delay.py:1891-1896)."""
    data, Ni, fsel, S0 = make_case(64, 29, 8, False, False, seed=11)
    _, _, (w1, w2, c2) = recorded_chain(data, 64, Ni, S0, "nuttall", fsel, 6, False)
    c2 = c2.copy()
    c2[2] = -c2[2]
    pb = gibbs.pack_batch([data, data], [Ni, Ni], [fsel, fsel], 64, np.stack([S0, S0]))
    res = gibbs.run_batch(pb, 6, draws=gibbs.pack_draws(pb, 6, [w1, w1], [w2, w2], [c2, c2]), save_samples=True,
                          cluster_size=C)
    ref = golden["fail_samples"]
    for b in range(2):
        assert int(res.n_done[b]) == len(ref) == 3 and not res.success[b]
        np.testing.assert_allclose(res.samples[-3:, b], ref, rtol=RTOL_SAMPLES)
        assert np.all(res.samples[:3, b] == 0.0)


@pytest.mark.parametrize("C", [2, 4])
def test_wiener_on_clusters(gibbs, C):
    """dgibbs_wiener (single-baseline latency path) on clusters: bit-identical to one CTA."""
    pb = _philox_batch(gibbs, 384, 3, 20, seed=9)
    ref, ok0 = gibbs.wiener_batch(pb, cluster_size=1)
    out, ok = gibbs.wiener_batch(pb, cluster_size=C)
    assert ok0.all() and ok.all()
    for a, b in zip(ref, out):
        assert np.array_equal(a, b)
