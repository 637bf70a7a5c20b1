"""Cross-power-spectrum sampler on the GPU (through the C ABI) against the reference's golden chains, the
CPU oracle with injected draws, and -- native RNG -- the oracle's posterior."""
import os
import sys

import numpy as np
import pytest

import oracle

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden_cross import CROSS_CASES, DRAW_SEED, make_cross_case  # noqa: E402

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
RTOL = 1e-6  # north_star: every sample agrees to rtol 1e-6 at fp64 under injected draws


@pytest.fixture(scope="module")
def gibbs():
    from draco_b200 import gibbs as g

    g.get_handle(0)
    return g


@pytest.fixture(scope="module")
def xgolden():
    return np.load(os.path.join(HERE, "golden", "reference_cross.npz"), allow_pickle=False)


def relerr(a, b):
    """max |a - b| relative to the largest entry of each [nd, nd] block's delay slice"""
    scale = np.abs(b).max(axis=(-3, -2), keepdims=True)
    return np.max(np.abs(a - b) / scale)


@pytest.mark.parametrize("case", CROSS_CASES, ids=[c[0] for c in CROSS_CASES])
def test_golden_chain(gibbs, xgolden, case):
    name, nd, N, M, T, window, niter = case
    data, Ni, fsel, S0 = make_cross_case(nd, N, M, T, seed=31)
    spec = gibbs.delay_spectrum_gibbs_cross(data, N, Ni, S0, window, fsel, niter, np.random.default_rng(DRAW_SEED))
    ref = xgolden[f"{name}_samples"]
    assert len(spec) == niter
    assert relerr(np.array(spec), ref) < RTOL


def test_batch_of_problems_matches_oracle(gibbs):
    """Several independent problems in one launch, each with its own injected draws."""
    nd, N, M, T, niter, B = 2, 32, 24, 6, 3, 5
    datas, Nis, S0s, refs, draws = [], [], [], [], []
    fsel = make_cross_case(nd, N, M, T, seed=100)[2]
    for b in range(B):
        data, Ni, _, S0 = make_cross_case(nd, N, M, T, seed=100 + b)
        rec = oracle.CrossRecordingDraws(np.random.default_rng(900 + b))
        refs.append(np.array(oracle.cross_gibbs_chain(data.copy(), N, Ni, S0.copy(), "nuttall", fsel, niter, rec)))
        draws.append(rec.packed(niter, N, nd, M, T))
        datas.append(data); Nis.append(Ni); S0s.append(S0)
    packed = tuple(np.stack([d[i] for d in draws]) for i in range(4))
    samples, n_done, ok = gibbs.cross_batch(np.stack(datas), N, np.stack(Nis), np.stack(S0s), "nuttall", fsel, niter,
                                            draws=packed)
    assert ok.all() and np.all(n_done == niter)
    for b in range(B):
        assert relerr(samples[b], refs[b]) < RTOL


def test_failure_raises_like_reference(gibbs):
    data, Ni, fsel, S0 = make_cross_case(2, 32, 32, 8, seed=31)
    with pytest.raises(RuntimeError, match="Exiting earlier as singular"):
        gibbs.delay_spectrum_gibbs_cross(data, 32, Ni, -S0, "nuttall", fsel, 2, np.random.default_rng(DRAW_SEED))
    with pytest.raises(ValueError, match="Length of frequency selection"):
        gibbs.delay_spectrum_gibbs_cross(data, 32, Ni, S0, "nuttall", fsel[:-1], 2, np.random.default_rng(1))


def test_single_dataset_agrees_with_power_spectrum_limit(gibbs):
    """nd = 1: samples are real positive 1x1 'matrices'."""
    data, Ni, fsel, S0 = make_cross_case(1, 16, 12, 5, seed=3)
    spec = gibbs.delay_spectrum_gibbs_cross(data, 16, Ni, S0, "hann", fsel, 4, None, seed=11)
    s = np.array(spec)
    assert s.shape == (4, 1, 1, 16) and np.all(s.real > 0) and np.all(np.abs(s.imag) < 1e-12 * s.real)


def test_philox_posterior_statistics(gibbs):
    """Native RNG: 48 device chains vs 48 oracle chains (numpy RNG) of the same problem; posterior means of
    log S_dd and of the cross-coherence agree within 3 sigma-ish bounds and KS tests pass."""
    from scipy import stats

    nd, N, M, T, niter, nb = 2, 16, 16, 12, 60, 48
    data, Ni, fsel, S0 = make_cross_case(nd, N, M, T, seed=77)
    samples, n_done, ok = gibbs.cross_batch(np.broadcast_to(data, (nb, nd, T, M)), N, Ni, S0, "nuttall", fsel, niter,
                                            seed=5, global_id=np.arange(nb))
    assert ok.all()

    def summarise(s):  # s [niter, nd, nd, N] -> per-chain means over the second half
        s = s[niter // 2:]
        p0, p1 = np.log(s[:, 0, 0].real), np.log(s[:, 1, 1].real)
        coh = (s[:, 0, 1] / np.sqrt(s[:, 0, 0].real * s[:, 1, 1].real)).real
        return np.stack([p0.mean(0), p1.mean(0), coh.mean(0)])

    dev = np.stack([summarise(samples[c]) for c in range(nb)])
    ora = np.stack([summarise(np.array(oracle.cross_gibbs_chain(data.copy(), N, Ni, S0.copy(), "nuttall", fsel, niter,
                                                                 np.random.default_rng(4000 + c)))) for c in range(nb)])
    se = np.sqrt(dev.var(axis=0, ddof=1) / nb + ora.var(axis=0, ddof=1) / nb)
    zsc = (dev.mean(axis=0) - ora.mean(axis=0)) / se
    assert np.mean(np.abs(zsc) < 3.0) >= 0.95 and np.max(np.abs(zsc)) < 4.5, zsc
    alpha = 0.01 / (3 * N)
    for q in range(3):
        for k in range(N):
            assert stats.ks_2samp(dev[:, q, k], ora[:, q, k]).pvalue > alpha, (q, k)
