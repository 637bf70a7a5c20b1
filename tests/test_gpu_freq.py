"""The frequency-space branch (_draw_signal_sample_t, delay.py:1821-1860; taken when M <= N/4): injected-draw
parity of the dedicated 2M x 2M kernel with the oracle, branch selection, and statistical agreement of its
native draws with the delay-space kernel."""
import numpy as np
import pytest

import oracle
from _cases import make_case, recorded_chain

pytestmark = pytest.mark.gpu
RTOL = 1e-6
DENSE, FREQ = 1, 2


@pytest.fixture(scope="module")
def gibbs():
    from draco_b200 import gibbs as g

    g.get_handle(0)
    return g


@pytest.mark.parametrize("N,M,T,window,niter", [(64, 12, 8, "nuttall", 5), (256, 40, 7, "blackman_harris", 4),
                                                (1024, 200, 6, "nuttall", 3), (512, 128, 5, None, 3)])
def test_injected_parity_multi_tile(gibbs, N, M, T, window, niter):
    """2M = 24 ... 400 (1 to 4 tiles of the 2M system), odd and even sample counts."""
    data, Ni, fsel, S0 = make_case(N, M, T, False, False, seed=300 + N)
    ref, ok, (w1, w2, c2) = recorded_chain(data, N, Ni, S0, window, fsel, niter, False)
    assert ok
    pb = gibbs.pack_batch([data], [Ni], [fsel], N, S0[None], window=window)
    draws = gibbs.pack_draws(pb, niter, [w1], [w2], [c2])
    res = gibbs.run_batch(pb, niter, draws=draws, save_samples=True)
    assert int(res.branch[0]) == FREQ and res.success[0]
    np.testing.assert_allclose(res.samples[:, 0], ref, rtol=RTOL)
    # the same posterior draw through the delay-space system (what the kernel replaced), forced
    forced = gibbs.run_batch(pb, niter, draws=draws, save_samples=True, force_branch=DENSE)
    assert int(forced.branch[0]) == DENSE
    np.testing.assert_allclose(forced.samples[:, 0], ref, rtol=RTOL)


def test_branch_rule_and_mixed_batch(gibbs):
    """len(fsel) > 0.25 N -> delay space, else frequency space (delay.py:1884-1886), per baseline of one batch."""
    N, niter = 128, 4
    shapes = [(32, 6), (33, 6), (10, 5), (64, 8), (20, 3)]
    datas, Nis, fsels, refs, draws = [], [], [], [], []
    for i, (M, T) in enumerate(shapes):
        data, Ni, fsel, S0 = make_case(N, M, T, False, False, seed=500 + i)
        s, ok, dr = recorded_chain(data, N, Ni, S0, "nuttall", fsel, niter, False, seed=600 + i)
        datas.append(data); Nis.append(Ni); fsels.append(fsel); refs.append(s); draws.append(dr)
    pb = gibbs.pack_batch(datas, Nis, fsels, N, np.full((len(shapes), N), 10.0))
    res = gibbs.run_batch(pb, niter, draws=gibbs.pack_draws(pb, niter, *zip(*draws)), save_samples=True)
    assert list(res.branch) == [FREQ, DENSE, FREQ, DENSE, FREQ]
    for b, s in enumerate(refs):
        np.testing.assert_allclose(res.samples[:, b], s, rtol=RTOL)
        np.testing.assert_allclose(res.spectrum[b], np.fft.fftshift(np.median(s[-2:], axis=0)), rtol=RTOL)


def test_failure_semantics_freq(gibbs):
    """A negative chi2 makes S negative; in this branch sqrt(S) is then NaN and so is the whole 2M system.
    Reference LAPACK dpotrf stops on a NaN pivot (`ajj <= 0 or isnan(ajj)`) -> LinAlgError -> chain truncated
    (delay.py:1891-1896); the device follows that.  (OpenBLAS 0.3.30's own potrf lets NaN through, so the
    oracle run here carries on with NaN samples: only the samples before the failure are compared.)"""
    N, M, T, niter = 64, 12, 8, 6
    data, Ni, fsel, S0 = make_case(N, M, T, False, False, seed=11)
    _, _, (w1, w2, c2) = recorded_chain(data, N, Ni, S0, "nuttall", fsel, niter, False)
    c2 = c2.copy()
    c2[1] = -c2[1]
    with np.errstate(invalid="ignore"):
        ref, _ = oracle.gibbs_chain(data, N, Ni, S0, "nuttall", fsel, niter, oracle.InjectedDraws(w1, w2, c2), False)
    assert np.all(np.array(ref[1]) < 0)
    pb = gibbs.pack_batch([data], [Ni], [fsel], N, S0[None])
    res = gibbs.run_batch(pb, niter, draws=gibbs.pack_draws(pb, niter, [w1], [w2], [c2]), save_samples=True)
    assert int(res.branch[0]) == FREQ and int(res.n_done[0]) == 2 and not res.success[0]
    np.testing.assert_allclose(res.samples[-2:, 0], np.array(ref[:2]), rtol=RTOL)
    assert np.all(res.samples[:-2, 0] == 0.0)


def test_native_draws_agree_with_delay_space_kernel(gibbs):
    """Device RNG: the frequency-space kernel and the delay-space kernel sample the same posterior."""
    from scipy import stats

    N, M, T, niter, nb = 256, 48, 12, 60, 48
    data, Ni, fsel, S0 = make_case(N, M, T, False, False, seed=42)
    pb = gibbs.pack_batch([data] * nb, [Ni] * nb, [fsel] * nb, N, np.full((nb, N), 10.0), global_id=np.arange(nb))
    a = gibbs.run_batch(pb, niter, seed=5, save_samples=True)
    b = gibbs.run_batch(pb, niter, seed=6, save_samples=True, force_branch=DENSE)
    assert np.all(a.branch == FREQ) and np.all(b.branch == DENSE) and a.success.all() and b.success.all()
    sa, sb = np.log(a.samples[niter // 2:]), np.log(b.samples[niter // 2:])
    ma, mb = sa.mean(axis=0), sb.mean(axis=0)
    se = np.sqrt(ma.var(axis=0, ddof=1) / nb + mb.var(axis=0, ddof=1) / nb)
    z = (ma.mean(axis=0) - mb.mean(axis=0)) / se
    assert np.mean(np.abs(z) < 3.0) >= 0.97 and np.max(np.abs(z)) < 4.7, (np.abs(z).max(), np.mean(np.abs(z) < 3))
    for k in range(0, N, 16):
        assert stats.ks_2samp(sa[-1, :, k], sb[-1, :, k]).pvalue > 0.01 / 16, k
