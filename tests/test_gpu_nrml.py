"""NRML / maximum-posterior estimator on the GPU: the device ``LogLikePS`` against the numpy oracle
and against outputs of the reference's own classes (tests/golden/reference_nrml.npz)."""
import os
import sys

import numpy as np
import pytest
import scipy.linalg as la

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(__file__)
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_golden_nrml import NRML_CASES, make_nrml_case  # noqa: E402

from oracle import nrml_oracle  # noqa: E402


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "reference_nrml.npz"))


def _close(a, b, rtol):
    np.testing.assert_allclose(a, b, rtol=rtol, atol=rtol * np.abs(b).max())


@pytest.mark.parametrize("k", range(len(NRML_CASES)))
def test_device_objective_matches_reference(golden, k):
    from draco_b200 import delayopt

    name, N, M, T, window, nflag, _ = NRML_CASES[k]
    data, Ni, fsel = make_nrml_case(N, M, T, nflag, 100 + k)
    _, X, F, Nm = nrml_oracle.maxpost_inputs(data, N, Ni, window, fsel)
    like = delayopt.LogLikePS(X, F, Nm, T, exact_hessian=True, bounds=(1e-15, 1e10))
    fish = delayopt.LogLikePS(X, F, Nm, T, exact_hessian=False, bounds=(1e-15, 1e10))
    assert like.MF.shape[0] == M - nflag  # fully masked channels are dropped (delayopt.py:103-104)
    for j in range(2):
        x = golden[f"{name}_x{j}"]
        np.testing.assert_allclose(like.value(x), golden[f"{name}_value{j}"], rtol=1e-10)
        _close(like.gradient(x), golden[f"{name}_grad{j}"], 1e-9)
        _close(like.hessian(x), golden[f"{name}_hess{j}"], 1e-9)
        _close(fish.hessian(x), golden[f"{name}_fisher{j}"], 1e-9)
    like.close()
    fish.close()


def test_device_objective_matches_oracle_midsize():
    """M = 200 channels, N = 512 delays: beyond the golden sizes, against the numpy oracle; bounds clip the input."""
    from draco_b200 import delayopt

    N, M, T = 512, 200, 40
    data, Ni, fsel = make_nrml_case(N, M, T, 5, 7)
    _, X, F, Nm = nrml_oracle.maxpost_inputs(data, N, Ni, "nuttall", fsel)
    dev = delayopt.LogLikePS(X, F, Nm, T, bounds=(1e-3, 1e3))
    ora = nrml_oracle.LogLike(X, F, Nm, T, bounds=(1e-3, 1e3))
    x = np.random.default_rng(3).normal(0.0, 4.0, N)  # some entries beyond log(1e3) = 6.9: clipped
    np.testing.assert_allclose(dev.value(x), ora.value(x), rtol=1e-10)
    _close(dev.gradient(x), ora.gradient(x), 1e-8)
    _close(dev.hessian(x), ora.hessian(x), 1e-8)
    assert dev.evals == 3 and dev.eval_flop > 0
    dev.close()


@pytest.mark.parametrize("window", ["nuttall", None])
def test_device_pinv_start_matches_scipy(window):
    """The eigendecomposition-based starting point against the reference's SVD pseudo-inverse (delayopt.py:505-506)."""
    from draco_b200 import delayopt

    N, M, T = 256, 100, 24
    data, Ni, fsel = make_nrml_case(N, M, T, 4, 11)
    dataw, X, F, Nm = nrml_oracle.maxpost_inputs(data, N, Ni, window, fsel)
    ref = (dataw @ la.pinv(F.T, rtol=1e-3)).var(axis=0)
    like = delayopt.LogLikePS(X, F, Nm, T)
    got = like.initial_guess(dataw[:, like._rows], rtol=1e-3)
    np.testing.assert_allclose(got, ref, rtol=1e-8, atol=1e-10 * ref.max())
    like.value(np.log(ref))  # the object still evaluates after its buffers served as scratch
    like.close()


def test_not_positive_definite_raises_linalgerror():
    from draco_b200 import delayopt

    N, M, T = 32, 17, 12
    data, Ni, fsel = make_nrml_case(N, M, T, 0, 1)
    _, X, F, Nm = nrml_oracle.maxpost_inputs(data, N, Ni, None, fsel)
    like = delayopt.LogLikePS(X, F, -np.abs(Nm) * 1e6, T)  # negative noise variances: C is indefinite
    with pytest.raises(la.LinAlgError):
        like.value(np.zeros(N))
    like.close()


@pytest.mark.parametrize("k", range(len(NRML_CASES)))
def test_maxpost_matches_reference_chain(golden, k):
    from draco_b200 import delayopt

    name, N, M, T, window, nflag, maxiter = NRML_CASES[k]
    data, Ni, fsel = make_nrml_case(N, M, T, nflag, 100 + k)
    stats = {}
    samples, ok = delayopt.delay_power_spectrum_maxpost(data, N, Ni, None, window=window, fsel=fsel, maxiter=maxiter,
                                                        tol=1e-3, stats=stats)
    ref = golden[f"{name}_samples"]
    assert ok == bool(golden[f"{name}_success"])
    assert len(samples) == len(ref)
    np.testing.assert_allclose(np.array(samples), ref, rtol=1e-6)
    assert stats["evals"] > 0
    samples_h, ok_h = delayopt.delay_power_spectrum_maxpost(data, N, Ni, None, window=window, fsel=fsel, maxiter=maxiter,
                                                            tol=1e-3, host_pinv=True)
    assert ok_h == ok and len(samples_h) == len(ref)
    np.testing.assert_allclose(np.array(samples_h), ref, rtol=1e-6)
    with pytest.raises(ValueError):
        delayopt.delay_power_spectrum_maxpost(data, N, Ni, fsel=fsel[:-1])


def test_nrml_task():
    from draco_b200 import delay
    from test_gpu_task import _stream

    ss, vis, wgt = _stream(nfreq=32, nstack=3, nra=10, dead=(1,))
    task = delay.DelayPowerSpectrumNRML.from_config(dict(sample_axis="ra", nsamp=12, save_samples=True, save_spectrum_mask=True))
    out = task.process(ss)
    delays, chan = __import__("oracle").calculate_delays(ss.freq)
    N = len(delays)
    import oracle

    for b in (0, 2):
        d, w, nzf, _ = oracle.cut_data(vis[b], wgt[b])
        samples, ok = nrml_oracle.maxpost_chain(d, N, w, None, window="nuttall", fsel=chan[nzf], maxiter=12, tol=1e-3)
        np.testing.assert_allclose(out.spectrum[:].local_array[b], np.fft.fftshift(samples[-1]), rtol=1e-6)
        assert bool(out.datasets["spectrum_mask"][:].local_array[b]) == (not ok)
    assert out.datasets["spectrum_mask"][:].local_array[1]
