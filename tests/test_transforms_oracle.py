"""The plain FFT delay transform and the delay-spectrum -> power-spectrum reduction: oracle against vectors produced by
the reference's own code (tests/golden/make_golden_transforms.py), and against the live reference when present."""
import os
import sys

import numpy as np
import pytest

import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_golden_transforms import FFT_CASES, VAR_CASES, fft_input, var_input  # noqa: E402


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "reference_transforms.npz"))


def test_fft_oracle_matches_reference_golden(golden):
    for i, (name, ns, N, window) in enumerate(FFT_CASES):
        x = fft_input(ns, N, 900 + i)
        keep = x.copy()
        got = oracle.delay_spectrum_fft(x, N, window)
        assert np.array_equal(x, keep)
        np.testing.assert_allclose(got, golden[name], rtol=0, atol=1e-14 * np.abs(golden[name]).max())


def test_to_power_spectrum_oracle_matches_reference_golden(golden):
    for i, (name, B, T, N, masked) in enumerate(VAR_CASES):
        spec, mask = var_input(B, T, N, masked, 950 + i)
        ps, flagged = oracle.delay_spectrum_to_power_spectrum(spec, mask)
        assert np.array_equal(ps, golden[f"{name}_spectrum"])
        assert np.array_equal(flagged, golden[f"{name}_mask"])
        if masked:
            assert flagged[1] and not ps[1].any() and not flagged[0]
    # without a mask: the plain variance (the reference itself raises TypeError there, see the oracle's docstring)
    spec, _ = var_input(3, 7, 5, False, 1)
    ps, flagged = oracle.delay_spectrum_to_power_spectrum(spec)
    np.testing.assert_allclose(ps, np.var(spec, axis=1))
    assert not flagged.any()


def test_live_reference_when_present(golden):
    from oracle import ref_loader

    if not ref_loader.available():
        pytest.skip("reference tree not present")
    import make_golden_transforms as mg

    ref_fft = mg.load_reference_fft()
    for i, (name, ns, N, window) in enumerate(FFT_CASES):
        assert np.array_equal(ref_fft(fft_input(ns, N, 900 + i), N, window), golden[name])
