"""CPU restatement of the plain FFT delay transform and of the delay-spectrum -> power-spectrum reduction.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Parity status: PINNED on tests/golden/reference_transforms.npz,
produced by the reference's own code (tests/golden/make_golden_transforms.py).
"""

from __future__ import annotations

import numpy as np

from .gibbs_oracle import cosine_sum_window


def delay_spectrum_fft(data, N, window="nuttall"):
    """``delay_spectrum_fft`` (draco/analysis/delay.py:2102-2130): window at ``x = i / N``, then an inverse FFT in
    numpy's convention (the reference calls caput's FFTW wrapper, same convention).  ``data[nsample, N]`` is not
    modified (the reference multiplies it in place)."""
    data = np.asarray(data, dtype=np.complex128)
    if window is not None:
        data = data * cosine_sum_window(np.arange(N) / N, window)[np.newaxis]
    return np.fft.ifft(data, axis=-1)


def delay_spectrum_to_power_spectrum(spec, mask=None):
    """The arithmetic of ``DelaySpectrumToPowerSpectrum.process`` (delay.py:1085-1104): variance over the sample axis
    of ``spec[baseline, sample, delay]`` restricted to the unmasked samples; NaN (no valid sample) -> 0 and flagged.
    Returns ``(ps[baseline, delay], flagged[baseline])``.  Without a mask the reference passes ``where=None`` to
    ``np.var``, which numpy rejects (TypeError); the evident intent, the plain variance, is returned here."""
    spec = np.asarray(spec)
    if mask is None:
        return np.var(spec, axis=1), np.zeros(spec.shape[0], dtype=bool)
    w = ~np.asarray(mask, dtype=bool)[..., np.newaxis]
    with np.errstate(all="ignore"):
        ps = np.var(spec, axis=1, where=w)
    nans = np.isnan(ps)
    ps[nans] = 0.0
    return ps, np.any(nans, axis=-1)
