"""CPU restatement of ``ConstructWienerDelayTransform.process`` (draco/analysis/powerspec.py:118-369).

TEST INFRASTRUCTURE ONLY: imported by ``tests/`` as the checker of the device
path behind ``include/dgwiener.h`` and by the CPU-baseline timing of ``scratch``/``profiles`` runs.
The product (``draco_b200/powerspec.py``) never imports it.

**Parity pinned:** ``tests/golden/make_golden_wiener_op.py`` executes the reference class itself
(exec of powerspec.py:118-369 over stand-in containers) and commits its operators as
``tests/golden/reference_wiener_op.npz``; ``tests/test_wiener_op_oracle.py`` checks this
restatement against them.

Array conventions (one polarisation, the unit the device call works on):
    freq [F] MHz, weight [nel, nra, F], beam_power [nel, F], filt / cov [nra, F, F]
    -> operator [nra, nel, ndelay, F] complex128 (the container stores it as complex64).
"""

from __future__ import annotations

import numpy as np
import scipy.linalg

from .gibbs_oracle import cosine_sum_window as window_generalised  # tools.py:547-601

def invert_no_zero(x):
    """caput.algorithms.invert_no_zero: 1/x where x != 0, else 0."""
    x = np.asarray(x)
    with np.errstate(divide="ignore", invalid="ignore"):
        return np.where(x == 0, 0.0, 1.0 / np.where(x == 0, 1.0, x))


def get_window(freq, window="uniform", lower=None, upper=None):
    """_get_window, powerspec.py:346-369."""
    frng = np.percentile(freq, [0, 100])
    if lower is not None:
        frng[0] = lower
    if upper is not None:
        frng[1] = upper
    x = (freq - frng[0]) / (frng[1] - frng[0])
    return window_generalised(x, window=window)


def delay_axis(freq, window):
    """Non-negative delays of the windowed band, powerspec.py:204-214.  Returns (tau, ntau)."""
    dfreq = np.median(np.abs(np.diff(freq)))
    ntau = int(np.sum(window > 0, dtype=int))
    tau = np.fft.fftshift(np.fft.fftfreq(ntau, d=dfreq))
    return tau[tau >= 0.0], ntau


def get_prior(tau, prior_amp=2.8e-5, prior_scale=0.0):
    """_get_prior, powerspec.py:325-344."""
    return prior_amp * np.exp(-2.0 * np.pi * prior_scale * np.abs(tau))


def fourier_operator(freq, tau, ntau):
    """Delay -> frequency DFT, powerspec.py:229."""
    return np.exp(2.0j * np.pi * np.outer(freq, tau)) / np.sqrt(ntau)


def construct_operator(freq, weight, beam_power, filt, cov, window, Sdiag, tau, ntau):
    """One polarisation of powerspec.py:229-319 (loops over elevation and right ascension).

    Raises ``numpy.linalg.LinAlgError`` where ``scipy.linalg.cho_factor`` does (powerspec.py:296).
    """
    nel, nra, nfreq = weight.shape
    win_mask = window > 0
    F = fourier_operator(freq, tau, ntau)  # freq, delay
    FT = F.T.conj()
    FSFT = (F * Sdiag[np.newaxis, :]) @ FT
    C = np.asarray(cov)
    K = np.asarray(filt)
    Cdiag = C[:, np.arange(nfreq), np.arange(nfreq)]
    out = np.zeros((nra, nel, tau.size, nfreq), dtype=np.complex128)
    for ee in range(nel):
        w = weight[ee]
        b = np.sqrt(beam_power[ee])
        r_noise = np.sqrt(invert_no_zero(w * Cdiag)) * win_mask
        N = C * (r_noise[:, :, np.newaxis] * r_noise[:, np.newaxis, :])
        M = win_mask * (w > 0)
        H = M[:, :, np.newaxis] * K
        HT = H.transpose(0, 2, 1).conj()
        A = H @ (FSFT * b[:, np.newaxis] * b) @ HT + N
        A_inv = np.zeros_like(A)
        for rr in range(nra):
            valid = np.flatnonzero(M[rr])
            if valid.size == 0:
                continue
            ix = np.ix_(valid, valid)
            cf = scipy.linalg.cho_factor(A[rr][ix], check_finite=False)
            A_inv[rr][ix] = scipy.linalg.cho_solve(cf, np.eye(valid.size), check_finite=False)
        RT = FT @ HT
        out[:, ee] = Sdiag[np.newaxis, :, np.newaxis] * (RT @ A_inv) * window
    return out


def apply_operator(op, map_, weight):
    """One polarisation of ``ApplyWienerDelayTransform.process`` (powerspec.py:430-455).

    op [nra, nel, ndelay, F] (complex64 as stored), map_ / weight [F, nra, nel] (the container's slices) ->
    spectrum [nel, nra, ndelay] complex128, weight [nel, nra, ndelay] float32.
    """
    nra, nel, nd, _ = op.shape
    spec = np.zeros((nel, nra, nd), dtype=np.complex128)
    sweight = np.zeros((nel, nra, nd), dtype=np.float32)
    for rr in range(nra):
        m = np.ascontiguousarray(map_[:, rr, :].T[..., np.newaxis])
        w = np.ascontiguousarray(weight[:, rr, :].T[..., np.newaxis])
        spec[:, rr, :] = np.matmul(op[rr], m)[..., 0]
        v = invert_no_zero(w)
        sweight[:, rr, :] = invert_no_zero(np.matmul(np.abs(op[rr]) ** 2, v)[..., 0])
    return spec, sweight
