"""numpy/scipy restatement of draco's delay power-spectrum Gibbs sampler (CPU oracle).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Every function cites the
reference lines whose behaviour it restates; paths are relative to ``/root/reference``.
Arithmetic is deliberately kept in the same association order as the reference so the
restatement agrees with it to a few ulp (checked against ``tests/golden``).

Third-party arithmetic used exactly as the reference uses it: LAPACK ``dpotrf`` /
``dpotrs`` through ``scipy.linalg.cho_factor`` / ``cho_solve`` (reference pins
``scipy>=1.13``; scipy 1.18.1 + OpenBLAS here) and ``numpy.random.Generator``.
"""

from __future__ import annotations

import numpy as np
import scipy.linalg as la

# ---------------------------------------------------------------------------
# small helpers
# ---------------------------------------------------------------------------

# Cosine-sum coefficient tables, draco/util/tools.py:584-595 (digit for digit).
_COSINE_SUM = {
    "uniform": (1.0, 0.0, 0.0, 0.0),
    "hann": (0.5, -0.5, 0.0, 0.0),
    "hanning": (0.5, -0.5, 0.0, 0.0),
    "hamming": (0.53836, -0.46164, 0.0, 0.0),
    "blackman": (0.42, -0.5, 0.08, 0.0),
    "nuttall": (0.355768, -0.487396, 0.144232, -0.012604),
    "blackman_nuttall": (0.3635819, -0.4891775, 0.1365995, -0.0106411),
    "blackman_harris": (0.35875, -0.48829, 0.14128, -0.01168),
}

WINDOW_NAMES = tuple(_COSINE_SUM)


def invert_no_zero(x):
    """1/x where x != 0, else 0 (caput.algorithms.invert_no_zero, re-exported at
    draco/util/tools.py:12)."""
    x = np.asarray(x)
    out = np.zeros_like(x, dtype=np.result_type(x.dtype, np.float64))
    nz = x != 0
    out[nz] = 1.0 / x[nz]
    return out


def cosine_sum_window(x, window="nuttall"):
    """Four-term cosine-sum window at positions ``x``; zero outside [0, 1].

    Restates draco/util/tools.py:547-601 for the enum members the delay tasks accept
    (draco/analysis/delay.py:408-420), plus 'triangular' and 'tukey-0.X'.
    """
    x = np.asarray(x, dtype=np.float64)
    if window == "triangular":
        w = 1.0 - 2.0 * np.abs(x - 0.5)
    elif window.startswith("tukey"):
        alpha = 0.5 * float(window.split("-")[1])
        w = np.ones_like(x)
        lo = x < alpha
        hi = x >= (1.0 - alpha)
        w[lo] = 0.5 * (1.0 + np.cos(np.pi * (x[lo] - alpha) / alpha))
        w[hi] = 0.5 * (1.0 + np.cos(np.pi * (x[hi] - (1.0 - alpha)) / alpha))
    else:
        coeff = np.array(_COSINE_SUM[window])
        phase = 2 * np.pi * np.arange(4)[:, None] * x[None, :]
        w = (coeff[:, None] * np.cos(phase)).sum(axis=0)
    return np.where((x >= 0) & (x <= 1), w, 0)


# ---------------------------------------------------------------------------
# Fourier matrices (draco/analysis/delay.py:1480-1613)
# ---------------------------------------------------------------------------


def _phase(N, fsel, default_count, freq_major=True):
    chan = np.arange(default_count) if fsel is None else np.array(fsel)
    t = np.arange(N)
    if freq_major:
        return 2 * np.pi * t[None, :] * chan[:, None] / N, chan
    return 2 * np.pi * t[:, None] * chan[None, :] / N, chan


def dft_r2c(N, fsel=None):
    """(2M, N) real matrix: row 2i = cos, row 2i+1 = -sin (delay.py:1480-1510)."""
    arg, chan = _phase(N, fsel, N // 2 + 1)
    out = np.zeros((2 * chan.shape[0], N))
    out[0::2] = np.cos(arg)
    out[1::2] = -np.sin(arg)
    return out


def dft_c2r(N, fsel=None):
    """(N, 2M) inverse matrix with the 1/N (x2 off DC/Nyquist) factor (delay.py:1513-1546)."""
    arg, chan = _phase(N, fsel, N // 2 + 1, freq_major=False)
    mul = np.where((chan == 0) | (chan == N // 2), 1.0, 2.0)[None, :] / N
    out = np.zeros((N, 2 * chan.shape[0]))
    out[:, 0::2] = np.cos(arg) * mul
    out[:, 1::2] = -np.sin(arg) * mul
    return out


def dft_c2c(N, fsel=None):
    """(2M, 2N) real 2x2-block form of the complex DFT, numpy.fft sign (delay.py:1549-1585)."""
    arg, chan = _phase(N, fsel, N)
    out = np.zeros((2 * chan.shape[0], 2 * N))
    c, s = np.cos(arg), np.sin(arg)
    out[0::2, 0::2] = c
    out[0::2, 1::2] = s
    out[1::2, 0::2] = -s
    out[1::2, 1::2] = c
    return out


def dft_complex(N, fsel=None):
    """(M, N) complex matrix exp(-2 pi i t f / N) (delay.py:1588-1613)."""
    chan = np.arange(N) if fsel is None else np.array(fsel)
    return np.exp(-2.0j * np.pi * np.arange(N)[None, :] * chan[:, None] / N)


def _as_alternating_real(a):
    """complex128 -> interleaved float64 view (delay.py:1616-1631)."""
    return np.ascontiguousarray(a, dtype=np.complex128).view(np.float64)


def _as_complex(a):
    """interleaved float64 -> complex128 view (delay.py:1634-1649)."""
    return np.ascontiguousarray(a, dtype=np.float64).view(np.complex128)


# ---------------------------------------------------------------------------
# Random-draw sources.  The reference takes a numpy Generator; deterministic parity
# is defined by INJECTED draws in the reference's call order:
# per iteration  standard_normal((n_sys, T)) -> standard_normal((2M, T)) ->
# chisquare(T, size=N).
# ---------------------------------------------------------------------------


class RecordingDraws:
    """Wraps a numpy Generator and keeps a copy of every array it hands out."""

    def __init__(self, rng):
        self._rng = rng
        self.normals = []
        self.chisq = []

    def standard_normal(self, shape):
        a = self._rng.standard_normal(shape)
        self.normals.append(a.copy())
        return a

    def chisquare(self, df, size=None):
        a = self._rng.chisquare(df, size=size)
        self.chisq.append(np.array(a, copy=True))
        return a

    def packed(self):
        """(w1[niter,n_sys,T], w2[niter,2M,T], chi2[niter,N]) of the recorded chain."""
        w1 = np.stack(self.normals[0::2]) if self.normals else np.zeros((0, 0, 0))
        w2 = np.stack(self.normals[1::2]) if self.normals else np.zeros((0, 0, 0))
        c2 = np.stack(self.chisq) if self.chisq else np.zeros((0, 0))
        return w1, w2, c2


class InjectedDraws:
    """Replays pre-generated draws: w1[i], w2[i], chi2[i] for iteration i."""

    def __init__(self, w1, w2, chi2):
        self._w1, self._w2, self._c2 = w1, w2, chi2
        self._in = 0
        self._ic = 0

    def standard_normal(self, shape):
        src = self._w1 if self._in % 2 == 0 else self._w2
        a = np.array(src[self._in // 2], dtype=np.float64, copy=True)
        self._in += 1
        if a.shape != tuple(shape):
            raise ValueError(f"injected normal draw has shape {a.shape}, wanted {shape}")
        return a

    def chisquare(self, df, size=None):
        a = np.array(self._c2[self._ic], dtype=np.float64, copy=True)
        self._ic += 1
        return a


# ---------------------------------------------------------------------------
# Setup: window, whiten, projection matrices  (delay.py:1657-1710)
# ---------------------------------------------------------------------------


def whiten_inputs(data, N, Ni, fsel, window, complex_timedomain):
    """Return ``(z, FTNih, FTNiF)`` for one baseline.

    ``data`` is (T, M) complex; ``z`` is the windowed, noise-whitened data as (2M, T)
    interleaved reals; ``FTNih = F^T sqrt(Ni_r)`` is (n_sys, 2M) and ``FTNiF`` is its
    Gram matrix.  DC and Nyquist channels carry no imaginary part in real mode
    (delay.py:1692-1699).
    """
    nchan_total = N if complex_timedomain else N // 2 + 1
    if fsel is None:
        fsel = np.arange(nchan_total)
    fsel = np.asarray(fsel)

    F = dft_c2c(N, fsel) if complex_timedomain else dft_r2c(N, fsel)
    z = _as_alternating_real(data).T.copy()

    if window is not None:
        w = np.repeat(cosine_sum_window(fsel / nchan_total, window=window), 2)
        F *= w[:, None]
        z *= w[:, None]

    if complex_timedomain:
        purely_real = np.zeros(fsel.shape, dtype=bool)
    else:
        purely_real = (fsel == 0) | (fsel == N // 2)

    Ni = np.asarray(Ni)
    Ni_r = np.zeros(2 * Ni.shape[0])
    Ni_r[0::2] = np.where(purely_real, Ni, Ni * 2)
    Ni_r[1::2] = np.where(purely_real, 0.0, Ni * 2)

    FTNih = F.T * Ni_r[None, :] ** 0.5
    FTNiF = FTNih @ FTNih.T
    z *= Ni_r[:, None] ** 0.5
    return z, FTNih, FTNiF


# ---------------------------------------------------------------------------
# The sampler  (delay.py:1713-1904)
# ---------------------------------------------------------------------------


def _signal_draw_delay_space(S, z, FTNih, FTNiF, N, rng, complex_timedomain):
    """Perturbed Wiener draw through the n_sys x n_sys system (delay.py:1778-1819).

    The factorisation precedes the draws, so a non-PD failure consumes no randoms.
    """
    Si = invert_no_zero(S)
    if complex_timedomain:
        Si = 2.0 * np.repeat(Si, 2)
    Ci = FTNiF.copy()
    Ci[np.diag_indices_from(Ci)] += Si
    fac = la.cho_factor(Ci, check_finite=False, lower=False, overwrite_a=True)

    T = z.shape[1]
    w1 = rng.standard_normal(((2 * N if complex_timedomain else N), T))
    w2 = rng.standard_normal(z.shape)
    rhs = w1 * (Si**0.5)[:, None] + FTNih @ (z + w2)
    return la.cho_solve(fac, rhs, check_finite=False, overwrite_b=True)


def _signal_draw_freq_space(S, z, FTNih, N, rng, complex_timedomain):
    """Same posterior through the 2M x 2M system (delay.py:1821-1860).

    Here the draws precede the factorisation.
    """
    Sh = S**0.5
    if complex_timedomain:
        Sh = (0.5**0.5) * np.repeat(Sh, 2)
    T = z.shape[1]
    w1 = rng.standard_normal(((2 * N if complex_timedomain else N), T))
    w2 = rng.standard_normal(z.shape)

    Rt = FTNih * Sh[:, None]
    R = Rt.T
    rhs = w2 - R @ w1
    rhs += z
    Ci = R @ Rt
    Ci[np.diag_indices_from(Ci)] += 1.0
    fac = la.cho_factor(Ci, check_finite=False, lower=False, overwrite_a=True)
    x = la.cho_solve(fac, rhs, check_finite=False, overwrite_b=True)
    return Sh[:, None] * ((Rt @ x) + w1)


def _spectrum_draw(d, rng, complex_timedomain):
    """Inverse-chi^2 draw given a signal sample (delay.py:1862-1880): variance over the
    sample axis (ddof 0, mean removed), df = T in both modes."""
    S_hat = d.var(axis=-1)
    if complex_timedomain:
        S_hat = S_hat[::2] + S_hat[1::2]
    df = d.shape[1]
    chi2 = rng.chisquare(df, size=S_hat.shape[0])
    return S_hat * df / chi2


def gibbs_chain(
    data,
    N,
    Ni,
    initial_S,
    window="nuttall",
    fsel=None,
    niter=20,
    rng=None,
    complex_timedomain=False,
):
    """Restatement of ``delay_power_spectrum_gibbs`` (delay.py:1713-1904).

    Returns ``(samples, success)``: a list of up to ``niter`` (N,) spectra in unshifted
    delay order; ``success`` is False when a factorisation hit a non-positive pivot,
    in which case the list is truncated at that iteration (delay.py:1891-1896).
    """
    if rng is None:
        rng = np.random.default_rng()
    if fsel is None:
        fsel = np.arange(N if complex_timedomain else N // 2 + 1)
    z, FTNih, FTNiF = whiten_inputs(data, N, Ni, fsel, window, complex_timedomain)

    use_delay_space = len(fsel) > 0.25 * N  # delay.py:1884-1886
    S = initial_S
    samples = []
    success = True
    for _ in range(niter):
        try:
            if use_delay_space:
                d = _signal_draw_delay_space(S, z, FTNih, FTNiF, N, rng, complex_timedomain)
            else:
                d = _signal_draw_freq_space(S, z, FTNih, N, rng, complex_timedomain)
        except np.linalg.LinAlgError:
            success = False
            break
        S = _spectrum_draw(d, rng, complex_timedomain)
        samples.append(S)
    return samples, success


def wiener_filter(delay_PS, data, N, Ni, window="nuttall", fsel=None, complex_timedomain=False):
    """Restatement of ``delay_spectrum_wiener_filter`` (delay.py:2132-2201): the mean of
    the signal draw, returned as (T, N) (complex in complex_timedomain mode)."""
    z, FTNih, FTNiF = whiten_inputs(data, N, Ni, fsel, window, complex_timedomain)
    rhs = FTNih @ z
    Si = invert_no_zero(delay_PS)
    if complex_timedomain:
        Si = 2.0 * np.repeat(Si, 2)
    FTNiF[np.diag_indices_from(FTNiF)] += Si
    fac = la.cho_factor(FTNiF, check_finite=False, lower=False)
    out = la.cho_solve(fac, rhs, check_finite=False).T
    return _as_complex(out) if complex_timedomain else out


# ---------------------------------------------------------------------------
# Host-side pieces of the task around the sampler
# ---------------------------------------------------------------------------


def calculate_delays(
    freq,
    freq_zero=None,
    freq_spacing=None,
    nfreq=None,
    skip_nyquist=True,
    complex_timedomain=False,
):
    """Delay grid (fftshifted) and channel indices (delay.py:461-512)."""
    freq = np.asarray(freq)
    f0 = freq[0] if freq_zero is None else freq_zero
    df = np.abs(np.diff(freq)).min() if freq_spacing is None else freq_spacing
    if complex_timedomain:
        nfreq = len(freq)
        channel_ind = np.arange(nfreq)
        ndelay = nfreq
    else:
        channel_ind = (np.abs(freq - f0) / df).astype(np.int64)
        if nfreq is None:
            nfreq = channel_ind[-1] + 1
            if skip_nyquist:
                nfreq += 1
        ndelay = 2 * (nfreq - 1)
    delays = np.fft.fftshift(np.fft.fftfreq(ndelay, d=df))
    return delays, channel_ind


def cut_data(
    data,
    weight,
    time_frac=0.0,
    freq_frac=0.0,
    remove_mean=True,
    scale_freq=False,
    use_average_weights=True,
    weight_boost=1.0,
):
    """Time/frequency cuts for one baseline (delay.py:516-597).

    ``data``/``weight`` are (T0, nfreq).  Returns ``None`` for a dead baseline (three
    exits: no positive weight, no surviving channel, all-zero data), else
    ``(data[T, M], weight[M], keep_freq[nfreq], keep_time[T0])``.
    """
    ntime, nfreq = data.shape[-2:]
    positive = weight > 0
    if not positive.any():
        return None
    keep_time = positive.mean(axis=-1).reshape(-1, ntime).mean(axis=0) > time_frac
    positive = positive[..., keep_time, :]
    keep_freq = positive.mean(axis=-2).reshape(-1, nfreq).mean(axis=0) > freq_frac
    if not keep_freq.any():
        return None
    data = data[..., keep_time, :][..., keep_freq]
    weight = weight[..., keep_time, :][..., keep_freq]
    if remove_mean:
        data = data - data.mean(axis=0, keepdims=True)
    if (data == 0.0).all():
        return None
    if scale_freq:
        rel = data.std(axis=-2)[..., None, :] / data.std(axis=(-1, -2))[..., None, None]
        data = data * invert_no_zero(rel)
    if use_average_weights:
        weight = np.mean(weight, axis=-2)
    weight = weight * weight_boost
    return data, weight, keep_freq, keep_time


def gibbs_estimate(
    data,
    weight,
    S0,
    ndelay,
    channel_ind,
    nsamp,
    rng,
    window="nuttall",
    median_frac=0.5,
    complex_timedomain=False,
):
    """``DelayPowerSpectrumGibbs._estimator`` (delay.py:1242-1267): run the chain, take
    the median of the trailing ``int(nsamp*median_frac)`` samples (``[-0:]`` is the
    whole chain), fftshift.  Empty chain -> the prior and ``success=False``."""
    samples, success = gibbs_chain(
        data,
        ndelay,
        weight,
        S0,
        window=window,
        fsel=channel_ind,
        niter=nsamp,
        rng=rng,
        complex_timedomain=complex_timedomain,
    )
    ntail = int(nsamp * median_frac)
    if samples:
        spec = np.fft.fftshift(np.median(samples[-ntail:], axis=0))
    else:
        success = False
        spec = S0
    return spec, samples, success
