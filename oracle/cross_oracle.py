"""numpy/scipy restatement of draco's delay CROSS-power-spectrum Gibbs sampler (CPU oracle).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Restates
``delay_spectrum_gibbs_cross`` (``draco/analysis/delay.py:1907-2099``) and the complex
normal / complex Wishart helpers it takes from ``caput.algorithms.random`` -- external
and absent here; their in-tree twins are ``draco/util/random.py:7-166``, which is what
this file follows.  Pinned against outputs of the reference functions themselves
(``tests/golden/make_golden.py`` executes them through ``oracle/ref_loader.py``).

Draw order per Gibbs iteration (verified by recording the reference's rng calls):
    standard_normal((N, nd, 2T))                    -> w1, complex (N, nd, T) / sqrt(2)
    standard_normal((nd, M, 2T))                    -> w2, complex (nd, M, T) / sqrt(2)
    then for every delay t = 0..N-1 (complex Wishart, Bartlett factor):
        standard_normal(nd(nd-1)/2)  (real parts of the strict lower triangle)
        standard_normal(nd(nd-1)/2)  (imaginary parts)
        gamma(T - i) for i = 0..nd-1  (squared diagonal)
"""

from __future__ import annotations

import numpy as np
import scipy.linalg as la

from .gibbs_oracle import cosine_sum_window, dft_complex


class CrossRecordingDraws:
    """Wraps a numpy Generator; keeps every variate it hands out, in call order."""

    def __init__(self, rng):
        self._rng = rng
        self.normals = []
        self.gammas = []

    def standard_normal(self, size=None, dtype=np.float64, out=None):
        if out is not None:
            self._rng.standard_normal(size, dtype=dtype, out=out)
            a = out
        else:
            a = self._rng.standard_normal(size, dtype=dtype)
        self.normals.append(np.array(a, copy=True))
        return a

    def gamma(self, shape, scale=1.0, size=None):
        a = self._rng.gamma(shape, scale, size)
        self.gammas.append(float(a))
        return a

    def packed(self, niter, N, nd, M, T):
        """``(w1[niter,N,nd,T] c128, w2[niter,nd,M,T] c128, wn[niter,N,2,L] f64, wg[niter,N,nd] f64)``
        with ``L = nd(nd-1)/2``: the draws of a recorded chain of ``niter`` complete iterations."""
        L = nd * (nd - 1) // 2
        per = 2 + 2 * N
        w1 = np.empty((niter, N, nd, T), dtype=np.complex128)
        w2 = np.empty((niter, nd, M, T), dtype=np.complex128)
        wn = np.empty((niter, N, 2, L))
        for i in range(niter):
            blk = self.normals[i * per : (i + 1) * per]
            w1[i] = blk[0].view(np.complex128) / 2**0.5
            w2[i] = blk[1].view(np.complex128) / 2**0.5
            for t in range(N):
                wn[i, t, 0] = blk[2 + 2 * t]
                wn[i, t, 1] = blk[3 + 2 * t]
        wg = np.array(self.gammas[: niter * N * nd]).reshape(niter, N, nd)
        return w1, w2, wn, wg


class CrossInjectedDraws:
    """Replays packed draws (see :meth:`CrossRecordingDraws.packed`) through the Generator calls the
    sampler makes."""

    def __init__(self, w1, w2, wn, wg):
        self._w1, self._w2, self._wn, self._wg = w1, w2, wn, wg
        niter, N, nd, _ = w1.shape
        self._N, self._nd = N, nd
        self._seq = []
        for i in range(niter):
            self._seq.append(np.ascontiguousarray(w1[i] * 2**0.5).view(np.float64))
            self._seq.append(np.ascontiguousarray(w2[i] * 2**0.5).view(np.float64))
            for t in range(N):
                self._seq.append(wn[i, t, 0])
                self._seq.append(wn[i, t, 1])
        self._in = 0
        self._ig = 0

    def standard_normal(self, size=None, dtype=np.float64, out=None):
        a = np.array(self._seq[self._in], dtype=np.float64, copy=True)
        self._in += 1
        if out is not None:
            out[...] = a.reshape(out.shape)
            return out
        return a

    def gamma(self, shape, scale=1.0, size=None):
        g = self._wg.reshape(-1)[self._ig]
        self._ig += 1
        return g


def standard_complex_normal(shape, rng):
    """draco/util/random.py:7-100: real normals viewed as complex, scaled by 1/sqrt(2)."""
    out = np.ndarray(shape, dtype=np.complex128)
    rsize = (*shape[:-1], shape[-1] * 2)
    rng.standard_normal(rsize, dtype=np.float64, out=out.view(np.float64))
    out *= 1.0 / 2**0.5
    return out


def standard_complex_wishart(m, n, rng):
    """draco/util/random.py:103-133 (Bartlett decomposition)."""
    T = np.zeros((m, m), dtype=np.complex128)
    T[np.tril_indices(m, k=-1)] = (rng.standard_normal(m * (m - 1) // 2) + 1.0j * rng.standard_normal(m * (m - 1) // 2)) / 2**0.5
    for i in range(m):
        T[i, i] = rng.gamma(n - i) ** 0.5
    return np.dot(T, T.T.conj())


def complex_wishart(C, n, rng):
    """draco/util/random.py:136-166."""
    L = la.cholesky(C, lower=True)
    A = standard_complex_wishart(C.shape[0], n, rng)
    return np.dot(L, np.dot(A, L.T.conj()))


def cross_whiten_inputs(data, N, Ni, fsel, window):
    """delay.py:1967-2001: returns ``(z[nd,M,T], FTNih[nd,N,M], FTNiF[nd,N,nd,N])``."""
    nd, nsamp, Nf = data.shape
    fsel = np.arange(Nf) if fsel is None else np.asarray(fsel)
    F = dft_complex(N, fsel)
    z = np.array(data.transpose(0, 2, 1), dtype=np.complex128, copy=True)
    if window is not None:
        w = cosine_sum_window(fsel * 1.0 / N, window=window)  # NB x = fsel / N here (delay.py:1982)
        F = F * w[:, None]
        z *= w[:, None]
    FTNih = F.T[None, :, :] * Ni[:, None, :] ** 0.5
    FTNiF = np.zeros((nd, N, nd, N), dtype=np.complex128)
    for ii in range(nd):
        FTNiF[ii, :, ii] = FTNih[ii] @ FTNih[ii].T.conj()
    z *= Ni[:, :, None] ** 0.5
    return z, FTNih, FTNiF


def cross_gibbs_chain(data, N, Ni, initial_S, window="nuttall", fsel=None, niter=20, rng=None):
    """``delay_spectrum_gibbs_cross`` (delay.py:1907-2099).

    ``data`` [nd, T, M] complex, ``Ni`` [nd, M], ``initial_S`` [nd, nd, N] complex.  Returns the list of
    ``[nd, nd, N]`` complex cross-spectrum samples; raises ``RuntimeError`` when a factorisation fails
    (delay.py:2096-2097).
    """
    if rng is None:
        rng = np.random.default_rng()
    nd, nsamp, Nf = data.shape
    if nd == 0:
        raise ValueError("Need at least one set of data")
    if fsel is not None and len(fsel) != Nf:
        raise ValueError("Length of frequency selection must match frequencies passed. " f"{len(fsel)} != {Nf}")
    z, FTNih, FTNiF = cross_whiten_inputs(data, N, np.asarray(Ni), fsel, window)

    def draw_signal(S):  # delay.py:2006-2058
        Si = np.empty_like(S)
        Sh = np.empty((N, nd, nd), dtype=S.dtype)
        for ii in range(N):
            Si[:, :, ii] = la.inv(S[:, :, ii])
            Sh[ii] = la.cholesky(S[:, :, ii], lower=False)
        Ci = FTNiF.copy()
        for ii in range(nd):
            for jj in range(nd):
                Ci[ii, :, jj] += np.diag(Si[ii, jj])
        w1 = standard_complex_normal((N, nd, nsamp), rng)
        w2 = standard_complex_normal(z.shape, rng)
        y = FTNih @ (z + w2)
        for ii in range(N):
            y[:, ii] += la.solve_triangular(Sh[ii], w1[ii], lower=False, check_finite=False)
        cf = la.cho_factor(Ci.reshape(nd * N, nd * N), check_finite=False)
        return la.cho_solve(cf, y.reshape(nd * N, nsamp), check_finite=False).reshape(nd, N, nsamp)

    def draw_spectrum(d):  # delay.py:2065-2083
        S = np.empty((nd, nd, N), dtype=np.complex128)
        for ii in range(N):
            S[:, :, ii] = np.cov(d[:, ii], bias=True)
        for ii in range(N):
            Sinv = la.inv(S[:, :, ii])
            Si_samp = complex_wishart(Sinv, nsamp, rng) / nsamp
            S[:, :, ii] = la.inv(Si_samp)
        return S

    spec = []
    S_samp = np.asarray(initial_S, dtype=np.complex128)
    try:
        for _ in range(niter):
            d_samp = draw_signal(S_samp)
            S_samp = draw_spectrum(d_samp)
            spec.append(S_samp)
    except la.LinAlgError as e:
        raise RuntimeError("Exiting earlier as singular") from e
    return spec
