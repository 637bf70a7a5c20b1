"""numpy/scipy restatement of draco's NRML / maximum-posterior delay power spectrum estimator (CPU oracle).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Restates ``delay_power_spectrum_maxpost``
(``draco/analysis/delayopt.py:424-561``), its objective ``LogLikePS`` (``delayopt.py:61-236``), the Gaussian-process
prior (``delayopt.py:239-355``) with the Matern-3/2 kernel (``draco/util/kernels.py:131-184, 316-351``) and the
sum of objectives (``delayopt.py:358-421``).  The optimiser itself is ``scipy.optimize.minimize(method="Newton-CG")``
exactly as in the reference: scipy is the reference's dependency, not part of the path being rebuilt.
Pinned against outputs of the reference classes themselves (``tests/golden/make_golden_nrml.py`` executes them
through ``oracle/ref_loader.py``).
"""

from __future__ import annotations

import numpy as np
import scipy.linalg as la
from scipy.optimize import minimize

from .gibbs_oracle import cosine_sum_window, invert_no_zero


def fourier_c2c(N, fsel=None):
    """delay.py:1588-1613: ``exp(-2 pi i t f / N)``, rows = channels, columns = delays."""
    f = np.arange(N) if fsel is None else np.asarray(fsel)
    return np.exp(-2.0j * np.pi * np.arange(N)[None, :] * f[:, None] / N)


def matern32_inverse(N, width=5, alpha=1.0, reg=1e-8):
    """``GaussianProcessPrior.__init__`` with ``kernel="matern", nu=1.5`` (delayopt.py:265-287,
    kernels.py:131-184): inverse of the regularised covariance of the points 0..N-1."""
    i = np.arange(N) / float(width)
    r = np.abs(i[:, None] - i[None, :]) * np.sqrt(3)
    C = (1.0 + r) * np.exp(-r)
    return la.inv(C + np.identity(N) * reg) / alpha**2


class LogLike:
    """Negative log-likelihood of the log power spectrum, its gradient and Hessian (delayopt.py:61-236)."""

    def __init__(self, X, MF, Nvar, nsamp, fsel=None, exact_hessian=True, bounds=(1e-10, 1e10)):
        if fsel is None:
            fsel = (MF != 0).any(axis=1)
        self.X = X[fsel][:, fsel]
        self.N = Nvar[fsel]
        self.MF = MF[fsel]
        self.nsamp = nsamp
        self.exact = exact_hessian
        self.lo, self.hi = sorted(np.log(b) for b in bounds)
        self._at = None

    def _prepare(self, x):
        if self._at is not None and np.array_equal(x, self._at_in):
            return
        self._at_in = np.array(x, copy=True)
        self._at = np.clip(x, self.lo, self.hi)
        S = np.exp(self._at)
        C = (self.MF * S[None, :]) @ self.MF.T.conj()
        C[np.diag_indices_from(C)] += self.N
        self.ch = la.cho_factor(C, check_finite=False)
        self.XC = self.X - C
        self.U = np.sqrt(S)[None, :] * self.MF
        self.Ut = la.cho_solve(self.ch, self.U, check_finite=False)
        self.XCUt = self.XC @ self.Ut

    def value(self, x):
        self._prepare(x)
        lndet = 2 * np.log(np.diag(self.ch[0])).real.sum()
        lndet += np.trace(la.cho_solve(self.ch, self.X, check_finite=False)).real
        return self.nsamp * lndet

    def gradient(self, x):
        self._prepare(x)
        return -self.nsamp * (self.Ut.conj() * self.XCUt).real.sum(axis=0)

    def hessian(self, x):
        self._prepare(x)
        G = self.U.T.conj() @ self.Ut
        H = (G * G.T.conj()).real
        if self.exact:
            K = self.Ut.T.conj() @ self.XCUt
            H += (2 * K * G.T).real
            H[np.diag_indices_from(H)] += -(self.Ut.conj() * self.XCUt).real.sum(axis=0)
        return self.nsamp * H


class GPPrior:
    """0.5 x^T Ci x (delayopt.py:239-355)."""

    def __init__(self, N, width=5, alpha=1.0, reg=1e-8):
        self.Ci = matern32_inverse(N, width, alpha, reg)

    def value(self, x):
        return 0.5 * float(x @ (self.Ci @ x))

    def gradient(self, x):
        return self.Ci @ x

    def hessian(self, x):
        return self.Ci


def maxpost_inputs(data, N, Ni, window="nuttall", fsel=None):
    """delayopt.py:474-503: windowed data covariance ``X``, masked windowed Fourier matrix ``F``, noise variances
    and the pseudo-inverse starting point."""
    nsamp, Nf = data.shape
    fsel = np.arange(Nf) if fsel is None else np.asarray(fsel)
    if len(fsel) != Nf:
        raise ValueError(f"Length of frequency selection must match frequencies passed. {len(fsel)} != {Nf}")
    F = fourier_c2c(N, fsel)
    data = np.array(data, dtype=np.complex128, copy=True)
    if window is not None:
        w = cosine_sum_window(fsel / N, window=window)
        F = F * w[:, None]
        data = data * w[None, :]
    X = (data.T @ data.conj()) / nsamp
    Nm = invert_no_zero(np.asarray(Ni, dtype=np.float64))
    F[np.asarray(Ni) == 0] = 0.0
    return data, X, F, Nm


def maxpost_chain(data, N, Ni, initial_S=None, window="nuttall", fsel=None, maxiter=100, tol=1e-3, bounds=(1e-15, 1e10),
                  loglike=None):
    """``delay_power_spectrum_maxpost`` (delayopt.py:424-561): list of spectrum iterates and the success flag.
    ``loglike(X, F, Nm, nsamp, bounds)`` may supply another implementation of the likelihood object (the GPU one)."""
    nsamp = data.shape[0]
    dataw, X, F, Nm = maxpost_inputs(data, N, Ni, window, fsel)
    if initial_S is None:
        initial_S = (dataw @ la.pinv(F.T, rtol=1e-3)).var(axis=0)
    samples = [initial_S]
    like = (loglike or (lambda *a: LogLike(a[0], a[1], a[2], a[3], exact_hessian=True, bounds=a[4])))(X, F, Nm, nsamp, bounds)
    prior = GPPrior(N, width=5, alpha=1.0)
    try:
        res = minimize(
            lambda x: like.value(x) + prior.value(x),
            x0=np.log(initial_S),
            jac=lambda x: like.gradient(x) + prior.gradient(x),
            hess=lambda x: like.hessian(x) + prior.hessian(x),
            method="Newton-CG",
            options={"maxiter": maxiter, "xtol": tol},
            callback=lambda xk: samples.append(np.exp(xk)),
        )
        success = bool(res.success)
    except (la.LinAlgError, ValueError):
        success = False
    return samples, success
