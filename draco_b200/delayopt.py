"""NRML / maximum-posterior delay power spectrum estimator with the likelihood on the GPU.

Host-side mirror of ``draco/analysis/delayopt.py``: same class and function names, same arguments, same error
behaviour.  ``LogLikePS`` (delayopt.py:61-236) evaluates value / gradient / Hessian on the device through
``libdgnrml.so`` (``include/dgnrml.h``); ``GaussianProcessPrior`` (delayopt.py:239-355) is an O(N^2) host
matrix-vector product as in the reference; ``delay_power_spectrum_maxpost`` (delayopt.py:424-561) drives
``scipy.optimize.minimize(method="Newton-CG")`` exactly as the reference does -- scipy is the reference's own
dependency, the optimiser's control flow is not part of the rebuilt path.

There is no CPU fallback: without the CUDA library (or a GPU) constructing ``LogLikePS`` raises.
"""

from __future__ import annotations

import ctypes as C
import os

import numpy as np
import scipy.linalg as la
from scipy.optimize import minimize
from scipy.spatial.distance import cdist

from . import _lib as _glib

HERE = os.path.dirname(os.path.abspath(__file__))
NRML_LIB_PATH = os.path.join(HERE, "libdgnrml.so")
NRML_EXPORTED = ["dgnrml_create", "dgnrml_eval", "dgnrml_initial_guess", "dgnrml_destroy", "dgnrml_last_error",
                 "dgnrml_last_timing"]
VALUE, GRADIENT, HESSIAN = 1, 2, 4

_nlib = None
_vp = C.c_void_p


def load_nrml():
    """``libdgnrml.so`` (built by ``draco_b200._build``); raises when it is missing."""
    global _nlib
    if _nlib is not None:
        return _nlib
    if not os.path.exists(NRML_LIB_PATH):
        raise RuntimeError(f"{NRML_LIB_PATH} is missing: build it with `python -m draco_b200._build`.  There is no CPU fallback.")
    lib = C.CDLL(NRML_LIB_PATH)
    lib.dgnrml_create.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_double, C.c_double, _vp, _vp,
                                  _vp, C.POINTER(_vp)]
    lib.dgnrml_eval.argtypes = [_vp, _vp, C.c_int32, C.POINTER(C.c_double), _vp, _vp]
    lib.dgnrml_destroy.argtypes = [_vp]
    lib.dgnrml_last_error.argtypes = [_vp]
    lib.dgnrml_last_error.restype = C.c_char_p
    lib.dgnrml_last_timing.argtypes = [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.dgnrml_initial_guess.argtypes = [_vp, _vp, C.c_int32, C.c_double, _vp]
    for n in ("dgnrml_create", "dgnrml_eval", "dgnrml_initial_guess", "dgnrml_destroy", "dgnrml_last_timing"):
        getattr(lib, n).restype = C.c_int
    _nlib = lib
    return lib


# cosine-sum window coefficients, draco/util/tools.py:584-595 (the table of csrc/setup_kernels.cu, host side)
_WINDOW_COEFF = {
    "uniform": (1.0, 0.0, 0.0, 0.0),
    "hann": (0.5, -0.5, 0.0, 0.0),
    "hanning": (0.5, -0.5, 0.0, 0.0),
    "hamming": (0.53836, -0.46164, 0.0, 0.0),
    "blackman": (0.42, -0.5, 0.08, 0.0),
    "nuttall": (0.355768, -0.487396, 0.144232, -0.012604),
    "blackman_nuttall": (0.3635819, -0.4891775, 0.1365995, -0.0106411),
    "blackman_harris": (0.35875, -0.48829, 0.14128, -0.01168),
}


def window_generalised(x, window="nuttall"):
    """Cosine-sum apodisation at positions ``x`` in [0, 1], zero outside (draco/util/tools.py:547-601)."""
    x = np.asarray(x, dtype=np.float64)
    if window not in _WINDOW_COEFF:
        raise ValueError(f"Unknown window function: {window}")
    t = 2 * np.pi * np.arange(4)[:, np.newaxis] * x[np.newaxis, :]
    w = (np.array(_WINDOW_COEFF[window])[:, np.newaxis] * np.cos(t)).sum(axis=0)
    return np.where((x >= 0) & (x <= 1), w, 0)


def _invert_no_zero(x):
    x = np.asarray(x, dtype=np.float64)
    out = np.zeros_like(x)
    np.divide(1.0, x, out=out, where=x != 0)
    return out


class LogLikePS:
    """Negative log-likelihood of the log delay power spectrum, with gradient and Hessian (delayopt.py:61-236).

    Same constructor as the reference class.  The arrays are selected with ``fsel`` on the host (delayopt.py:103-110)
    and uploaded once; every evaluation runs on ``cuda:device``.  A covariance that is not positive definite raises
    ``scipy.linalg.LinAlgError`` as ``la.cho_factor`` does in the reference (delayopt.py:149).
    """

    def __init__(self, X, MF, N, nsamp, fsel=None, exact_hessian=True, bounds=(1e-10, 1e10), *, device=0):
        if fsel is None:
            fsel = (MF != 0).any(axis=1)
        self._rows = fsel
        self.X = np.ascontiguousarray(X[fsel][:, fsel], dtype=np.complex128)
        self.N = np.ascontiguousarray(N[fsel], dtype=np.float64)
        self.MF = np.ascontiguousarray(MF[fsel], dtype=np.complex128)
        self.nsamp = int(nsamp)
        self.exact_hessian = bool(exact_hessian)
        self._logbounds = tuple(sorted(np.log(x) for x in bounds))
        self._lib = load_nrml()
        self._h = _vp()
        self.eval_ms = 0.0
        self.eval_flop = 0.0
        self.evals = 0
        M, nd = self.MF.shape
        rc = self._lib.dgnrml_create(int(device), M, nd, self.nsamp, int(self.exact_hessian), float(self._logbounds[0]),
                                     float(self._logbounds[1]), self.X.ctypes.data, self.MF.ctypes.data,
                                     self.N.ctypes.data, C.byref(self._h))
        if rc != 0:
            msg = self._lib.dgnrml_last_error(self._h).decode() if self._h else "no CUDA device"
            self.close()
            raise _glib.DGibbsError(f"dgnrml_create failed ({rc}): {msg}")
        self._ndelay = nd

    def _eval(self, x, what):
        x = np.ascontiguousarray(x, dtype=np.float64)
        if x.shape != (self._ndelay,):
            raise ValueError(f"expected {self._ndelay} parameters, got {x.shape}")
        val = C.c_double(0.0)
        g = np.empty(self._ndelay) if what & GRADIENT else None
        H = np.empty((self._ndelay, self._ndelay)) if what & HESSIAN else None
        rc = self._lib.dgnrml_eval(self._h, x.ctypes.data, what, C.byref(val), g.ctypes.data if g is not None else None,
                                   H.ctypes.data if H is not None else None)
        if rc == 1:
            raise la.LinAlgError(self._lib.dgnrml_last_error(self._h).decode())
        if rc != 0:
            raise _glib.DGibbsError(f"dgnrml_eval failed ({rc}): {self._lib.dgnrml_last_error(self._h).decode()}")
        ms, fl = C.c_double(0.0), C.c_double(0.0)
        self._lib.dgnrml_last_timing(self._h, C.byref(ms), C.byref(fl))
        self.eval_ms += ms.value
        self.eval_flop += fl.value
        self.evals += 1
        return val.value, g, H

    def initial_guess(self, data, rtol=1e-3):
        """``(data @ la.pinv(MF.T, rtol=rtol)).var(axis=0)`` (delayopt.py:505-506) on the device; ``data`` [T, M] are
        the windowed samples of the channels this object keeps."""
        data = np.ascontiguousarray(data, dtype=np.complex128)
        if data.ndim != 2 or data.shape[1] != self.MF.shape[0]:
            raise ValueError(f"data must be [T, {self.MF.shape[0]}], got {data.shape}")
        S0 = np.empty(self._ndelay)
        rc = self._lib.dgnrml_initial_guess(self._h, data.ctypes.data, data.shape[0], float(rtol), S0.ctypes.data)
        if rc == 1:
            raise la.LinAlgError(self._lib.dgnrml_last_error(self._h).decode())
        if rc != 0:
            raise _glib.DGibbsError(f"dgnrml_initial_guess failed ({rc}): {self._lib.dgnrml_last_error(self._h).decode()}")
        return S0

    def value(self, x):
        return self._eval(x, VALUE)[0]

    def gradient(self, x):
        return self._eval(x, GRADIENT)[1]

    def hessian(self, x):
        return self._eval(x, HESSIAN)[2]

    def close(self):
        if getattr(self, "_h", None) and self._h:
            self._lib.dgnrml_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def matern_kernel(N, width, alpha, nu):
    """draco/util/kernels.py:131-184 for a square kernel over ``arange(N)``."""
    if nu not in {1.5, 2.5}:
        raise ValueError(f"Invalid value `nu`={nu}. Only values of (1.5, 2.5) are currently supported.")
    i = np.arange(int(N)) / float(width)
    dist = cdist(i[:, np.newaxis], i[:, np.newaxis], metric="euclidean")
    if nu == 1.5:
        dist *= np.sqrt(3)
        Cm = 1.0 + dist
    else:
        dist *= np.sqrt(5)
        Cm = 1.0 + dist + dist**2 / 3.0
    Cm *= np.exp(-dist)
    Cm *= alpha**2
    return Cm


_prior_cache: dict = {}
_fourier_cache: dict = {}


class GaussianProcessPrior:
    """Gaussian-process prior on the parameters (delayopt.py:239-355); the Matern kernel the estimator uses.
    The inverse covariance depends only on the constructor arguments, so it is kept between baselines (the reference
    rebuilds and inverts the N x N kernel for every baseline, delayopt.py:518-523)."""

    def __init__(self, N, *, width=5, alpha=1, kernel="matern", reg=1e-8, nu=1.5):
        if kernel != "matern":
            raise ValueError("only the 'matern' kernel of delay_power_spectrum_maxpost is implemented")
        key = (int(N), int(width), float(alpha), float(reg), float(nu))
        if key not in _prior_cache:
            _prior_cache.clear()
            Cm = matern_kernel(int(N), int(width), 1.0, nu)
            Ci = la.inv(Cm + np.identity(N) * reg) / alpha**2
            Ci.setflags(write=False)
            _prior_cache[key] = Ci
        self.Ci = _prior_cache[key]

    def value(self, x):
        return 0.5 * float(x @ (self.Ci @ x))

    def gradient(self, x):
        return self.Ci @ x

    def hessian(self, x):
        return self.Ci


class AddFunctions:
    """Sum of objectives (delayopt.py:358-421)."""

    def __init__(self, functions):
        if len(functions) <= 0:
            raise ValueError("At least one function must be supplied.")
        self.functions = functions

    def value(self, x):
        return sum(f.value(x) for f in self.functions)

    def gradient(self, x):
        g = self.functions[0].gradient(x)
        for f in self.functions[1:]:
            g = g + f.gradient(x)
        return g

    def hessian(self, x):
        h = self.functions[0].hessian(x)
        for f in self.functions[1:]:
            h = h + f.hessian(x)
        return h


def delay_power_spectrum_maxpost(data, N, Ni, initial_S=None, window="nuttall", fsel=None, maxiter=100, tol=1e-3,
                                 bounds=(1e-15, 1e10), *, device=0, stats=None, host_pinv=False):
    """Maximum-posterior delay power spectrum (delayopt.py:424-561): ``(samples, success)``.

    ``stats`` (optional dict) receives the device time, flop and number of likelihood evaluations.  The
    pseudo-inverse starting point (delayopt.py:505-506) is computed on the device from a Hermitian
    eigendecomposition unless ``host_pinv`` asks for scipy's SVD-based ``la.pinv`` as in the reference."""
    nsamp, Nf = data.shape
    if fsel is None:
        fsel = np.arange(Nf)
    elif len(fsel) != Nf:
        raise ValueError(f"Length of frequency selection must match frequencies passed. {len(fsel)} != {data.shape[-1]}")
    fsel = np.asarray(fsel)
    fkey = (int(N), fsel.tobytes())
    if fkey not in _fourier_cache:  # consecutive baselines usually share their channel selection
        _fourier_cache.clear()
        _fourier_cache[fkey] = np.exp(-2.0j * np.pi * np.arange(N)[np.newaxis, :] * fsel[:, np.newaxis] / N)
    F = _fourier_cache[fkey].copy()  # fourier_matrix, delay.py:1588-1613
    data = np.array(data, dtype=np.complex128, copy=True)
    if window is not None:
        w = window_generalised(fsel / N, window=window)
        F *= w[:, np.newaxis]
        data *= w[np.newaxis, :]
    X = (data.T @ data.conj()) / nsamp
    Nm = _invert_no_zero(Ni)
    F[np.asarray(Ni) == 0] = 0.0
    if initial_S is None and host_pinv:
        initial_S = (data @ la.pinv(F.T, rtol=1e-3)).var(axis=0)
    like = LogLikePS(X, F, Nm, nsamp, exact_hessian=True, bounds=bounds, device=device)
    if initial_S is None:
        initial_S = like.initial_guess(data[:, like._rows], rtol=1e-3)  # masked channels have zero rows in F
    samples = [initial_S]
    optfunc = AddFunctions([like, GaussianProcessPrior(N, width=5, alpha=1.0, kernel="matern", nu=1.5)])
    try:
        res = minimize(optfunc.value, x0=np.log(initial_S), jac=optfunc.gradient, hess=optfunc.hessian, method="Newton-CG",
                       options={"maxiter": maxiter, "xtol": tol}, callback=lambda xk: samples.append(np.exp(xk)))
        success = bool(res.success)
    except (la.LinAlgError, ValueError):
        success = False
    finally:
        if stats is not None:
            stats.update(ms=like.eval_ms, flop=like.eval_flop, evals=like.evals)
        like.close()
    return samples, success
