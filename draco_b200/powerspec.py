"""caput-task face of the Wiener delay-transform operator on the GPU.

Drop-in for ``draco.analysis.powerspec.ConstructWienerDelayTransform`` (powerspec.py:118-369): same class name,
same YAML parameters and defaults (powerspec.py:152-171), same input (a ``RingMap`` holding ``filter``,
``freq_cov``, ``weight``, ``dirty_beam_power``) and the same ``DelayTransformOperator`` output
(draco/core/containers.py:2185-2203: ``filter [pol, ra, el, delay, freq]`` complex64, attrs ``window``,
``window_lower_freq``, ``window_upper_freq``).  The window, the delay axis, the prior and the Fourier operator are
computed on the host exactly as the reference computes them; the loops over elevation and right ascension
(powerspec.py:245-319) -- per pixel two F x F x F complex products, a Cholesky factorisation, the solves and the
projection -- run on the device through ``dgwiener_construct`` (``include/dgwiener.h``, ``libdgnrml.so``), one call
per polarisation.  There is no CPU fallback.
"""

from __future__ import annotations

import ctypes as C

import numpy as np

from . import _compat, _lib as _glib
from .delayopt import load_nrml, window_generalised

if _compat.HAVE_CAPUT:  # pragma: no cover - caput is absent from this image
    from caput import config
    from caput.pipeline import tasklib

    _ContainerTask = tasklib.base.ContainerTask
    try:
        from draco.core import containers
    except Exception:
        containers = _compat
else:
    config = _compat.config
    _ContainerTask = _compat.ContainerTask
    containers = _compat

WIENER_EXPORTED = ["dgwiener_construct", "dgwiener_sizeof", "dgwiener_last_error"]
_vp = C.c_void_p


class WienerProblem(C.Structure):
    """``dgwiener_problem`` of include/dgwiener.h."""

    _fields_ = [
        ("device_id", C.c_int32), ("nfreq", C.c_int32), ("ndelay", C.c_int32), ("nra", C.c_int32), ("nel", C.c_int32),
        ("filter_is_complex", C.c_int32), ("cov_is_complex", C.c_int32), ("solver", C.c_int32),
        ("fourier", _vp), ("prior", _vp), ("window", _vp), ("filter", _vp), ("freq_cov", _vp), ("weight", _vp),
        ("beam_power", _vp), ("out", _vp),
        ("out_stride_ra", C.c_int64), ("out_stride_el", C.c_int64), ("workspace_bytes", C.c_int64),
        ("bad_el", C.c_int32), ("bad_ra", C.c_int32),
        ("ms_device", C.c_double), ("ms_compute", C.c_double), ("flop", C.c_double), ("launches", C.c_int64),
    ]


_wlib = None


def load_wiener():
    """The ``dgwiener_*`` entry points of ``libdgnrml.so``; raises when the library has not been built."""
    global _wlib
    if _wlib is None:
        lib = load_nrml()
        lib.dgwiener_construct.argtypes = [C.POINTER(WienerProblem)]
        lib.dgwiener_construct.restype = C.c_int
        lib.dgwiener_sizeof.restype = C.c_int
        lib.dgwiener_last_error.restype = C.c_char_p
        if lib.dgwiener_sizeof() != C.sizeof(WienerProblem):
            raise RuntimeError(f"ABI mismatch: dgwiener_problem is {C.sizeof(WienerProblem)} bytes here, "
                               f"{lib.dgwiener_sizeof()} in libdgnrml.so; rebuild the library")
        _wlib = lib
    return _wlib


def _window(x, name):
    """tools.window_generalised, draco/util/tools.py:547-601 (the enum of powerspec.py:155-169)."""
    x = np.asarray(x, dtype=np.float64)
    if name.startswith("tukey"):
        alpha = 0.5 * float(name.split("-")[1])
        w = np.ones_like(x)
        lo, hi = x < alpha, x >= (1.0 - alpha)
        w[lo] = 0.5 * (1.0 + np.cos(np.pi * (x[lo] - alpha) / alpha))
        w[hi] = 0.5 * (1.0 + np.cos(np.pi * (x[hi] - (1.0 - alpha)) / alpha))
        return np.where((x >= 0) & (x <= 1), w, 0)
    return window_generalised(x, window=name)


def wiener_delay_operator(fourier, prior, window, filt, freq_cov, weight, beam_power, out=None, device=0, solver=0,
                          workspace_bytes=0, timing=None):
    """Wiener delay-transform operators of one polarisation on the GPU (powerspec.py:245-319).

    fourier [F, ndelay] complex (delay -> frequency DFT, powerspec.py:229), prior [ndelay], window [F],
    filt / freq_cov [nra, F, F] (float64 or complex128), weight [nel, nra, F], beam_power [nel, F].
    Returns (or fills ``out``, any complex64 array whose [ra, el] blocks are contiguous [ndelay, F]) the operator
    [nra, nel, ndelay, F].  Raises ``numpy.linalg.LinAlgError`` where ``scipy.linalg.cho_factor`` would.
    """
    lib = load_wiener()
    fourier = np.ascontiguousarray(fourier, dtype=np.complex128)
    nfreq, ndelay = fourier.shape
    prior = np.ascontiguousarray(prior, dtype=np.float64)
    window = np.ascontiguousarray(window, dtype=np.float64)
    filt = np.ascontiguousarray(filt, dtype=np.complex128 if np.iscomplexobj(filt) else np.float64)
    freq_cov = np.ascontiguousarray(freq_cov, dtype=np.complex128 if np.iscomplexobj(freq_cov) else np.float64)
    weight = np.ascontiguousarray(weight, dtype=np.float64)
    beam_power = np.ascontiguousarray(beam_power, dtype=np.float64)
    nel, nra, _ = weight.shape
    if prior.shape != (ndelay,) or window.shape != (nfreq,):
        raise ValueError("prior / window do not match the Fourier operator")
    if filt.shape != (nra, nfreq, nfreq) or freq_cov.shape != (nra, nfreq, nfreq):
        raise ValueError(f"filter / freq_cov must be [{nra}, {nfreq}, {nfreq}]")
    if weight.shape != (nel, nra, nfreq) or beam_power.shape != (nel, nfreq):
        raise ValueError("weight must be [nel, nra, nfreq] and beam_power [nel, nfreq]")
    if out is None:
        out = np.zeros((nra, nel, ndelay, nfreq), dtype=np.complex64)
    if out.dtype != np.complex64 or out.shape != (nra, nel, ndelay, nfreq):
        raise ValueError("out must be complex64 [nra, nel, ndelay, nfreq]")
    es = out.itemsize
    if out.size and (out.strides[3] != es or out.strides[2] != es * nfreq or out.strides[0] % es or out.strides[1] % es):
        raise ValueError("the [ndelay, nfreq] blocks of out must be contiguous")
    p = WienerProblem(
        device_id=int(device), nfreq=nfreq, ndelay=ndelay, nra=nra, nel=nel,
        filter_is_complex=int(np.iscomplexobj(filt)), cov_is_complex=int(np.iscomplexobj(freq_cov)), solver=int(solver),
        fourier=fourier.ctypes.data, prior=prior.ctypes.data, window=window.ctypes.data, filter=filt.ctypes.data,
        freq_cov=freq_cov.ctypes.data, weight=weight.ctypes.data, beam_power=beam_power.ctypes.data,
        out=out.ctypes.data, out_stride_ra=out.strides[0] // es if out.size else 0,
        out_stride_el=out.strides[1] // es if out.size else 0, workspace_bytes=int(workspace_bytes))
    rc = lib.dgwiener_construct(C.byref(p))
    if rc == 1:
        raise np.linalg.LinAlgError(lib.dgwiener_last_error().decode())
    if rc != 0:
        raise _glib.DGibbsError(f"dgwiener_construct failed ({rc}): {lib.dgwiener_last_error().decode()}")
    if timing is not None:
        timing.update(ms_device=p.ms_device, ms_compute=p.ms_compute, flop=p.flop, launches=p.launches)
    return out


class ConstructWienerDelayTransform(_ContainerTask):
    """Construct a Wiener filter that transforms maps from frequency into delay space (powerspec.py:118-369).

    Parameters as in the reference (powerspec.py:131-171): ``prior_amp``, ``prior_scale``, ``window``,
    ``window_lower_freq``, ``window_upper_freq``; plus ``device`` (CUDA device, default rank % device count).
    """

    prior_amp = config.Property(proptype=float, default=2.8e-5)
    prior_scale = config.Property(proptype=float, default=0.0)
    window = config.enum(
        ["uniform", "hann", "hanning", "hamming", "blackman", "nuttall", "blackman_nuttall", "blackman_harris",
         "tukey-0.5", "None"],
        default="uniform",
    )
    window_lower_freq = config.Property(proptype=float)
    window_upper_freq = config.Property(proptype=float)
    device = config.Property(proptype=int, default=None)

    def _device(self):
        if self.device is not None:
            return int(self.device)
        import torch

        rank, _ = _compat.comm_rank_size()
        n = torch.cuda.device_count()
        return rank % n if n else 0

    def process(self, data):
        """Build the ``DelayTransformOperator`` of a filtered ring map (powerspec.py:173-322)."""
        data.redistribute("el")
        npol, nfreq, nra, nel_local = data.weight.local_shape

        freq = np.asarray(data.freq)
        dfreq = np.median(np.abs(np.diff(freq)))
        window = self._get_window(freq)
        win_mask = window > 0

        # non-negative delays of the windowed band, in micro-seconds (powerspec.py:210-214)
        ntau = np.sum(win_mask, dtype=int)
        tau = np.fft.fftshift(np.fft.fftfreq(ntau, d=dfreq))
        tau = tau[tau >= 0.0]

        out = containers.DelayTransformOperator(delay=tau, axes_from=data, attrs_from=data)
        out.redistribute("el")
        D = out.filter[:].local_array
        D[:] = 0.0
        for attr in ["window", "window_lower_freq", "window_upper_freq"]:
            out.attrs[attr] = getattr(self, attr)

        # delay -> frequency DFT and the diagonal signal prior (powerspec.py:229-233)
        F = np.exp(2.0j * np.pi * np.outer(freq, tau)) / np.sqrt(ntau)
        Sdiag = self._get_prior(tau)

        wall = data.weight[:].local_array
        ball = data.dirty_beam_power[0].local_array
        device = self._device()
        self.last_timing = []
        for pp in range(npol):
            self.log.info(f"Polarisation {pp} of {npol}")
            # filter / freq_cov have no elevation axis: every rank needs all of them (powerspec.py:248-252)
            C_ = _gathered(data.freq_cov[pp]).transpose(2, 0, 1)
            K = _gathered(data.filter[pp]).transpose(2, 0, 1)
            t = {}
            wiener_delay_operator(
                F, Sdiag, window, K, C_, np.ascontiguousarray(wall[pp].transpose(2, 1, 0)),
                np.ascontiguousarray(ball[pp].T), out=D[pp], device=device, timing=t)
            self.last_timing.append(t)
        return out

    def _get_prior(self, delay):
        """``prior_amp * exp(-2 pi prior_scale |delay|)`` (powerspec.py:325-344)."""
        return self.prior_amp * np.exp(-2.0 * np.pi * self.prior_scale * np.abs(delay))

    def _get_window(self, freq):
        """Spectral window over [window_lower_freq, window_upper_freq] (powerspec.py:346-369)."""
        frng = np.percentile(freq, [0, 100])
        if self.window_lower_freq is not None:
            frng[0] = self.window_lower_freq
        if self.window_upper_freq is not None:
            frng[1] = self.window_upper_freq
        self.log.info(f"Applying a {self.window} window spanning {frng[0]:0.2f} - {frng[1]:0.2f} MHz.")
        x = (freq - frng[0]) / (frng[1] - frng[0])
        return _window(x, self.window)


def _gathered(arr):
    """The full array behind a distributed slice (``MPIArray.allgather``), or the array itself."""
    gather = getattr(arr, "allgather", None)
    return np.asarray(gather() if gather is not None else arr)
