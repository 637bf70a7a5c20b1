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

WIENER_EXPORTED = ["dgwiener_construct", "dgwiener_apply", "dgwiener_sizeof", "dgwiener_apply_sizeof", "dgwiener_last_error",
                   "dgwiener_gemm_ozaki"]
_vp = C.c_void_p


class WienerProblem(C.Structure):
    """``dgwiener_problem`` of include/dgwiener.h."""

    _fields_ = [
        ("device_id", C.c_int32), ("nfreq", C.c_int32), ("ndelay", C.c_int32), ("nra", C.c_int32), ("nel", C.c_int32),
        ("filter_is_complex", C.c_int32), ("cov_is_complex", C.c_int32), ("solver", C.c_int32),
        ("fourier", _vp), ("prior", _vp), ("window", _vp), ("filter", _vp), ("freq_cov", _vp), ("weight", _vp),
        ("beam_power", _vp), ("out", _vp),
        ("out_stride_ra", C.c_int64), ("out_stride_el", C.c_int64), ("workspace_bytes", C.c_int64),
        ("out_on_device", C.c_int32), ("reserved", C.c_int32),
        ("bad_el", C.c_int32), ("bad_ra", C.c_int32),
        ("ms_device", C.c_double), ("ms_compute", C.c_double), ("flop", C.c_double), ("launches", C.c_int64),
    ]


class WienerApplyProblem(C.Structure):
    """``dgwiener_apply_problem`` of include/dgwiener.h."""

    _fields_ = [
        ("device_id", C.c_int32), ("nfreq", C.c_int32), ("ndelay", C.c_int32), ("nra", C.c_int32), ("nel", C.c_int32),
        ("op_on_device", C.c_int32), ("op", _vp), ("op_stride_ra", C.c_int64), ("op_stride_el", C.c_int64),
        ("map", _vp), ("weight", _vp), ("spectrum", _vp), ("spectrum_weight", _vp), ("workspace_bytes", C.c_int64),
        ("ms_kernel", C.c_double), ("bytes", C.c_double), ("launches", C.c_int64),
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
        lib.dgwiener_apply.argtypes = [C.POINTER(WienerApplyProblem)]
        lib.dgwiener_apply.restype = C.c_int
        lib.dgwiener_apply_sizeof.restype = C.c_int
        lib.dgwiener_gemm_ozaki.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                            C.POINTER(C.c_double)]
        lib.dgwiener_gemm_ozaki.restype = C.c_int
        for st, fn in ((WienerProblem, lib.dgwiener_sizeof), (WienerApplyProblem, lib.dgwiener_apply_sizeof)):
            if fn() != C.sizeof(st):
                raise RuntimeError(f"ABI mismatch: {st.__name__} is {C.sizeof(st)} bytes here, {fn()} in libdgnrml.so; "
                                   "rebuild the library")
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


def _operator_view(op, shape):
    """(pointer, ra stride, el stride, on_device) of an operator [nra, nel, ndelay, nfreq] complex64 held either in a
    numpy array or in a CUDA ``torch`` tensor; its [ndelay, nfreq] blocks must be contiguous."""
    if isinstance(op, np.ndarray):
        if op.dtype != np.complex64 or op.shape != tuple(shape):
            raise ValueError(f"the operator must be complex64 {list(shape)}")
        es = op.itemsize
        st = [x // es for x in op.strides]
        bad = any(x % es for x in op.strides)
        ptr, on_dev = op.ctypes.data, False
    else:
        import torch

        if not (isinstance(op, torch.Tensor) and op.is_cuda and op.dtype == torch.complex64 and tuple(op.shape) == tuple(shape)):
            raise ValueError(f"the operator must be a numpy array or a CUDA torch tensor, complex64 {list(shape)}")
        st, bad = list(op.stride()), False
        ptr, on_dev = op.data_ptr(), True
    if int(np.prod(shape)) and (bad or st[3] != 1 or st[2] != shape[3]):
        raise ValueError("the [ndelay, nfreq] blocks of the operator must be contiguous")
    return ptr, (st[0] if int(np.prod(shape)) else 0), (st[1] if int(np.prod(shape)) else 0), on_dev


def wiener_delay_operator(fourier, prior, window, filt, freq_cov, weight, beam_power, out=None, device=0, solver=0,
                          workspace_bytes=0, timing=None):
    """Wiener delay-transform operators of one polarisation on the GPU (powerspec.py:245-319).

    fourier [F, ndelay] complex (delay -> frequency DFT, powerspec.py:229), prior [ndelay], window [F],
    filt / freq_cov [nra, F, F] (float64 or complex128), weight [nel, nra, F], beam_power [nel, F].
    Returns (or fills ``out``, any complex64 numpy array or CUDA torch tensor whose [ra, el] blocks are contiguous
    [ndelay, F]; a device tensor keeps the operator in HBM for ``apply_wiener_operator``) the operator
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
    ptr, sra, sel, on_dev = _operator_view(out, (nra, nel, ndelay, nfreq))
    if on_dev:
        import torch

        device = out.device.index
        torch.cuda.synchronize(out.device)  # the library works on its own stream
    p = WienerProblem(
        device_id=int(device), nfreq=nfreq, ndelay=ndelay, nra=nra, nel=nel,
        filter_is_complex=int(np.iscomplexobj(filt)), cov_is_complex=int(np.iscomplexobj(freq_cov)), solver=int(solver),
        fourier=fourier.ctypes.data, prior=prior.ctypes.data, window=window.ctypes.data, filter=filt.ctypes.data,
        freq_cov=freq_cov.ctypes.data, weight=weight.ctypes.data, beam_power=beam_power.ctypes.data,
        out=ptr, out_stride_ra=sra, out_stride_el=sel, workspace_bytes=int(workspace_bytes), out_on_device=int(on_dev))
    rc = lib.dgwiener_construct(C.byref(p))
    if rc == 1:
        raise np.linalg.LinAlgError(lib.dgwiener_last_error().decode())
    if rc != 0:
        raise _glib.DGibbsError(f"dgwiener_construct failed ({rc}): {lib.dgwiener_last_error().decode()}")
    if timing is not None:
        timing.update(ms_device=p.ms_device, ms_compute=p.ms_compute, flop=p.flop, launches=p.launches)
    return out


def gemm_nt_int8(A, B, device=0, timing=None):
    """``A @ B.T`` (last two axes; a leading batch axis is allowed) for float64 operands on the INT8 tensor cores: the
    contraction kernel of ``solver=4`` by itself (``dgwiener_gemm_ozaki``, ``csrc/ozaki_gemm.cuh``): 8 signed 7-bit slices per
    row and 256 columns, ``tcgen05.mma kind::i8``, int32 accumulators in TMEM -- results of fp64 quality (error ~1e-15 of the
    product of the row maxima).  Shapes: ``A [.., M, K]``, ``B [.., N, K]`` with ``M % 128 == N % 64 == K % 128 == 0``."""
    A = np.ascontiguousarray(A, dtype=np.float64)
    B = np.ascontiguousarray(B, dtype=np.float64)
    if A.ndim not in (2, 3) or B.ndim != A.ndim or A.shape[-1] != B.shape[-1] or (A.ndim == 3 and A.shape[0] != B.shape[0]):
        raise ValueError(f"gemm_nt_int8: incompatible shapes {A.shape} and {B.shape}")
    batch = A.shape[0] if A.ndim == 3 else 1
    M, K, N = A.shape[-2], A.shape[-1], B.shape[-2]
    if M % 128 or N % 64 or K % 128:
        raise ValueError("gemm_nt_int8: M % 128, N % 64 and K % 128 must be 0")
    out = np.empty(A.shape[:-1] + (N,), dtype=np.float64)
    ms = C.c_double(0.0)
    lib = load_wiener()
    rc = lib.dgwiener_gemm_ozaki(int(device), A.ctypes.data, B.ctypes.data, out.ctypes.data, M, N, K, batch, C.byref(ms))
    if rc != 0:
        raise _glib.DGibbsError(f"dgwiener_gemm_ozaki failed ({rc}): {lib.dgwiener_last_error().decode()}")
    if timing is not None:
        timing["ms_device"] = ms.value
    return out


def apply_wiener_operator(op, map_, weight, device=0, workspace_bytes=0, timing=None):
    """Apply the operators of one polarisation to its map on the GPU (powerspec.py:430-452).

    op [nra, nel, ndelay, F] complex64 (numpy, or a CUDA torch tensor: no upload), map_ / weight [F, nra, nel] (the
    container's own slices).  Returns ``spectrum [nel, nra, ndelay]`` complex128 and its ``weight`` float32.
    """
    lib = load_wiener()
    map_ = np.ascontiguousarray(map_, dtype=np.float64)
    weight = np.ascontiguousarray(weight, dtype=np.float64)
    nfreq, nra, nel = map_.shape
    if weight.shape != map_.shape:
        raise ValueError("map and weight must both be [nfreq, nra, nel]")
    if len(op.shape) != 4 or tuple(op.shape[:2]) != (nra, nel) or op.shape[3] != nfreq:
        raise ValueError(f"the operator must be [{nra}, {nel}, ndelay, {nfreq}]")
    ndelay = int(op.shape[2])
    ptr, sra, sel, on_dev = _operator_view(op, (nra, nel, ndelay, nfreq))
    spec = np.zeros((nel, nra, ndelay), dtype=np.complex128)
    sw = np.zeros((nel, nra, ndelay), dtype=np.float32)
    if on_dev:
        device = op.device.index
    p = WienerApplyProblem(
        device_id=int(device), nfreq=nfreq, ndelay=ndelay, nra=nra, nel=nel, op_on_device=int(on_dev), op=ptr,
        op_stride_ra=sra, op_stride_el=sel, map=map_.ctypes.data, weight=weight.ctypes.data, spectrum=spec.ctypes.data,
        spectrum_weight=sw.ctypes.data, workspace_bytes=int(workspace_bytes))
    if on_dev:
        import torch

        torch.cuda.synchronize(op.device)  # the library works on its own stream
    rc = lib.dgwiener_apply(C.byref(p))
    if rc != 0:
        raise _glib.DGibbsError(f"dgwiener_apply failed ({rc}): {lib.dgwiener_last_error().decode()}")
    if timing is not None:
        timing.update(ms_kernel=p.ms_kernel, bytes=p.bytes, launches=p.launches)
    return spec, sw


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


class ApplyWienerDelayTransform(_ContainerTask):
    """Apply a precomputed Wiener filter to project a ring map into delay space (powerspec.py:372-458).

    Same inputs and output as the reference: ``process(data, operator)`` with a ``RingMap`` (``map``, ``weight``) and
    the ``DelayTransformOperator`` of ``ConstructWienerDelayTransform``; returns a ``DelayTransform`` whose
    ``baseline`` axis flattens (pol, el) and whose ``sample`` axis is the right ascension, with a ``weight`` dataset.
    """

    device = config.Property(proptype=int, default=None)
    _device = ConstructWienerDelayTransform._device

    def process(self, data, operator):
        for cont in (data, operator):
            cont.redistribute("ra")
        w_in = data.weight[:]
        npol, nel = w_in.local_shape[0], w_in.local_shape[3]

        # output: (pol, el) flattened into `baseline`, right ascension as `sample` (powerspec.py:406-434)
        flat_axes = np.array(["pol", "el"])
        out = containers.DelayTransform(attrs_from=data, baseline=npol * nel, sample=data.index_map["ra"],
                                        delay=operator.index_map["delay"])
        out.add_dataset("weight")
        out.redistribute("sample")
        for name in flat_axes:
            out.create_index_map(name, data.index_map[name])
        out.attrs["baseline_axes"] = flat_axes
        out.attrs["freq"] = data.freq
        for suffix in ("", "_lower_freq", "_upper_freq"):
            out.attrs["window_los" + suffix] = operator.attrs["window" + suffix]

        weight, maps = w_in.local_array, data.map[0].local_array
        filt = operator.filter[:].local_array
        spec, sweight = out.spectrum[:].local_array, out.weight[:].local_array
        device = self._device()
        self.last_timing = []
        for pp in range(npol):
            eslc = slice(pp * nel, (pp + 1) * nel)
            t = {}
            sp, sw = apply_wiener_operator(filt[pp], maps[pp], weight[pp], device=device, timing=t)
            spec[eslc] = sp
            sweight[eslc] = sw
            self.last_timing.append(t)
        return out
