"""Python face of the device sampler: same names and argument meaning as the reference's
module-level functions in ``draco/analysis/delay.py`` (cited per function), plus the
batched entry point the task class uses.  All arithmetic happens in ``libdgibbs.so``.
"""

from __future__ import annotations

import ctypes as C
import dataclasses

import numpy as np

from . import _lib
from ._lib import RNG_INJECTED, RNG_PHILOX, RNG_ZERO, WINDOW_IDS, DGibbsError, Handle  # noqa: F401
from ._lib import BRANCH_DELAY_SPACE, BRANCH_DIAGONAL, BRANCH_FREQ_SPACE  # noqa: F401

_handles: dict = {}


def get_handle(device=0) -> Handle:
    """Per-process handle cache (one handle per device)."""
    h = _handles.get(device)
    if h is None:
        h = _handles[device] = Handle(device)
    return h


def rng_description() -> str:
    """Which generator ``rng_mode = RNG_PHILOX`` uses in the loaded build (``dgibbs_rng_info``)."""
    return _lib.load().dgibbs_rng_info().decode()


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


def _is_c64(a):
    """complex64 visibilities (numpy complex64, or a CUDA float32 tensor of interleaved re / im)."""
    if isinstance(a, np.ndarray):
        return a.dtype == np.complex64
    return str(a.dtype) == "torch.float32"


def _is_device(a):
    return not isinstance(a, np.ndarray) and hasattr(a, "is_cuda") and a.is_cuda


@dataclasses.dataclass
class PackedBatch:
    """Ragged batch of cut baselines in the layout ``include/dgibbs.h`` describes."""

    N: int
    T: object  # int32 [B]
    M: object  # int32 [B]
    data_off: object  # int64 [B+1]
    chan_off: object  # int64 [B+1]
    data: object  # complex128 flat
    Ni: object  # float64 flat
    fsel: object  # int64 flat
    S0: object  # float64 [B, N]
    global_id: object = None  # int64 [B]
    complex_timedomain: bool = False
    window: str | None = "nuttall"

    @property
    def B(self):
        return int(self.T.shape[0])

    @property
    def on_device(self):
        return _is_device(self.data)

    def to_device(self, device=0):
        """Move every array to ``cuda:device`` (torch tensors); inputs then stay resident in HBM."""
        import torch

        dev = torch.device("cuda", device)
        kw = {}
        for f in dataclasses.fields(self):
            v = getattr(self, f.name)
            if isinstance(v, np.ndarray):
                if v.dtype == np.complex128:
                    v = torch.from_numpy(v.view(np.float64)).to(dev)
                elif v.dtype == np.complex64:
                    v = torch.from_numpy(v.view(np.float32)).to(dev)
                else:
                    v = torch.from_numpy(np.ascontiguousarray(v).copy() if not v.flags.writeable else v).to(dev)
            kw[f.name] = v
        return PackedBatch(**kw)


def pack_batch(datas, Nis, fsels, N, S0, window="nuttall", complex_timedomain=False, global_id=None) -> PackedBatch:
    """Pack per-baseline ``(data[T,M] c128, Ni[M], fsel[M])`` triples (the outputs of
    ``_cut_data`` and ``channel_ind[nzf]``, delay.py:1156-1167) into one ragged batch."""
    B = len(datas)
    T = np.array([d.shape[0] for d in datas], dtype=np.int32)
    M = np.array([d.shape[1] for d in datas], dtype=np.int32)
    data_off = np.zeros(B + 1, dtype=np.int64)
    np.cumsum(T.astype(np.int64) * M, out=data_off[1:])
    chan_off = np.zeros(B + 1, dtype=np.int64)
    np.cumsum(M, out=chan_off[1:])
    data = np.empty(int(data_off[-1]), dtype=np.complex128)
    Ni = np.empty(int(chan_off[-1]), dtype=np.float64)
    fsel = np.empty(int(chan_off[-1]), dtype=np.int64)
    for b in range(B):
        if Nis[b].ndim != 1:
            # same restriction as the reference sampler (delay.py:1697 assumes 1-D Ni)
            raise ValueError("the Gibbs sampler needs time-averaged weights (use_average_weights=True)")
        data[data_off[b] : data_off[b + 1]] = np.asarray(datas[b], dtype=np.complex128).ravel()
        Ni[chan_off[b] : chan_off[b + 1]] = Nis[b]
        fsel[chan_off[b] : chan_off[b + 1]] = fsels[b]
    S0 = np.ascontiguousarray(np.broadcast_to(np.asarray(S0, dtype=np.float64), (B, N)))
    gid = None if global_id is None else np.ascontiguousarray(global_id, dtype=np.int64)
    return PackedBatch(N, T, M, data_off, chan_off, data, Ni, fsel, S0, gid, bool(complex_timedomain), window)


def pack_uniform(vis, weight, fsel, N, S0, window="nuttall", complex_timedomain=False, global_id=None) -> PackedBatch:
    """Fast path of :func:`pack_batch` when every baseline has the same (T, M):
    ``vis[B,T,M]`` c128 (complex64 is kept as it is: ``data_c64`` of the C ABI), ``weight[B,M]``, common ``fsel[M]``."""
    B, T, M = vis.shape
    data_off = np.arange(B + 1, dtype=np.int64) * (T * M)
    chan_off = np.arange(B + 1, dtype=np.int64) * M
    S0 = np.ascontiguousarray(np.broadcast_to(np.asarray(S0, dtype=np.float64), (B, N)))
    gid = None if global_id is None else np.ascontiguousarray(global_id, dtype=np.int64)
    return PackedBatch(
        N,
        np.full(B, T, dtype=np.int32),
        np.full(B, M, dtype=np.int32),
        data_off,
        chan_off,
        np.ascontiguousarray(vis, dtype=np.complex64 if vis.dtype == np.complex64 else np.complex128).reshape(-1),
        np.ascontiguousarray(weight, dtype=np.float64).reshape(-1),
        np.ascontiguousarray(np.broadcast_to(np.asarray(fsel, dtype=np.int64), (B, M))).reshape(-1),
        S0,
        gid,
        bool(complex_timedomain),
        window,
    )


def _upload_as_it_lies(a, dev):
    """``a`` (numpy) as a contiguous CUDA tensor.  A transposed view of a C-contiguous block -- what ``flatten_axes``
    (delay.py:2238-2302) hands over for a container stored [freq, stack, ra] -- is shipped in its memory order and
    permuted on the device (a strided host copy of 1 GB costs seconds; the device permutation well under a millisecond)."""
    import torch

    if a.flags.c_contiguous or a.ndim < 2 or 0 in a.strides:
        return torch.from_numpy(np.ascontiguousarray(a)).to(dev, non_blocking=True)
    order = np.argsort([-st for st in a.strides], kind="stable")
    block = a.transpose(order)
    if not block.flags.c_contiguous:
        return torch.from_numpy(np.ascontiguousarray(a)).to(dev, non_blocking=True)
    return torch.from_numpy(block).to(dev, non_blocking=True).permute(*np.argsort(order).tolist()).contiguous()


@dataclasses.dataclass
class CutResult:
    """Outcome of :func:`cut_batch`: ``batch`` holds the surviving baselines (device resident), ``index`` their
    positions in the raw slab; ``alive``/``keep_t``/``keep_f`` are host arrays over the whole slab."""

    batch: PackedBatch | None
    index: np.ndarray  # int64 [B_alive]
    alive: np.ndarray  # bool [B]
    keep_t: np.ndarray  # bool [B, T]
    keep_f: np.ndarray  # bool [B, nfreq]


def cut_batch(vis, weight, channel_ind, N, S0, *, time_frac=-1.0, freq_frac=-1.0, remove_mean=False, scale_freq=False,
              weight_boost=1.0, window="nuttall", complex_timedomain=False, global_id=None, handle: Handle | None = None,
              device=0) -> CutResult:
    """``DelayTransformBase._cut_data`` (delay.py:516-597) for a whole raw slab on the GPU, packed in place as
    the ragged batch :func:`run_batch` / :func:`wiener_batch` consume (replaces the loop delay.py:1146-1167).

    ``vis`` [B, T, nfreq] complex128 or complex64 (the dtype draco's containers store, core/containers.py:503; it is
    uploaded as it is -- half the bytes -- and cut in numpy's complex64 arithmetic, bit for bit) and ``weight``
    [B, T, nfreq] float32/float64 (numpy, or CUDA torch tensors with ``vis`` viewed as float64 / float32 [..., 2]);
    ``channel_ind`` [nfreq]; ``S0`` broadcastable to [B, N].  Weights are always averaged over the kept samples
    (``use_average_weights=True``).  ``scale_freq`` needs complex128 input."""
    import torch

    h = handle or get_handle(device)
    dev = torch.device("cuda", h.device if hasattr(h, "device") else device)

    def to_dev(a, dtype=None):
        if isinstance(a, np.ndarray):
            if dtype is not None and a.dtype != dtype:
                a = np.ascontiguousarray(a, dtype=dtype)
            t = _upload_as_it_lies(a, dev)
            return torch.view_as_real(t) if t.is_complex() else t
        return a.contiguous()

    if isinstance(vis, np.ndarray):
        if vis.dtype not in (np.complex64, np.complex128):
            vis = np.asarray(vis, dtype=np.complex128)
        B, T, F = vis.shape
        c64 = vis.dtype == np.complex64
    else:
        B, T, F = vis.shape[:3]
        c64 = vis.dtype == torch.float32
        if not c64 and vis.dtype != torch.float64:
            raise ValueError("device-resident vis must be float64 or float32 [B, T, nfreq, 2]")
    if c64 and scale_freq:
        raise ValueError("scale_freq on complex64 data is not served by the device cut; use the host _cut_data")
    if isinstance(weight, np.ndarray) and weight.dtype not in (np.float32, np.float64):
        weight = weight.astype(np.float64)
    d_vis, d_w = to_dev(vis), to_dev(weight)
    if tuple(d_w.shape) != (B, T, F):
        raise ValueError(f"weight shape {tuple(d_w.shape)} does not match the data {(B, T, F)}")
    w32 = d_w.dtype == torch.float32
    if not w32 and d_w.dtype != torch.float64:
        raise ValueError("weight must be float32 or float64")
    d_cind = to_dev(np.asarray(channel_ind, dtype=np.int64))
    d_out = torch.empty((B, T * F, 2), dtype=torch.float64, device=dev)
    d_Ni = torch.empty((B, F), dtype=torch.float64, device=dev)
    d_fsel = torch.empty((B, F), dtype=torch.int64, device=dev)
    d_T = torch.empty(B, dtype=torch.int32, device=dev)
    d_M = torch.empty(B, dtype=torch.int32, device=dev)
    d_alive = torch.empty(B, dtype=torch.uint8, device=dev)
    d_kt = torch.zeros((B, T), dtype=torch.uint8, device=dev)
    d_kf = torch.zeros((B, F), dtype=torch.uint8, device=dev)
    c = _lib.Cut()
    c.struct_size = C.sizeof(_lib.Cut)
    c.on_device = 1
    c.B, c.T, c.nfreq = B, T, F
    c.weight_f32 = 1 if w32 else 0
    c.data_c64 = 1 if c64 else 0
    c.remove_mean, c.scale_freq = int(bool(remove_mean)), int(bool(scale_freq))
    c.time_frac, c.freq_frac, c.weight_boost = float(time_frac), float(freq_frac), float(weight_boost)
    c.data, c.weight, c.channel_ind = _ptr(d_vis), _ptr(d_w), _ptr(d_cind)
    c.data_out, c.Ni_out, c.fsel_out = _ptr(d_out), _ptr(d_Ni), _ptr(d_fsel)
    c.T_out, c.M_out, c.alive, c.keep_t, c.keep_f = _ptr(d_T), _ptr(d_M), _ptr(d_alive), _ptr(d_kt), _ptr(d_kf)
    torch.cuda.current_stream(dev).synchronize()  # the uploads above ran on torch's stream
    h.check(h.lib.dgibbs_cut_data(h._h, C.byref(c)), "dgibbs_cut_data")
    alive = d_alive.cpu().numpy().astype(bool)
    Tk, Mk = d_T.cpu().numpy(), d_M.cpu().numpy()
    keep_t, keep_f = d_kt.cpu().numpy().astype(bool), d_kf.cpu().numpy().astype(bool)
    index = np.flatnonzero(alive).astype(np.int64)
    if index.size == 0:
        return CutResult(None, index, alive, keep_t, keep_f)
    data_off = np.concatenate([index * (T * F), [B * T * F]]).astype(np.int64)
    chan_off = np.concatenate([index * F, [B * F]]).astype(np.int64)
    S0 = np.ascontiguousarray(np.broadcast_to(np.asarray(S0, dtype=np.float64), (B, N))[index])
    gid = index if global_id is None else np.ascontiguousarray(np.asarray(global_id, dtype=np.int64)[index])

    def up(a):
        return torch.from_numpy(np.ascontiguousarray(a)).to(dev)

    pb = PackedBatch(int(N), up(Tk[index]), up(Mk[index]), up(data_off), up(chan_off), d_out.view(-1), d_Ni.view(-1),
                     d_fsel.view(-1), up(S0), up(gid), bool(complex_timedomain), window)
    return CutResult(pb, index, alive, keep_t, keep_f)



def _problem(pb: PackedBatch, niter, rng_mode, seed, median_frac, draws=None, force_branch=0, cluster_size=0):
    p = _lib.Problem()
    p.struct_size = C.sizeof(_lib.Problem)
    p.on_device = 1 if pb.on_device else 0
    p.B, p.N, p.niter = pb.B, int(pb.N), int(niter)
    if pb.window not in WINDOW_IDS:
        raise ValueError(f"window {pb.window!r} is not one of {sorted(k for k in WINDOW_IDS if k)}")
    p.window_id = WINDOW_IDS[pb.window]
    p.complex_timedomain = 1 if pb.complex_timedomain else 0
    p.rng_mode = rng_mode
    p.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    p.median_frac = float(median_frac)
    p.force_branch = int(force_branch)
    p.cluster_size = int(cluster_size)
    for name in ("T", "M", "data_off", "chan_off", "data", "Ni", "fsel", "S0", "global_id"):
        setattr(p, name, _ptr(getattr(pb, name)))
    p.data_c64 = 1 if _is_c64(pb.data) else 0
    keep = []
    if draws is not None:
        w1, w1_off, w2, w2_off, chi2 = draws
        keep = [w1, w1_off, w2, w2_off, chi2]
        p.w1, p.w1_off, p.w2, p.w2_off, p.chi2 = map(_ptr, keep)
    return p, keep


def pack_draws(pb: PackedBatch, niter, w1s, w2s, chi2s):
    """Concatenate per-baseline injected draws: ``w1s[b]`` (niter, n_sys, T), ``w2s[b]``
    (niter, 2M, T), ``chi2s[b]`` (niter, N) -- the reference's shapes and order."""
    B = pb.B
    w1_off = np.zeros(B + 1, dtype=np.int64)
    w2_off = np.zeros(B + 1, dtype=np.int64)
    np.cumsum([int(np.asarray(w).size) for w in w1s], out=w1_off[1:])
    np.cumsum([int(np.asarray(w).size) for w in w2s], out=w2_off[1:])
    w1 = np.concatenate([np.asarray(w, dtype=np.float64).ravel() for w in w1s]) if niter else np.zeros(1)
    w2 = np.concatenate([np.asarray(w, dtype=np.float64).ravel() for w in w2s]) if niter else np.zeros(1)
    chi2 = np.ascontiguousarray(np.stack([np.asarray(c, dtype=np.float64) for c in chi2s])) if niter else np.zeros(1)
    return w1, w1_off, w2, w2_off, chi2


@dataclasses.dataclass
class BatchResult:
    spectrum: object  # [B, N] fftshifted median
    samples: object  # [niter, B, N] or None
    n_done: object  # [B]
    success: object  # [B] bool
    branch: object  # [B]
    final_S: object
    timing: dict


def run_batch(
    pb: PackedBatch,
    niter,
    *,
    handle: Handle | None = None,
    rng_mode=RNG_PHILOX,
    seed=0,
    median_frac=0.5,
    draws=None,
    save_samples=False,
    want_final_S=False,
    force_branch=0,
    cluster_size=0,
) -> BatchResult:
    """Run every chain of a packed batch on the GPU (replaces the loop delay.py:1146-1178).

    ``cluster_size``: CTAs that share one chain of the dense branch (0 = planned per launch so that the last wave
    fills the GPU; 1, 2, 4 force it).  The samples do not depend on it."""
    device = pb.data.device.index if pb.on_device else 0
    h = handle or get_handle(device or 0)
    B, N = pb.B, int(pb.N)
    if draws is not None:
        rng_mode = RNG_INJECTED
    if rng_mode == RNG_INJECTED and draws is None and niter > 0:
        raise ValueError("injected rng mode needs draws")
    if pb.on_device:
        import torch

        dev = pb.data.device
        if draws is not None:
            draws = tuple(torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in draws)
        spectrum = torch.empty((B, N), dtype=torch.float64, device=dev)
        samples = torch.empty((niter, B, N), dtype=torch.float64, device=dev) if save_samples else None
        n_done = torch.empty(B, dtype=torch.int32, device=dev)
        success = torch.empty(B, dtype=torch.uint8, device=dev)
        branch = torch.empty(B, dtype=torch.uint8, device=dev)
        final_S = torch.empty((B, N), dtype=torch.float64, device=dev) if want_final_S else None
    else:
        spectrum = np.empty((B, N), dtype=np.float64)
        samples = np.empty((niter, B, N), dtype=np.float64) if save_samples else None
        n_done = np.empty(B, dtype=np.int32)
        success = np.empty(B, dtype=np.uint8)
        branch = np.empty(B, dtype=np.uint8)
        final_S = np.empty((B, N), dtype=np.float64) if want_final_S else None
    p, keep = _problem(pb, niter, rng_mode, seed, median_frac, draws, force_branch, cluster_size)
    r = _lib.Result()
    r.struct_size = C.sizeof(_lib.Result)
    r.on_device = p.on_device
    r.spectrum, r.samples, r.n_done = _ptr(spectrum), _ptr(samples), _ptr(n_done)
    r.success, r.branch, r.final_S = _ptr(success), _ptr(branch), _ptr(final_S)
    h.check(h.lib.dgibbs_run(h._h, C.byref(p), C.byref(r)), "dgibbs_run")
    del keep
    ok = success.bool() if pb.on_device else success.astype(bool)
    return BatchResult(spectrum, samples, n_done, ok, branch, final_S, h.timing())


# ---------------------------------------------------------------------------------------
# Reference-named single-baseline functions
# ---------------------------------------------------------------------------------------


def _default_fsel(N, complex_timedomain):
    return np.arange(N if complex_timedomain else N // 2 + 1)


def _host_draws(rng, niter, n_sys, M, T, N):
    """Draw (w1, w2, chi2) for one baseline from a numpy Generator in the reference's call
    order (delay.py:1806-1809, 1878): per iteration normal(n_sys,T), normal(2M,T), chisquare(T,N)."""
    w1 = np.empty((niter, n_sys, T))
    w2 = np.empty((niter, 2 * M, T))
    c2 = np.empty((niter, N))
    for i in range(niter):
        w1[i] = rng.standard_normal((n_sys, T))
        w2[i] = rng.standard_normal((2 * M, T))
        c2[i] = rng.chisquare(T, size=N)
    return w1, w2, c2


def delay_power_spectrum_gibbs(
    data,
    N,
    Ni,
    initial_S,
    window="nuttall",
    fsel=None,
    niter=20,
    rng=None,
    complex_timedomain=False,
    *,
    device=0,
    seed=None,
):
    """Drop-in for ``delay_power_spectrum_gibbs`` (delay.py:1713-1904): returns
    ``(list of (N,) spectrum samples in unshifted delay order, success)``.

    ``rng``: a ``numpy.random.Generator`` selects deterministic mode -- the Gaussian and
    chi^2 variates are drawn from it on the host in the reference's order and injected, so
    the same seed reproduces the reference's chain to rounding.  ``rng=None`` uses the
    device Philox generator (``seed``; default taken from OS entropy).
    """
    data = np.asarray(data)
    fsel = _default_fsel(N, complex_timedomain) if fsel is None else np.asarray(fsel)
    T, M = data.shape
    pb = pack_batch([data], [np.asarray(Ni, dtype=np.float64)], [fsel], N, np.asarray(initial_S)[None, :], window,
                    complex_timedomain)
    if rng is not None:
        n_sys = 2 * N if complex_timedomain else N
        w1, w2, c2 = _host_draws(rng, niter, n_sys, M, T, N)
        draws = pack_draws(pb, niter, [w1], [w2], [c2])
        res = run_batch(pb, niter, handle=get_handle(device), draws=draws, save_samples=True)
    else:
        if seed is None:
            seed = int(np.random.SeedSequence().generate_state(1, dtype=np.uint64)[0])
        res = run_batch(pb, niter, handle=get_handle(device), rng_mode=RNG_PHILOX, seed=seed, save_samples=True)
    nd = int(res.n_done[0])
    spec = [res.samples[niter - nd + i, 0].copy() for i in range(nd)]
    return spec, bool(res.success[0])


def delay_spectrum_wiener_filter(delay_PS, data, N, Ni, window="nuttall", fsel=None, complex_timedomain=False, *,
                                 device=0):
    """Drop-in for ``delay_spectrum_wiener_filter`` (delay.py:2132-2201): ``(T, N)`` float64,
    or complex128 in complex_timedomain mode.  Raises ``np.linalg.LinAlgError`` when the
    Wiener covariance is not positive definite, like ``scipy.linalg.cho_factor``."""
    data = np.asarray(data)
    fsel = _default_fsel(N, complex_timedomain) if fsel is None else np.asarray(fsel)
    pb = pack_batch([data], [np.asarray(Ni, dtype=np.float64)], [fsel], N, np.asarray(delay_PS)[None, :], window,
                    complex_timedomain)
    out, ok = wiener_batch(pb, handle=get_handle(device))
    if not ok[0]:
        raise np.linalg.LinAlgError("Wiener covariance is not positive definite")
    y = out[0]
    return y.view(np.complex128) if complex_timedomain else y


def wiener_batch(pb: PackedBatch, *, handle: Handle | None = None, cluster_size=0):
    """Wiener-filter every baseline of a packed batch (host arrays, or the device-resident batch
    :func:`cut_batch` leaves behind); returns a list of (T_b, n_sys) host arrays and a success flag per baseline."""
    n_sys = (2 if pb.complex_timedomain else 1) * int(pb.N)
    p, keep = _problem(pb, 1, RNG_ZERO, 0, 0.5, cluster_size=cluster_size)
    if pb.on_device:
        import torch

        dev = pb.data.device
        h = handle or get_handle(dev.index or 0)
        Ts = pb.T.cpu().numpy()
        d_out = torch.empty(int(Ts.astype(np.int64).sum()) * n_sys, dtype=torch.float64, device=dev)
        d_ok = torch.empty(pb.B, dtype=torch.uint8, device=dev)
        h.check(h.lib.dgibbs_wiener(h._h, C.byref(p), _ptr(d_out), _ptr(d_ok)), "dgibbs_wiener")
        out, ok = d_out.cpu().numpy(), d_ok.cpu().numpy()
    else:
        h = handle or get_handle(0)
        Ts = np.asarray(pb.T)
        out = np.empty(int(Ts.astype(np.int64).sum()) * n_sys, dtype=np.float64)
        ok = np.empty(pb.B, dtype=np.uint8)
        h.check(h.lib.dgibbs_wiener(h._h, C.byref(p), out.ctypes.data, ok.ctypes.data), "dgibbs_wiener")
    del keep
    res, o = [], 0
    for b in range(pb.B):
        n = int(Ts[b]) * n_sys
        res.append(out[o : o + n].reshape(int(Ts[b]), n_sys))
        o += n
    return res, ok.astype(bool)


def fft_rows(rows, window="nuttall", *, shift=False, handle: Handle | None = None, device=0, timing=None):
    """``delay_spectrum_fft`` (delay.py:2102-2130) for a stack of rows ``[..., N]`` complex128 on the GPU
    (``dgibbs_delay_fft``): ``ifft(rows * window(i / N))``, optionally fftshifted along the delay axis (the task's
    ``np.fft.fftshift``, delay.py:980).  numpy in / numpy out, or a CUDA float64 tensor ``[..., N, 2]`` in / out."""
    if window not in WINDOW_IDS:
        raise ValueError(f"window {window!r} is not one of {sorted(k for k in WINDOW_IDS if k)}")
    ms = C.c_double(0.0)
    if isinstance(rows, np.ndarray):
        x = np.ascontiguousarray(rows, dtype=np.complex128)
        N = x.shape[-1]
        out = np.empty_like(x)
        h = handle or get_handle(device)
        h.check(h.lib.dgibbs_delay_fft(h._h, 0, x.size // max(N, 1), N, WINDOW_IDS[window], int(bool(shift)), _ptr(x),
                                       _ptr(out), C.byref(ms)), "dgibbs_delay_fft")
    else:
        import torch

        x = rows.contiguous()
        if x.dtype != torch.float64 or x.shape[-1] != 2:
            raise ValueError("device rows must be float64 [..., N, 2]")
        N = x.shape[-2]
        out = torch.empty_like(x)
        h = handle or get_handle(x.device.index or 0)
        torch.cuda.current_stream(x.device).synchronize()
        h.check(h.lib.dgibbs_delay_fft(h._h, 1, x.numel() // max(2 * N, 1), N, WINDOW_IDS[window], int(bool(shift)),
                                       _ptr(x), _ptr(out), C.byref(ms)), "dgibbs_delay_fft")
    if timing is not None:
        timing["ms_kernel"] = ms.value
    return out


def fft_rows_noalloc(rows, out, window="nuttall", *, shift=False, handle: Handle | None = None, sync=True):
    """:func:`fft_rows` through ``dgibbs_delay_fft_noalloc``: CUDA float64 tensors ``rows`` / ``out`` ``[..., N, 2]`` owned by
    the caller, N = 2^m in [8, 4096]; the library allocates nothing (its tables live in a static device array).  No
    counterpart in the reference.  ``sync=False`` only enqueues the
    kernels on the handle's stream.  Calls that overlap on one GPU race on the shared tables."""
    import torch

    if window not in WINDOW_IDS:
        raise ValueError(f"window {window!r} is not one of {sorted(k for k in WINDOW_IDS if k)}")
    for t in (rows, out):
        if not _is_device(t) or t.dtype != torch.float64 or t.shape[-1] != 2 or not t.is_contiguous():
            raise ValueError("rows and out must be contiguous CUDA float64 tensors [..., N, 2]")
    if rows.shape != out.shape:
        raise ValueError("rows and out must have the same shape")
    N = rows.shape[-2]
    h = handle or get_handle(rows.device.index or 0)
    h.check(h.lib.dgibbs_delay_fft_noalloc(h._h, rows.numel() // max(2 * N, 1), N, WINDOW_IDS[window], int(bool(shift)),
                                           _ptr(rows), _ptr(out), int(bool(sync))), "dgibbs_delay_fft_noalloc")
    return out


def delay_spectrum_fft(data, N, window="nuttall", *, device=0):
    """Drop-in for ``delay_spectrum_fft`` (delay.py:2102-2130): ``data[nsample, N]`` -> ``[nsample, N]`` complex128,
    unshifted.  (The reference multiplies ``data`` by the window in place; the input is left untouched here.)"""
    data = np.asarray(data)
    if data.shape[-1] != N:  # the reference's `data *= window` cannot broadcast otherwise
        raise ValueError(f"operands could not be broadcast together with shapes {data.shape} (1,{N})")
    return fft_rows(data, window, shift=False, device=device)


def sample_variance(spec, mask=None, *, handle: Handle | None = None, device=0, timing=None):
    """``np.var(spec, axis=1, where=~mask[..., None])`` of a ``[B, T, N]`` complex128 delay spectrum on the GPU
    (``dgibbs_sample_variance``; DelaySpectrumToPowerSpectrum.process, delay.py:1061-1106).  Returns ``(var[B, N],
    bad[B])``; with a mask, NaN variances are replaced by 0 and flagged in ``bad`` (delay.py:1099-1104)."""
    x = np.ascontiguousarray(spec, dtype=np.complex128)
    if x.ndim != 3:
        raise ValueError("spec must be [baseline, sample, delay]")
    B, T, N = x.shape
    var = np.empty((B, N), dtype=np.float64)
    bad = np.zeros(B, dtype=np.uint8)
    mk = None if mask is None else np.ascontiguousarray(np.asarray(mask).astype(bool).reshape(B, T), dtype=np.uint8)
    ms = C.c_double(0.0)
    h = handle or get_handle(device)
    for b0 in range(0, B, 65535):
        b1 = min(B, b0 + 65535)
        h.check(h.lib.dgibbs_sample_variance(h._h, 0, b1 - b0, T, N, _ptr(x[b0:b1]), None if mk is None else _ptr(mk[b0:b1]),
                                             _ptr(var[b0:b1]), _ptr(bad[b0:b1]), C.byref(ms)), "dgibbs_sample_variance")
    if timing is not None:
        timing["ms_kernel"] = ms.value
    return var, bad.astype(bool)


def compute_delay_spectrum_inputs(data, N, Ni, fsel, window, complex_timedomain, *, device=0):
    """Device evaluation of ``_compute_delay_spectrum_inputs`` (delay.py:1657-1710):
    ``(data_w (2M,T), FTNih (n_sys,2M), FTNiF (n_sys,n_sys))``."""
    data = np.asarray(data)
    fsel = _default_fsel(N, complex_timedomain) if fsel is None else np.asarray(fsel)
    T, M = data.shape
    n_sys = 2 * N if complex_timedomain else N
    pb = pack_batch([data], [np.asarray(Ni, dtype=np.float64)], [fsel], N, np.ones((1, N)), window, complex_timedomain)
    h = get_handle(device)
    dw = np.empty((2 * M, T))
    fh = np.empty((n_sys, 2 * M))
    ff = np.empty((n_sys, n_sys))
    p, _ = _problem(pb, 0, RNG_ZERO, 0, 0.5)
    h.check(
        h.lib.dgibbs_setup_inputs(h._h, C.byref(p), 0, dw.ctypes.data, fh.ctypes.data, ff.ctypes.data),
        "dgibbs_setup_inputs",
    )
    return dw, fh, ff


def _fourier(kind, N, fsel, default_count, shape_fn, device):
    fa = np.arange(default_count, dtype=np.int64) if fsel is None else np.ascontiguousarray(fsel, dtype=np.int64)
    M = fa.shape[0]
    out = np.empty(shape_fn(M), dtype=np.float64)
    h = get_handle(device)
    h.check(h.lib.dgibbs_fourier_matrix(h._h, kind, int(N), fa.ctypes.data, M, out.ctypes.data), "dgibbs_fourier_matrix")
    return out


def fourier_matrix_r2c(N, fsel=None, *, device=0):
    """delay.py:1480-1510."""
    return _fourier(0, N, fsel, N // 2 + 1, lambda M: (2 * M, N), device)


def fourier_matrix_c2r(N, fsel=None, *, device=0):
    """delay.py:1513-1546."""
    return _fourier(1, N, fsel, N // 2 + 1, lambda M: (N, 2 * M), device)


def fourier_matrix_c2c(N, fsel=None, *, device=0):
    """delay.py:1549-1585."""
    return _fourier(2, N, fsel, N, lambda M: (2 * M, 2 * N), device)


def fourier_matrix(N, fsel=None, *, device=0):
    """delay.py:1588-1613 (complex128 (M, N))."""
    return _fourier(3, N, fsel, N, lambda M: (M, 2 * N), device).view(np.complex128)


# ---------------------------------------------------------------------------------------
# Cross-power-spectrum sampler (delay.py:1907-2099)
# ---------------------------------------------------------------------------------------


def _host_cross_draws(rng, niter, N, nd, M, T):
    """Draw one problem's variates from a numpy Generator in the reference's call order
    (delay.py:2029-2030 and, per delay, draco/util/random.py:121-130)."""
    L = nd * (nd - 1) // 2
    w1 = np.empty((niter, N, nd, T), dtype=np.complex128)
    w2 = np.empty((niter, nd, M, T), dtype=np.complex128)
    wn = np.empty((niter, N, 2, L))
    wg = np.empty((niter, N, nd))
    for i in range(niter):
        rng.standard_normal((N, nd, 2 * T), out=w1[i].view(np.float64))
        w1[i] *= 1.0 / 2**0.5
        rng.standard_normal((nd, M, 2 * T), out=w2[i].view(np.float64))
        w2[i] *= 1.0 / 2**0.5
        for t in range(N):
            wn[i, t, 0] = rng.standard_normal(L)
            wn[i, t, 1] = rng.standard_normal(L)
            for j in range(nd):
                wg[i, t, j] = rng.gamma(T - j)
    return w1, w2, wn, wg


def cross_batch(data, N, Ni, initial_S, window="nuttall", fsel=None, niter=20, *, draws=None, seed=0, global_id=None,
                handle: Handle | None = None, rng_mode=None):
    """Run ``B`` independent cross-spectrum chains on the GPU.

    ``data`` [B, nd, T, M] c128, ``Ni`` [B, nd, M], ``initial_S`` [B, nd, nd, N] c128, common ``fsel`` [M].
    ``draws = (w1, w2, wn, wg)`` (shapes in ``include/dgibbs.h``) selects the injected mode.
    Returns ``(samples [B, niter, nd, nd, N] c128, n_done [B], success [B])``."""
    data = np.ascontiguousarray(data, dtype=np.complex128)
    B, nd, T, M = data.shape
    if nd > _lib.CROSS_MAX_DATASETS:
        raise ValueError(f"at most {_lib.CROSS_MAX_DATASETS} datasets are supported")
    Ni = np.ascontiguousarray(np.broadcast_to(np.asarray(Ni, dtype=np.float64), (B, nd, M)))
    fsel = np.arange(M, dtype=np.int64) if fsel is None else np.ascontiguousarray(fsel, dtype=np.int64)
    if len(fsel) != M:  # delay.py:1962-1966
        raise ValueError(f"Length of frequency selection must match frequencies passed. {len(fsel)} != {M}")
    S0 = np.ascontiguousarray(np.broadcast_to(np.asarray(initial_S, dtype=np.complex128), (B, nd, nd, N)))
    if window not in WINDOW_IDS:
        raise ValueError(f"window {window!r} is not one of {sorted(k for k in WINDOW_IDS if k)}")
    h = handle or get_handle(0)
    p = _lib.CrossProblem()
    p.struct_size = C.sizeof(_lib.CrossProblem)
    p.B, p.nd, p.N, p.T, p.M, p.niter = B, nd, int(N), T, M, int(niter)
    p.window_id = WINDOW_IDS[window]
    p.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    keep = [data, Ni, fsel, S0]
    p.data, p.Ni, p.fsel, p.S0 = (a.ctypes.data for a in keep)
    if global_id is not None:
        gid = np.ascontiguousarray(global_id, dtype=np.int64)
        keep.append(gid)
        p.global_id = gid.ctypes.data
    if draws is not None:
        w1, w2, wn, wg = draws
        L = nd * (nd - 1) // 2
        w1 = np.ascontiguousarray(w1, dtype=np.complex128).reshape(B, niter, N, nd, T)
        w2 = np.ascontiguousarray(w2, dtype=np.complex128).reshape(B, niter, nd, M, T)
        wn = np.ascontiguousarray(wn, dtype=np.float64).reshape(B, niter, N, 2, L)
        wg = np.ascontiguousarray(wg, dtype=np.float64).reshape(B, niter, N, nd)
        if wn.size == 0:
            wn = np.zeros(1)
        keep += [w1, w2, wn, wg]
        p.w1, p.w2, p.wn, p.wg = w1.ctypes.data, w2.ctypes.data, wn.ctypes.data, wg.ctypes.data
        p.rng_mode = RNG_INJECTED
    else:
        p.rng_mode = RNG_PHILOX if rng_mode is None else rng_mode
    samples = np.empty((B, niter, nd, nd, N), dtype=np.complex128)
    n_done = np.empty(B, dtype=np.int32)
    ok = np.empty(B, dtype=np.uint8)
    h.check(h.lib.dgibbs_cross_run(h._h, C.byref(p), samples.ctypes.data, n_done.ctypes.data, ok.ctypes.data),
            "dgibbs_cross_run")
    del keep
    return samples, n_done, ok.astype(bool)


def delay_spectrum_gibbs_cross(data, N, Ni, initial_S, window="nuttall", fsel=None, niter=20, rng=None, *, device=0,
                               seed=None):
    """Drop-in for ``delay_spectrum_gibbs_cross`` (delay.py:1907-2099): ``data`` [nd, T, M] c128,
    ``Ni`` [nd, M], ``initial_S`` [nd, nd, N]; returns the list of ``[nd, nd, N]`` complex samples and raises
    ``RuntimeError("Exiting earlier as singular")`` where the reference does (delay.py:2096-2097).

    ``rng``: a ``numpy.random.Generator`` selects deterministic mode (its variates are drawn on the host in the
    reference's order and injected); ``rng=None`` uses the device Philox generator."""
    data = np.asarray(data)
    nd, T, M = data.shape
    if nd == 0:
        raise ValueError("Need at least one set of data")
    draws = None
    if rng is not None:
        draws = tuple(a[None] for a in _host_cross_draws(rng, niter, N, nd, M, T))
    elif seed is None:
        seed = int(np.random.SeedSequence().generate_state(1, dtype=np.uint64)[0])
    samples, n_done, ok = cross_batch(data[None], N, np.asarray(Ni)[None], np.asarray(initial_S)[None], window, fsel, niter,
                                      draws=draws, seed=seed or 0, handle=get_handle(device))
    if not ok[0]:
        raise RuntimeError("Exiting earlier as singular")
    return [samples[0, i].copy() for i in range(niter)]
