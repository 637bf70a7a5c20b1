"""caput-task face of the B200 delay power-spectrum estimator.

Drop-in for ``draco.analysis.delay.DelayPowerSpectrumGibbs`` (delay.py:1218-1267) and, as the next
row of the scope table, ``DelaySpectrumWienerFilter`` (delay.py:982-1053): same class names, same
YAML parameters and defaults (delay.py:403-429, 693-694, 761-763, 831, 1235-1236), same
``setup`` / ``process`` contract and the same ``DelaySpectrum`` / ``DelayTransform`` output layout
(draco/core/containers.py:2049-2182).  Only the per-baseline Python loop of ``_evaluate``
(delay.py:1146-1178) is replaced: all local baselines are cut on the host, packed into one ragged
batch and run on the GPU through the C ABI (``include/dgibbs.h``), the sampler staying on the
device for all ``nsamp`` iterations.

With caput installed the classes derive from caput's task base classes; otherwise (this image)
from the stand-ins in ``_compat``.  Historical names (``DelayPowerSpectrumStochasticEstimator``,
``DelaySpectrumEstimator``) and the old ``average_axis`` parameter are accepted as aliases.
"""

from __future__ import annotations

import numpy as np

from . import _compat, gibbs
from ._compat import comm_rank_size

if _compat.HAVE_CAPUT:  # pragma: no cover - caput is absent from this image
    from caput import config
    from caput.pipeline import tasklib

    _ContainerTask = tasklib.base.ContainerTask
    _RandomTask = tasklib.random.RandomTask
    try:
        from draco.core import containers
    except Exception:  # draco itself not importable: keep our container stand-ins
        containers = _compat
else:
    config = _compat.config
    _ContainerTask = _compat.ContainerTask
    _RandomTask = _compat.RandomTask
    containers = _compat

_WINDOWS = ["uniform", "hann", "hanning", "hamming", "blackman", "nuttall", "blackman_nuttall", "blackman_harris"]


def _invert_no_zero(x):
    x = np.asarray(x, dtype=np.float64)
    out = np.zeros_like(x)
    np.divide(1.0, x, out=out, where=x != 0)
    return out


# --------------------------------------------------------------------------- array helpers
def match_axes(dset1, dset2):
    """View of ``dset2`` with length-1 axes inserted where ``dset1`` has axes ``dset2`` lacks
    (delay.py:2209-2235)."""
    axes1, axes2 = list(dset1.attrs["axis"]), list(dset2.attrs["axis"])
    sel = tuple(slice(None) if ax in axes2 else np.newaxis for ax in axes1)
    arr = dset2[:]
    arr = getattr(arr, "local_array", arr)
    return arr[sel]


def flatten_axes(dset, axes_to_keep, match_dset=None):
    """Move ``axes_to_keep`` to the back, flatten all other axes into one leading ``baseline`` axis
    and distribute it over ranks in contiguous blocks (delay.py:2238-2302).  Returns the
    distributed array and the names of the flattened axes, slowest first."""
    data_axes = list(dset.attrs["axis"])
    for ax in axes_to_keep:
        if ax not in data_axes:
            raise ValueError(f"Specified axis={ax!r} not present in dataset.")
    arr = dset[:]
    comm = getattr(arr, "comm", None)
    if getattr(comm, "size", 1) > 1:
        # a real caput MPIArray distributed over several MPI ranks: its local block is NOT the global array this
        # stand-in splits.  The reference's own flatten_axes (delay.py:2278-2298) does the all-to-all; keep it.
        raise RuntimeError("draco_b200.delay.flatten_axes got an MPIArray distributed over several MPI ranks.  Under "
                           "caput + MPI run the reference's task classes with `draco_b200.install()` (only _evaluate / "
                           "_estimator are rebound; the redistribution stays caput's), not the stand-alone classes.")
    arr = np.asarray(getattr(arr, "local_array", arr))
    if match_dset is not None and tuple(data_axes) != tuple(match_dset.attrs["axis"]):
        full = match_dset[:]
        full = np.asarray(getattr(full, "local_array", full))
        arr = np.broadcast_to(match_axes(match_dset, dset), full.shape)
        data_axes = list(match_dset.attrs["axis"])
    keep = [data_axes.index(ax) for ax in axes_to_keep]
    other = [i for i in range(len(data_axes)) if i not in keep]
    arr = np.transpose(arr, other + keep)
    nflat = int(np.prod(arr.shape[: len(other)], dtype=np.int64)) if other else 1
    arr = arr.reshape((nflat, *arr.shape[len(other):]))
    return _compat.DistArray.from_global(arr, axis=0), [data_axes[i] for i in other]


# --------------------------------------------------------------------------- base classes
class DelayTransformBase(_ContainerTask):
    """Frequency -> delay base task: the parameter set of delay.py:403-429 and the host-side
    helpers ``_calculate_delays`` (delay.py:461-512) and ``_cut_data`` (delay.py:516-597)."""

    freq_zero = config.Property(proptype=float, default=None)
    freq_spacing = config.Property(proptype=float, default=None)
    nfreq = config.Property(proptype=int, default=None)
    skip_nyquist = config.Property(proptype=bool, default=True)
    apply_window = config.Property(proptype=bool, default=True)
    window = config.enum(_WINDOWS, default="nuttall")
    complex_timedomain = config.Property(proptype=bool, default=False)
    use_average_weights = config.Property(proptype=bool, default=True)
    weight_boost = config.Property(proptype=float, default=1.0)
    freq_frac = config.Property(proptype=float, default=0.0)
    time_frac = config.Property(proptype=float, default=0.0)
    remove_mean = config.Property(proptype=bool, default=True)
    scale_freq = config.Property(proptype=bool, default=False)

    # new, optional: defaults change nothing for an existing config
    device = config.Property(proptype=int, default=None)  # CUDA device; default = local rank
    batch_baselines = config.Property(proptype=int, default=0)  # 0 = all local baselines in one batch

    def process(self, ss):
        """Estimate the delay spectrum or power spectrum of ``ss`` (delay.py:431-459)."""
        delays, channel_ind = self._calculate_delays(ss)
        data_view, weight_view, coord_axes = self._prepare_inputs(ss)
        out_cont = self._create_output(ss, delays, coord_axes)
        out_cont.attrs["window_los"] = self.window if self.apply_window else "None"
        return self._evaluate(data_view, weight_view, out_cont, delays, channel_ind)

    def _device(self):
        """CUDA device of this rank: the ``device`` parameter, else local rank modulo the visible GPUs (the MPI
        communicator of a caput task when there is one, else torch.distributed / the launcher's environment)."""
        if self.device is not None:
            return int(self.device)
        import os

        import torch

        n = torch.cuda.device_count()
        for var in ("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "MV2_COMM_WORLD_LOCAL_RANK", "SLURM_LOCALID"):
            if os.environ.get(var, "").isdigit():
                return int(os.environ[var]) % n if n else 0
        comm = getattr(self, "comm", None)
        rank = comm.rank if hasattr(comm, "rank") else comm_rank_size()[0]
        return rank % n if n else 0

    def _calculate_delays(self, ss):
        if isinstance(ss, (list, tuple)):
            if not ss:
                raise TypeError("Could not find a frequency axis in the input.")
            ss = ss[0]
        freq = np.asarray(ss.freq)
        f0 = freq[0] if self.freq_zero is None else self.freq_zero
        spacing = np.abs(np.diff(freq)).min() if self.freq_spacing is None else self.freq_spacing
        if self.complex_timedomain:
            nchan = len(freq)
            channel_ind = np.arange(nchan)
            ndelay = nchan
        else:
            channel_ind = (np.abs(freq - f0) / spacing).astype(np.int64)
            nchan = self.nfreq
            if nchan is None:
                nchan = channel_ind[-1] + 1 + (1 if self.skip_nyquist else 0)
            ndelay = 2 * (nchan - 1)
        delays = np.fft.fftshift(np.fft.fftfreq(ndelay, d=spacing))
        return delays, channel_ind

    def _cut_data(self, data, weight):
        """Time/frequency cuts of one baseline; ``None`` for a dead baseline (three exits)."""
        ntime, nfreq = data.shape[-2:]
        positive = weight > 0
        if not positive.any():
            return None
        keep_t = positive.mean(axis=-1).reshape(-1, ntime).mean(axis=0) > self.time_frac
        positive = positive[..., keep_t, :]
        keep_f = positive.mean(axis=-2).reshape(-1, nfreq).mean(axis=0) > self.freq_frac
        if not keep_f.any():
            return None
        data = data[..., keep_t, :][..., keep_f]
        weight = weight[..., keep_t, :][..., keep_f]
        if self.remove_mean:
            data = data - data.mean(axis=0, keepdims=True)
        if (data == 0.0).all():
            return None
        if self.scale_freq:
            rel = data.std(axis=-2)[..., np.newaxis, :] / data.std(axis=(-1, -2))[..., np.newaxis, np.newaxis]
            data = data * _invert_no_zero(rel)
        if self.use_average_weights:
            weight = np.mean(weight, axis=-2)
        weight = weight * self.weight_boost
        return data, weight, keep_f, keep_t

    def _prepare_inputs(self, ss):
        raise NotImplementedError

    def _create_output(self, ss, delays, coord_axes):
        raise NotImplementedError

    def _evaluate(self, data_view, weight_view, out_cont, delays, channel_ind):
        raise NotImplementedError


class GeneralInputContainerMixin:
    """Collapse every axis except ``freq`` and ``sample_axis`` into ``baseline`` (delay.py:675-741)."""

    dataset = config.Property(proptype=str, default=None)
    sample_axis = config.Property(proptype=str, default=None)
    average_axis = config.Property(proptype=str, default=None)  # historical name of sample_axis

    def _sample_axis(self):
        return self.sample_axis if self.sample_axis is not None else self.average_axis

    def _prepare_inputs(self, ss):
        ss.redistribute("freq")
        if self.dataset is not None:
            if self.dataset not in ss.datasets:
                raise ValueError(f"Specified dataset to delay transform ({self.dataset}) not in "
                                 f"container of type {type(ss)}.")
            data_dset = ss[self.dataset]
        else:
            data_dset = ss.data
        sax = self._sample_axis()
        if sax not in ss.axes or sax not in list(data_dset.attrs["axis"]):
            raise ValueError(f"Specified sample axis ({sax}) not in container of type {type(ss)}.")
        data_view, bl_axes = flatten_axes(data_dset, [sax, "freq"])
        weight_view, _ = flatten_axes(ss.weight, [sax, "freq"], match_dset=data_dset)
        return data_view, weight_view, bl_axes


class DelayPowerSpectrumContainerMixin(GeneralInputContainerMixin):
    """Creates the ``DelaySpectrum`` output (delay.py:744-818)."""

    nsamp = config.Property(proptype=int, default=1)
    save_samples = config.Property(proptype=bool, default=False)
    save_spectrum_mask = config.Property(proptype=bool, default=False)

    def _create_output(self, ss, delays, coord_axes):
        if isinstance(coord_axes, np.ndarray):
            bl = coord_axes
        elif len(coord_axes) == 1:
            bl = ss.index_map[coord_axes[0]]
        else:
            bl = int(np.prod([len(ss.index_map[ax]) for ax in coord_axes]))
        out = containers.DelaySpectrum(baseline=bl, delay=delays, sample=self.nsamp, attrs_from=ss)
        out.redistribute("baseline")
        out.spectrum[:] = 0.0
        if isinstance(coord_axes, list):
            for ax in coord_axes:
                out.create_index_map(ax, ss.index_map[ax])
            out.attrs["baseline_axes"] = coord_axes
        if self.save_samples:
            out.add_dataset("spectrum_samples")
        if self.save_spectrum_mask:
            out.add_dataset("spectrum_mask")
            out.datasets["spectrum_mask"][:] = 0
        out.attrs["freq"] = ss.freq
        return out


class DelaySpectrumContainerMixin(GeneralInputContainerMixin):
    """Creates the ``DelayTransform`` output (delay.py:821-866)."""

    save_spectrum_mask = config.Property(proptype=bool, default=False)

    def _create_output(self, ss, delays, coord_axes):
        nbase = int(np.prod([len(ss.index_map[ax]) for ax in coord_axes]))
        out = containers.DelayTransform(baseline=nbase, sample=ss.index_map[self._sample_axis()], delay=delays,
                                        attrs_from=ss, weight_boost=self.weight_boost)
        out.redistribute("baseline")
        out.spectrum[:] = 0.0
        for ax in coord_axes:
            out.create_index_map(ax, ss.index_map[ax])
        out.attrs["baseline_axes"] = coord_axes
        if self.save_spectrum_mask:
            out.add_dataset("spectrum_mask")
            out.datasets["spectrum_mask"][:] = 0
        out.attrs["freq"] = ss.freq
        return out


def _local_array(view):
    return getattr(view, "local_array", view)


def _local(dset):
    """The local numpy slab of a (caput or stand-in) distributed dataset."""
    arr = dset[:]
    return getattr(arr, "local_array", arr)


class _BatchedEvaluate:
    """Shared host side of the batched device evaluation: cut every local baseline exactly as the
    reference loop does (delay.py:1146-1167), keep the survivors, remember the dead ones."""

    def _cut_all(self, data_view, weight_view, out_cont, channel_ind):
        jobs = []
        for lbi, bi in out_cont.spectrum[:].enumerate(axis=0):
            t = self._cut_data(data_view.local_array[lbi], weight_view.local_array[lbi])
            if t is None:
                if self.save_spectrum_mask:
                    _local(out_cont.datasets["spectrum_mask"])[lbi] = 1
                continue
            data, weight, nzf, nzt = t
            jobs.append((lbi, bi, data, weight, channel_ind[nzf], nzt))
        return jobs

    def _device_cut_serves(self, data_view):
        """complex64 slabs (draco's container dtype) are cut on the device in numpy's complex64 arithmetic, except
        under ``scale_freq`` (numpy's float32 std is not reproduced there): those go through the host ``_cut_data``."""
        return not (self.scale_freq and _local_array(data_view).dtype == np.complex64)

    def _chunks(self, jobs):
        step = self.batch_baselines if self.batch_baselines and self.batch_baselines > 0 else max(1, len(jobs))
        for i in range(0, len(jobs), step):
            yield jobs[i : i + step]


class DelayPowerSpectrumBase(_BatchedEvaluate, DelayPowerSpectrumContainerMixin, DelayTransformBase):
    """Delay power spectrum estimation base (delay.py:1114-1215)."""

    def _get_prior(self, nbase, ndelay, dtype):
        raise NotImplementedError

    def _estimator(self, data, weight, S, ndelay, channel_ind):
        raise NotImplementedError


class DelayPowerSpectrumGibbs(DelayPowerSpectrumBase, _RandomTask):
    """Gibbs-sampler estimate of the delay power spectrum; the spectrum returned is the median of
    the last ``median_frac`` of the chain (delay.py:1218-1267), computed on the GPU.

    Extra optional parameter ``rng_mode``: ``"philox"`` (default) draws on the device with
    Philox4x32-10 keyed on (seed, global baseline index, iteration), so results do not depend on
    the number of ranks; ``"injected"`` draws every Gaussian / chi^2 variate from ``self.rng`` on
    the host in the reference's order and ships them to the device (slow, meant for validation).  All ``nsamp``
    iterations of a baseline are drawn up front, so the host generator's stream equals the reference's only while
    every chain runs to completion: after a chain that stops early (delay.py:1891-1896) the reference has consumed
    fewer variates than this task, and the baselines that follow see different draws.
    """

    initial_amplitude = config.Property(proptype=float, default=10.0)
    median_frac = config.Property(proptype=float, default=0.5)
    rng_mode = config.enum(["philox", "injected"], default="philox")
    device_cut = config.Property(proptype=bool, default=True)

    _ncalls = 0

    def _get_prior(self, nbase, ndelay, dtype):
        return np.ones((nbase, ndelay), dtype=dtype) * self.initial_amplitude

    def _window(self):
        return self.window if self.apply_window else None

    def _estimator(self, data, weight, S, ndelay, channel_ind):
        """Single-baseline form kept for API parity with delay.py:1242-1267."""
        samples, success = gibbs.delay_power_spectrum_gibbs(
            data, ndelay, weight, S, window=self._window(), fsel=channel_ind, niter=self.nsamp,
            rng=self.rng if self.rng_mode == "injected" else None, complex_timedomain=self.complex_timedomain,
            device=self._device(), seed=self._device_seed())
        ntail = int(self.nsamp * self.median_frac)
        if samples:
            spec = np.fft.fftshift(np.median(samples[-ntail:], axis=0))
        else:
            success = False
            spec = S
        return spec, samples, success

    def _device_seed(self):
        """One 64-bit Philox seed shared by all ranks (results independent of the sharding)."""
        from . import dist as ddist

        if getattr(self, "seed", None) is not None:
            base = self.seed
        else:
            base = int(np.random.SeedSequence().generate_state(1, dtype=np.uint64)[0] >> 1)
            comm = getattr(self, "comm", None)
            if hasattr(comm, "bcast"):  # a caput task under MPI
                base = comm.bcast(base, root=0)
            else:
                base = ddist.broadcast_int(base, device=self._device())
        self._ncalls += 1
        return int(np.random.SeedSequence([int(base), self._ncalls]).generate_state(1, dtype=np.uint64)[0])

    def _evaluate(self, data_view, weight_view, out_cont, delays, channel_ind):
        """Batched replacement of delay.py:1117-1185."""
        from . import dist as ddist

        if not self.use_average_weights:
            # the reference sampler cannot work with per-sample weights either (delay.py:1697)
            raise ValueError("DelayPowerSpectrumGibbs needs use_average_weights=True")
        nbase = out_cont.spectrum.global_shape[0]
        ndelay = len(delays)
        prior = self._get_prior(nbase, ndelay, delays.dtype)
        spec_out = _local(out_cont.spectrum)
        mask_out = _local(out_cont.datasets["spectrum_mask"]) if self.save_spectrum_mask else None
        samp_out = _local(out_cont.datasets["spectrum_samples"]) if self.save_samples else None
        handle = gibbs.get_handle(self._device())
        seed = self._device_seed()
        if self.device_cut and self.rng_mode == "philox" and self._device_cut_serves(data_view):
            return self._evaluate_device_cut(data_view, weight_view, out_cont, channel_ind, ndelay, prior, handle, seed)
        jobs = self._cut_all(data_view, weight_view, out_cont, channel_ind)
        nsys = (2 if self.complex_timedomain else 1) * ndelay
        for chunk in self._chunks(jobs):
            lbis = [j[0] for j in chunk]
            pb = gibbs.pack_batch([j[2] for j in chunk], [np.asarray(j[3], dtype=np.float64) for j in chunk],
                                  [j[4] for j in chunk], ndelay, prior[lbis], window=self._window(),
                                  complex_timedomain=self.complex_timedomain,
                                  global_id=np.array([j[1] for j in chunk], dtype=np.int64))
            draws = None
            if self.rng_mode == "injected":
                per = [gibbs._host_draws(self.rng, self.nsamp, nsys, j[2].shape[1], j[2].shape[0], ndelay) for j in chunk]
                draws = gibbs.pack_draws(pb, self.nsamp, *zip(*per))
            res = gibbs.run_batch(pb, self.nsamp, handle=handle, seed=seed, median_frac=self.median_frac, draws=draws,
                                  save_samples=self.save_samples)
            for n, (lbi, bi, *_rest) in enumerate(chunk):
                spec_out[lbi] = res.spectrum[n]
                if mask_out is not None and not res.success[n]:
                    mask_out[lbi] = 1
                if samp_out is not None:
                    samp_out[:, lbi] = res.samples[:, n]  # already front-padded with zeros (delay.py:1177-1178)
            self.log.debug(f"device batch of {len(chunk)} baselines: {res.timing['ms_sampler']:.1f} ms in the sampler")
        if self.save_spectrum_mask:
            n_conv = nbase - ddist.allreduce_count(int(mask_out.sum()), device=self._device())  # delay.py:1182
            self.log.debug(f"{n_conv}/{nbase} unflagged baselines.")
        return out_cont


    def _evaluate_device_cut(self, data_view, weight_view, out_cont, channel_ind, ndelay, prior, handle, seed):
        """Same as the loop above with ``_cut_data`` done on the device for the whole local slab
        (``dgibbs_cut_data``, delay.py:516-597): the raw [B, T, nfreq] slab goes up once, the cut / mean-subtracted /
        compacted batch stays in HBM and feeds the sampler in place."""
        from . import dist as ddist

        # no host-side ascontiguousarray: cut_batch ships the slab as it lies in memory and permutes on the device
        vis = np.asarray(_local_array(data_view))
        wgt = np.asarray(_local_array(weight_view))
        spec_out = _local(out_cont.spectrum)
        mask_out = _local(out_cont.datasets["spectrum_mask"]) if self.save_spectrum_mask else None
        samp_out = _local(out_cont.datasets["spectrum_samples"]) if self.save_samples else None
        gids = np.array([bi for _, bi in out_cont.spectrum[:].enumerate(axis=0)], dtype=np.int64)
        nloc = vis.shape[0]
        step = self.batch_baselines if self.batch_baselines and self.batch_baselines > 0 else max(1, nloc)
        for s0 in range(0, nloc, step):
            s1 = min(nloc, s0 + step)
            cr = gibbs.cut_batch(vis[s0:s1], wgt[s0:s1], channel_ind, ndelay, prior[s0:s1], time_frac=self.time_frac,
                                 freq_frac=self.freq_frac, remove_mean=self.remove_mean, scale_freq=self.scale_freq,
                                 weight_boost=self.weight_boost, window=self._window(),
                                 complex_timedomain=self.complex_timedomain, global_id=gids[s0:s1], handle=handle)
            if mask_out is not None:
                mask_out[s0:s1][~cr.alive] = 1
            if cr.batch is None:
                continue
            res = gibbs.run_batch(cr.batch, self.nsamp, handle=handle, seed=seed, median_frac=self.median_frac,
                                  save_samples=self.save_samples)
            lbis = s0 + cr.index
            spec_out[lbis] = res.spectrum.cpu().numpy()
            if mask_out is not None:
                mask_out[lbis[~res.success.cpu().numpy()]] = 1
            if samp_out is not None:
                samp_out[:, lbis] = res.samples.cpu().numpy()
            self.log.debug(f"device batch of {len(lbis)} baselines: {res.timing['ms_sampler']:.1f} ms in the sampler")
        if self.save_spectrum_mask:
            nbase = out_cont.spectrum.global_shape[0]
            n_conv = nbase - ddist.allreduce_count(int(mask_out.sum()), device=self._device())  # delay.py:1182
            self.log.debug(f"{n_conv}/{nbase} unflagged baselines.")
        return out_cont


class DelayPowerSpectrumNRML(DelayPowerSpectrumBase):
    """NRML / maximum-posterior estimate of the delay power spectrum (delay.py:1270-1301): the spectrum is the last
    iterate of ``delay_power_spectrum_maxpost`` (delayopt.py:424-561).  The per-baseline loop and scipy's Newton-CG
    stay on the host as in the reference (delay.py:1146-1178); every likelihood / gradient / Hessian evaluation runs
    on the GPU (``draco_b200.delayopt.LogLikePS`` -> ``libdgnrml.so``)."""

    maxpost_tol = config.Property(proptype=float, default=1e-3)

    def _get_prior(self, nbase, ndelay, dtype):
        return [None] * nbase

    def _estimator(self, data, weight, S, ndelay, channel_ind):
        from . import delayopt

        samples, success = delayopt.delay_power_spectrum_maxpost(
            data, ndelay, weight, S, window=self.window if self.apply_window else None, fsel=channel_ind,
            maxiter=self.nsamp, tol=self.maxpost_tol, device=self._device())
        return np.fft.fftshift(samples[-1]), samples, success

    def _evaluate(self, data_view, weight_view, out_cont, delays, channel_ind):
        from . import dist as ddist

        nbase = out_cont.spectrum.global_shape[0]
        ndelay = len(delays)
        prior = self._get_prior(nbase, ndelay, delays.dtype)
        spec_out = _local(out_cont.spectrum)
        mask_out = _local(out_cont.datasets["spectrum_mask"]) if self.save_spectrum_mask else None
        samp_out = _local(out_cont.datasets["spectrum_samples"]) if self.save_samples else None
        for lbi, bi, data, weight, fsel, _nzt in self._cut_all(data_view, weight_view, out_cont, channel_ind):
            spec, samples, success = self._estimator(data, np.asarray(weight, dtype=np.float64), prior[lbi], ndelay, fsel)
            spec_out[lbi] = spec
            if mask_out is not None and not success:
                mask_out[lbi] = 1
            if samp_out is not None:
                # delay.py:1175-1178; a run that uses all `nsamp` iterations returns nsamp + 1 iterates (the start
                # included), which the reference cannot store (broadcast error): keep the last `nsamp` of them
                arr = np.array(samples)[-samp_out.shape[0]:]
                samp_out[:, lbi] = 0.0
                samp_out[-len(arr):, lbi] = arr
        if self.save_spectrum_mask:
            n_conv = nbase - ddist.allreduce_count(int(mask_out.sum()), device=self._device())  # delay.py:1182
            self.log.debug(f"{n_conv}/{nbase} unflagged baselines.")
        return out_cont


class DelayCrossPowerSpectrumEstimator(DelayPowerSpectrumGibbs):
    """Delay cross power spectrum of several compatible ``FreqContainer`` inputs (delay.py:1304-1447):
    ``process(sslist)`` returns a ``DelayCrossSpectrum`` with the pair-wise cross spectra.  Every local
    baseline's chain (``delay_spectrum_gibbs_cross``, delay.py:1907-2099) runs on the GPU in one batch
    per group of baselines that share a channel selection.

    Two defects of the reference wrapper that stop it from running are not reproduced: the frequency
    check ``(ss.freq != freq_ref).all()`` (delay.py:1328; any mismatch must raise) and the call
    ``super()._prepare_inputs(self, ss)`` (delay.py:1330; ``self`` passed twice).  Like the reference, the
    output datasets are float64 (containers.py:2493-2506): the real part of the median is stored; and like
    the reference the median is passed through ``np.fft.fftshift`` with no ``axes`` argument (delay.py:1442),
    which also rolls the two dataset axes.
    """

    def _prepare_inputs(self, sslist):
        if len(sslist) == 0:
            raise ValueError("No datasets passed.")
        freq_ref = np.asarray(sslist[0].freq)
        data_views, weight_views, coord_axes = [], [], None
        for ss in sslist:
            ss.redistribute("freq")
            f = np.asarray(ss.freq)
            if f.shape != freq_ref.shape or (f != freq_ref).any():
                raise ValueError("Input containers must have the same frequencies.")
            dv, wv, ca = GeneralInputContainerMixin._prepare_inputs(self, ss)
            if coord_axes is not None and not coord_axes == ca:
                raise ValueError("Different axes found for the input containers.")
            data_views.append(dv)
            weight_views.append(wv)
            coord_axes = ca
        return data_views, weight_views, coord_axes

    def _create_output(self, ss, delays, coord_axes):
        ssref, ndata = ss[0], len(ss)
        if len(coord_axes) == 1:
            bl = ssref.index_map[coord_axes[0]]
        else:
            bl = int(np.prod([len(ssref.index_map[ax]) for ax in coord_axes]))
        out = containers.DelayCrossSpectrum(baseline=bl, dataset=ndata, delay=delays, sample=self.nsamp, attrs_from=ssref)
        out.redistribute("baseline")
        out.spectrum[:] = 0.0
        if isinstance(coord_axes, list):
            for ax in coord_axes:
                out.create_index_map(ax, ssref.index_map[ax])
            out.attrs["baseline_axes"] = coord_axes
        if self.save_samples:
            out.add_dataset("spectrum_samples")
        out.attrs["freq"] = ssref.freq
        return out

    def _evaluate(self, data_view, weight_view, out_cont, delays, channel_ind):
        """Batched replacement of the loop delay.py:1421-1445."""
        ndata, ndelay = len(data_view), len(delays)
        nbase = out_cont.spectrum.global_shape[-2]
        initial_S = self._get_prior(nbase, ndelay, delays.dtype)
        if initial_S.ndim == 2:  # delay.py:1402-1407
            initial_S = np.identity(ndata)[np.newaxis, ..., np.newaxis] * initial_S[:, np.newaxis, np.newaxis]
        elif initial_S.ndim != 4 or initial_S.shape[1] != ndata:
            raise ValueError(f"Expected an initial sample with dimension 4 and {ndata} datasets. "
                             f"Got sample with dimension {initial_S.ndim} and shape {initial_S.shape}.")
        spec_out = _local(out_cont.spectrum)
        samp_out = _local(out_cont.datasets["spectrum_samples"]) if self.save_samples else None
        # cut every local baseline, group by (channel selection, sample count): one device batch per group
        groups = {}
        for lbi, bi in out_cont.spectrum[:].enumerate(axis=-2):
            data = np.array([_local_array(d)[lbi] for d in data_view])
            weight = np.array([_local_array(w)[lbi] for w in weight_view])
            t = self._cut_data(data, weight)
            if t is None:
                continue
            data, weight, nzf, _ = t
            key = (nzf.tobytes(), data.shape[1])
            groups.setdefault(key, []).append((lbi, bi, data, weight, channel_ind[nzf]))
        handle = gibbs.get_handle(self._device())
        seed = self._device_seed()
        for jobs in groups.values():
            data = np.stack([j[2] for j in jobs])
            T, M = data.shape[2], data.shape[3]
            draws = None
            # one generator shared by the local baselines (delay.py:1418); drawn group by group, which is the
            # reference's loop order only when every local baseline has the same channel selection and sample count
            if self.rng_mode == "injected":
                per = [gibbs._host_cross_draws(self.rng, self.nsamp, ndelay, ndata, M, T) for _ in jobs]
                draws = tuple(np.stack([p[i] for p in per]) for i in range(4))
            samples, n_done, ok = gibbs.cross_batch(
                data, ndelay, np.stack([j[3] for j in jobs]), initial_S[[j[0] for j in jobs]],
                window=self._window(), fsel=jobs[0][4], niter=self.nsamp, draws=draws, seed=seed,
                global_id=np.array([j[1] for j in jobs], dtype=np.int64), handle=handle)
            if not ok.all():  # delay.py:2096-2097
                raise RuntimeError("Exiting earlier as singular")
            for n, (lbi, bi, *_r) in enumerate(jobs):
                spec = samples[n]
                spec_av = np.median(spec[-(self.nsamp // 2):], axis=0)  # delay.py:1438-1441
                spec_out[..., lbi, :] = np.fft.fftshift(spec_av).real  # all axes, as delay.py:1442
                if samp_out is not None:
                    samp_out[..., lbi, :] = spec.real
        return out_cont


# historical task names (BASELINE.json / old YAML files); see SURVEY.md section 0.3
DelayPowerSpectrumStochasticEstimator = DelayPowerSpectrumGibbs
DelaySpectrumEstimator = DelayPowerSpectrumGibbs


class DelaySpectrumBase(_BatchedEvaluate, DelaySpectrumContainerMixin, DelayTransformBase):
    """Delay spectrum estimation base (delay.py:874-957)."""


class DelaySpectrumWienerFilter(DelaySpectrumBase):
    """Wiener-filtered delay spectrum given a delay power spectrum (delay.py:982-1024), one
    deterministic device solve per baseline with the sampler's factorisation kernels."""

    def setup(self, dps=None):
        self.dps = dps
        super().setup()

    def _get_prior(self, nbase):
        return _local(self.dps.spectrum)

    def _estimator(self, data, weight, S, ndelay, channel_ind):
        y = gibbs.delay_spectrum_wiener_filter(np.fft.fftshift(S), data, ndelay, weight,
                                               window=self.window if self.apply_window else None, fsel=channel_ind,
                                               complex_timedomain=self.complex_timedomain, device=self._device())
        return np.fft.fftshift(y, axes=-1)

    device_cut = config.Property(proptype=bool, default=True)

    def _evaluate(self, data_view, weight_view, out_cont, delays, channel_ind):
        nbase = out_cont.spectrum.global_shape[0]
        ndelay = len(delays)
        prior = self._get_prior(nbase)
        spec_out = _local(out_cont.spectrum)
        mask_out = _local(out_cont.datasets["spectrum_mask"]) if self.save_spectrum_mask else None
        handle = gibbs.get_handle(self._device())
        window = self.window if self.apply_window else None

        def store(lbi, bi, y, ok, nzt):
            if not ok:
                raise np.linalg.LinAlgError(f"Wiener covariance of baseline {bi} is not positive definite")
            y = y.view(np.complex128) if self.complex_timedomain else y
            spec_out[lbi, nzt] = np.fft.fftshift(y, axes=-1)
            if mask_out is not None:
                mask_out[lbi][~nzt] = 1

        if self.device_cut and self.use_average_weights and self._device_cut_serves(data_view):
            # _cut_data on the device for the whole local slab (dgibbs_cut_data); the batch stays in HBM
            vis = np.asarray(_local_array(data_view))
            wgt = np.asarray(_local_array(weight_view))
            gids = np.array([bi for _, bi in out_cont.spectrum[:].enumerate(axis=0)], dtype=np.int64)
            nloc = vis.shape[0]
            step = self.batch_baselines if self.batch_baselines and self.batch_baselines > 0 else max(1, nloc)
            for s0 in range(0, nloc, step):
                s1 = min(nloc, s0 + step)
                S = np.fft.fftshift(np.asarray(prior[s0:s1]), axes=-1)
                cr = gibbs.cut_batch(vis[s0:s1], wgt[s0:s1], channel_ind, ndelay, S, time_frac=self.time_frac,
                                     freq_frac=self.freq_frac, remove_mean=self.remove_mean, scale_freq=self.scale_freq,
                                     weight_boost=self.weight_boost, window=window,
                                     complex_timedomain=self.complex_timedomain, global_id=gids[s0:s1], handle=handle)
                if mask_out is not None and self.save_spectrum_mask:
                    for lb in np.flatnonzero(~cr.alive):
                        mask_out[s0 + lb] = 1  # delay.py:1160-1162 (a skipped baseline flags all its samples)
                if cr.batch is None:
                    continue
                ys, ok = gibbs.wiener_batch(cr.batch, handle=handle)
                for n, lb in enumerate(cr.index):
                    store(s0 + lb, gids[s0 + lb], ys[n], ok[n], cr.keep_t[lb])
            return out_cont
        jobs = self._cut_all(data_view, weight_view, out_cont, channel_ind)
        for chunk in self._chunks(jobs):
            S = np.stack([np.fft.fftshift(prior[j[0]]) for j in chunk])
            pb = gibbs.pack_batch([j[2] for j in chunk], [np.asarray(j[3], dtype=np.float64) for j in chunk],
                                  [j[4] for j in chunk], ndelay, S, window=window,
                                  complex_timedomain=self.complex_timedomain)
            ys, ok = gibbs.wiener_batch(pb, handle=handle)
            for n, (lbi, bi, _d, _w, _f, nzt) in enumerate(chunk):
                store(lbi, bi, ys[n], ok[n], nzt)
        return out_cont


class DelaySpectrumFFT(DelaySpectrumBase):
    """Delay spectrum of a general container via a windowed inverse FFT (delay.py:960-980 over
    ``delay_spectrum_fft``, delay.py:2102-2130).  As in the reference, the transform ignores masking and needs the full
    channel set (``complex_timedomain: true`` and no frequency cut: the window has one entry per delay).  All local
    baselines go through ONE device call (``dgibbs_delay_fft``: window, inverse FFT, 1/N and the fftshift fused)."""

    device_cut = config.Property(proptype=bool, default=True)

    def _get_prior(self, nbase):
        return [None] * nbase

    def _estimator(self, data, weight, S, ndelay, channel_ind):
        y_spec = gibbs.delay_spectrum_fft(data, ndelay, self.window if self.apply_window else None, device=self._device())
        return np.fft.fftshift(y_spec, axes=-1)

    @staticmethod
    def _check_width(nchan, ndelay):
        if nchan != ndelay:  # the reference's `data *= window` (delay.py:2127) fails to broadcast
            raise ValueError(f"operands could not be broadcast together with shapes (nsample,{nchan}) (1,{ndelay})")

    def _evaluate(self, data_view, weight_view, out_cont, delays, channel_ind):
        """Batched replacement of the loop delay.py:899-930."""
        ndelay = len(delays)
        spec_out = _local(out_cont.spectrum)
        mask_out = _local(out_cont.datasets["spectrum_mask"]) if self.save_spectrum_mask else None
        handle = gibbs.get_handle(self._device())
        window = self.window if self.apply_window else None
        if self.device_cut and self._device_cut_serves(data_view):
            vis = np.asarray(_local_array(data_view))
            wgt = np.asarray(_local_array(weight_view))
            nloc, T, F = vis.shape
            step = self.batch_baselines if self.batch_baselines and self.batch_baselines > 0 else max(1, nloc)
            for s0 in range(0, nloc, step):
                s1 = min(nloc, s0 + step)
                cr = gibbs.cut_batch(vis[s0:s1], wgt[s0:s1], channel_ind, ndelay, 1.0, time_frac=self.time_frac,
                                     freq_frac=self.freq_frac, remove_mean=self.remove_mean, scale_freq=self.scale_freq,
                                     weight_boost=self.weight_boost, window=window,
                                     complex_timedomain=self.complex_timedomain, handle=handle)
                if mask_out is not None:
                    mask_out[s0:s1][~cr.alive] = 1  # delay.py:917-920
                if cr.batch is None:
                    continue
                for m in cr.batch.M.cpu().numpy():
                    self._check_width(int(m), ndelay)
                self._check_width(F, ndelay)
                # every surviving baseline kept all F = ndelay channels: its kept samples are the leading rows of its
                # [T, F] slot of the cut slab, which is transformed as it lies in HBM
                y = gibbs.fft_rows(cr.batch.data.view(-1, F, 2), window, shift=True, handle=handle)
                y = y.cpu().numpy().view(np.complex128).reshape(s1 - s0, T, F)
                for lb in cr.index:
                    nzt = cr.keep_t[lb]
                    spec_out[s0 + lb, nzt] = y[lb, : int(nzt.sum())]
                    if mask_out is not None:
                        mask_out[s0 + lb][~nzt] = 1  # delay.py:929-930
            return out_cont
        jobs = self._cut_all(data_view, weight_view, out_cont, channel_ind)
        for chunk in self._chunks(jobs):
            for j in chunk:
                self._check_width(j[2].shape[1], ndelay)
            rows = np.concatenate([np.asarray(j[2], dtype=np.complex128) for j in chunk], axis=0)
            y = gibbs.fft_rows(rows, window, shift=True, handle=handle)
            o = 0
            for lbi, bi, d, _w, _f, nzt in chunk:
                spec_out[lbi, nzt] = y[o : o + d.shape[0]]
                o += d.shape[0]
                if mask_out is not None:
                    mask_out[lbi][~nzt] = 1
        return out_cont


class DelaySpectrumToPowerSpectrum(_ContainerTask):
    """Delay power spectrum = variance over the sample axis of a delay spectrum (delay.py:1061-1106), one device
    reduction (``dgibbs_sample_variance``) over the local ``[baseline, sample, delay]`` slab."""

    device = config.Property(proptype=int, default=None)

    def process(self, dspec):
        dspec.redistribute("baseline")
        pspec = containers.DelaySpectrum(attrs_from=dspec, axes_from=dspec)
        pspec.redistribute("baseline")
        mask = None
        if "spectrum_mask" in dspec.datasets:
            mask = _local(dspec.datasets["spectrum_mask"])
            pspec.add_dataset("spectrum_mask")
            pspec.datasets["spectrum_mask"][:] = 0
        dev = self.device
        if dev is None:
            import torch

            dev = comm_rank_size()[0] % max(1, torch.cuda.device_count())
        var, bad = gibbs.sample_variance(_local(dspec.spectrum), mask, device=dev)
        _local(pspec.spectrum)[:] = var
        if mask is not None:
            _local(pspec.datasets["spectrum_mask"])[:] = bad
        return pspec


class DelaySpectrumWienerFilterIteratePS(DelaySpectrumWienerFilter):
    """Variant whose power spectrum arrives with every ``process`` call (delay.py:1027-1053)."""

    def process(self, ss, dps):
        self.dps = dps
        return super().process(ss)


class DelayPowerSpectrumStokesIEstimator(DelayPowerSpectrumGibbs):
    """Deprecated (delay.py:1451-1461)."""

    def setup(self, requires=None):
        raise DeprecationWarning("`DelayPowerSpectrumStokesIEstimator` is deprecated. Use `draco.transform.StokesI` to "
                                 "generate Stokes I visibilities, then use `DelayPowerSpectrumGibbs` or "
                                 "`DelayPowerSpectrumNRML`.")


class DelayPowerSpectrumGeneralEstimator(DelayPowerSpectrumGibbs):
    """Deprecated (delay.py:1464-1472)."""

    def setup(self, requires=None):
        raise DeprecationWarning("`DelayPowerSpectrumGeneralEstimator` is deprecated. Use `DelayPowerSpectrumGibbs` or "
                                 "`DelayPowerSpectrumNRML`.")
