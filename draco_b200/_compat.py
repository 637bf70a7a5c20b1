"""Minimal in-repo stand-ins for the caput / draco symbols the delay tasks touch.

caput, mpi4py and h5py are not installed in the build container nor on the GPU box (probed,
SURVEY.md section 0/8b), so the task classes in ``draco_b200.delay`` are written against
caput's API and fall back to these ~15 symbols when ``import caput`` fails.  Only the surface
used by ``draco/analysis/delay.py:347-866, 1114-1267`` is provided:

* ``config.Property`` / ``config.enum`` descriptors (class attributes filled from YAML ``params``),
* ``ContainerTask`` (``.log``, ``setup`` / ``process``, ``read_config``) and ``RandomTask`` (``.rng``),
* ``DistArray``: the slice of ``caput.mpiarray.MPIArray`` the hot path uses (``local_array``,
  ``global_shape``, ``local_shape``, ``enumerate(axis)``, distributed axis 0 blocks),
* ``Container`` / ``FreqContainer`` / ``SiderealStream`` / ``DelaySpectrum`` / ``DelayTransform`` with
  ``index_map``, ``attrs``, ``datasets``, ``axes``, dataset ``attrs['axis']`` (the output layout contract
  of ``draco/core/containers.py:2049-2182``),
* ``RingMap`` / ``DelayTransformOperator``: the datasets ``draco/analysis/powerspec.py:118-320`` reads and writes.

Distribution mirrors caput: the flattened baseline axis is split in contiguous balanced blocks,
one per rank; ranks come from ``torch.distributed`` when it is initialised (NCCL on GPUs, gloo in
the CPU tests), else a single rank.
"""

from __future__ import annotations

import logging

import numpy as np

from .dist import split_range

HAVE_CAPUT = False
try:  # pragma: no cover - not available in this image
    import caput  # noqa: F401

    HAVE_CAPUT = True
except Exception:
    HAVE_CAPUT = False


def comm_rank_size():
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except Exception:
        pass
    return 0, 1


# --------------------------------------------------------------------------- config
class _NoDefault:
    pass


class Property:
    """``caput.config.Property``: a typed class attribute filled from a config dict."""

    def __init__(self, default=_NoDefault, proptype=None, key=None):
        self.default = default
        self.proptype = (lambda x: x) if proptype is None else proptype
        self.key = key
        self.name = None

    def __set_name__(self, owner, name):
        self.name = name
        if self.key is None:
            self.key = name

    def __get__(self, obj, objtype=None):
        if obj is None:
            return self
        if self.name in obj.__dict__:
            return obj.__dict__[self.name]
        if self.default is _NoDefault:
            return None
        return self.default

    def __set__(self, obj, value):
        obj.__dict__[self.name] = None if value is None else self.proptype(value)


def enum(options, default=None):
    def check(val):
        if val not in options:
            raise ValueError(f"Input {val!r} is not one of the allowed values {options}")
        return val

    return Property(default=default, proptype=check)


class config:  # namespace mirroring ``from caput import config``
    Property = Property
    enum = staticmethod(enum)


class Reader:
    """``caput.config.Reader``: ``read_config`` / ``from_config`` over ``Property`` attributes."""

    @classmethod
    def _properties(cls):
        props = {}
        for klass in reversed(cls.__mro__):
            for k, v in vars(klass).items():
                if isinstance(v, Property):
                    props[v.key] = k
        return props

    def read_config(self, params, strict=False):
        props = self._properties()
        for key, value in (params or {}).items():
            if key in props:
                setattr(self, props[key], value)
            elif strict:
                raise ValueError(f"unknown config parameter {key!r} for {type(self).__name__}")
        return self

    @classmethod
    def from_config(cls, params=None, **kwargs):
        obj = cls()
        obj.read_config({**(params or {}), **kwargs}, strict=True)
        return obj


class ContainerTask(Reader):
    """``caput.pipeline.tasklib.base.ContainerTask`` (the part the delay tasks use)."""

    def __init__(self):
        self.log = logging.getLogger(f"{type(self).__module__}.{type(self).__name__}")

    def setup(self, *args, **kwargs):
        pass

    def process(self, *args, **kwargs):
        raise NotImplementedError


class RandomTask(ContainerTask):
    """``caput.pipeline.tasklib.random.RandomTask``: provides ``self.rng``.  The per-rank seeding
    rule of caput is not recoverable from the reference tree; here rank r uses
    ``SeedSequence(seed).spawn`` child r, and ``seed=None`` takes OS entropy."""

    seed = Property(proptype=int, default=None)
    _rng = None

    @property
    def rng(self):
        if self._rng is None:
            rank, size = comm_rank_size()
            ss = np.random.SeedSequence(self.seed)
            self._rng = np.random.default_rng(ss.spawn(size)[rank] if size > 1 else ss)
        return self._rng


# --------------------------------------------------------------------------- arrays / containers
class DistArray:
    """Block-distributed array (axis 0), the used slice of ``caput.mpiarray.MPIArray``."""

    def __init__(self, local, global_len=None, offset=0, axis=0):
        self.local_array = np.asarray(local)
        self.axis = axis
        self.local_offset = int(offset)
        n0 = self.local_array.shape[axis] if global_len is None else int(global_len)
        gs = list(self.local_array.shape)
        gs[axis] = n0
        self.global_shape = tuple(gs)

    @classmethod
    def from_global(cls, arr, axis=0):
        """Take this rank's contiguous block of a replicated global array."""
        rank, size = comm_rank_size()
        a, b = split_range(arr.shape[axis], size, rank)
        sl = [slice(None)] * arr.ndim
        sl[axis] = slice(a, b)
        return cls(arr[tuple(sl)], arr.shape[axis], a, axis)

    @property
    def local_shape(self):
        return self.local_array.shape

    @property
    def shape(self):
        return self.local_array.shape

    @property
    def dtype(self):
        return self.local_array.dtype

    def enumerate(self, axis=0):
        """(local index, global index) pairs along ``axis`` (``MPIArray.enumerate``)."""
        axis = axis % self.local_array.ndim
        n = self.local_array.shape[axis]
        off = self.local_offset if axis == self.axis % self.local_array.ndim else 0
        return [(i, off + i) for i in range(n)]

    def __getitem__(self, key):
        if isinstance(key, slice) and key == slice(None):
            return self
        return self.local_array[key]

    def __setitem__(self, key, value):
        self.local_array[key] = value

    def sum(self):
        return self.local_array.sum()


class Dataset:
    """A dataset: ``[:]`` gives the (distributed) array, ``attrs['axis']`` names the axes."""

    def __init__(self, data, axes):
        self.data = data
        self.attrs = {"axis": np.array(axes)}

    def __getitem__(self, key):
        if isinstance(key, slice) and key == slice(None):
            return self.data
        target = self.data.local_array if isinstance(self.data, DistArray) else self.data
        return target[key]

    def __setitem__(self, key, value):
        target = self.data.local_array if isinstance(self.data, DistArray) else self.data
        target[key] = value

    @property
    def local_array(self):
        return self.data.local_array if isinstance(self.data, DistArray) else self.data

    @property
    def global_shape(self):
        return self.data.global_shape if isinstance(self.data, DistArray) else self.data.shape

    @property
    def local_shape(self):
        return self.local_array.shape


class Container:
    """Bare-bones ``ContainerPrototype``: ``index_map``, ``attrs``, ``datasets``."""

    _axes = ()
    _dataset_spec: dict = {}

    def __init__(self, attrs_from=None, **axes):
        self.index_map = {}
        self.attrs = {}
        self.datasets = {}
        if attrs_from is not None:
            self.attrs.update(getattr(attrs_from, "attrs", {}))
        for name, val in axes.items():
            self.index_map[name] = np.arange(val) if np.isscalar(val) else np.asarray(val)

    @property
    def axes(self):
        return tuple(self.index_map)

    def __getitem__(self, name):
        return self.datasets[name]

    def create_index_map(self, name, value):
        self.index_map[name] = np.asarray(value)

    def redistribute(self, axis):
        return self

    def _make(self, name, dist_axis=None):
        spec = self._dataset_spec[name]
        shape = [len(self.index_map[a]) for a in spec["axes"]]
        if dist_axis is not None and dist_axis in spec["axes"]:
            ax = spec["axes"].index(dist_axis)
            rank, size = comm_rank_size()
            a, b = split_range(shape[ax], size, rank)
            lshape = list(shape)
            lshape[ax] = b - a
            arr = DistArray(np.zeros(lshape, dtype=spec["dtype"]), shape[ax], a, ax)
        else:
            arr = np.zeros(shape, dtype=spec["dtype"])
        self.datasets[name] = Dataset(arr, spec["axes"])
        return self.datasets[name]

    def add_dataset(self, name):
        return self._make(name, self._dataset_spec[name].get("distributed_axis"))


class FreqContainer(Container):
    @property
    def freq(self):
        fm = self.index_map["freq"]
        try:
            return fm[:]["centre"]
        except (IndexError, ValueError):
            return fm[:]


class SiderealStream(FreqContainer):
    """Visibility stack ``vis[freq, stack, ra]`` + ``vis_weight`` (draco/core/containers.py)."""

    _dataset_spec = {
        "vis": {"axes": ["freq", "stack", "ra"], "dtype": np.complex128, "distributed_axis": "freq"},
        "vis_weight": {"axes": ["freq", "stack", "ra"], "dtype": np.float64, "distributed_axis": "freq"},
    }

    def __init__(self, freq, stack, ra, attrs_from=None, native_dtypes=False):
        super().__init__(attrs_from=attrs_from, freq=freq, stack=stack, ra=ra)
        if native_dtypes:  # what draco's own container stores (core/containers.py:503-520): complex64 / float32
            self._dataset_spec = {"vis": {**self._dataset_spec["vis"], "dtype": np.complex64},
                                  "vis_weight": {**self._dataset_spec["vis_weight"], "dtype": np.float32}}
        # the stand-in keeps inputs replicated; the baseline split happens in flatten_axes
        self._make("vis")
        self._make("vis_weight")

    @property
    def vis(self):
        return self.datasets["vis"]

    @property
    def data(self):
        return self.datasets["vis"]

    @property
    def weight(self):
        return self.datasets["vis_weight"]


class DelaySpectrum(Container):
    """Delay power spectrum container (draco/core/containers.py:2049-2110)."""

    _dataset_spec = {
        "spectrum": {"axes": ["baseline", "delay"], "dtype": np.float64, "distributed_axis": "baseline"},
        "spectrum_samples": {"axes": ["sample", "baseline", "delay"], "dtype": np.float64,
                             "distributed_axis": "baseline"},
        "spectrum_mask": {"axes": ["baseline"], "dtype": bool, "distributed_axis": "baseline"},
    }

    def __init__(self, baseline=None, delay=None, sample=1, attrs_from=None, weight_boost=1.0, axes_from=None):
        if axes_from is not None:  # memh5's axes_from: take the axis definitions of another container
            baseline = axes_from.index_map["baseline"] if baseline is None else baseline
            delay = axes_from.index_map["delay"] if delay is None else delay
            if "sample" in axes_from.index_map:
                sample = axes_from.index_map["sample"]
            weight_boost = getattr(axes_from, "attrs", {}).get("weight_boost", weight_boost)
        super().__init__(attrs_from=attrs_from, baseline=baseline, sample=sample, delay=delay)
        self.attrs["weight_boost"] = weight_boost
        self.add_dataset("spectrum")

    @property
    def spectrum(self):
        return self.datasets["spectrum"]

    @property
    def delay(self):
        return self.index_map["delay"]

    @property
    def weight_boost(self):
        return self.attrs["weight_boost"]

    @property
    def freq(self):
        return self.attrs["freq"]


class DelayCrossSpectrum(DelaySpectrum):
    """Delay cross power spectrum container (draco/core/containers.py:2488-2513).  NB the reference
    declares the datasets float64: only the real part of the complex cross spectra is kept on assignment."""

    _dataset_spec = {
        "spectrum": {"axes": ["dataset", "dataset", "baseline", "delay"], "dtype": np.float64,
                     "distributed_axis": "baseline"},
        "spectrum_samples": {"axes": ["sample", "dataset", "dataset", "baseline", "delay"], "dtype": np.float64,
                             "distributed_axis": "baseline"},
    }

    def __init__(self, baseline, dataset, delay, sample=1, attrs_from=None, weight_boost=1.0):
        Container.__init__(self, attrs_from=attrs_from, baseline=baseline, dataset=dataset, sample=sample, delay=delay)
        self.attrs["weight_boost"] = weight_boost
        self.add_dataset("spectrum")


class DelayTransform(Container):
    """Delay spectrum container (draco/core/containers.py:2113-2182)."""

    _dataset_spec = {
        "spectrum": {"axes": ["baseline", "sample", "delay"], "dtype": np.complex128, "distributed_axis": "baseline"},
        "weight": {"axes": ["baseline", "sample", "delay"], "dtype": np.float32, "distributed_axis": "baseline"},
        "spectrum_mask": {"axes": ["baseline", "sample"], "dtype": bool, "distributed_axis": "baseline"},
    }

    def __init__(self, baseline, sample, delay, attrs_from=None, weight_boost=1.0):
        super().__init__(attrs_from=attrs_from, baseline=baseline, sample=sample, delay=delay)
        self.attrs["weight_boost"] = weight_boost
        self.add_dataset("spectrum")

    @property
    def spectrum(self):
        return self.datasets["spectrum"]

    @property
    def weight(self):
        return self.datasets["weight"]

    @property
    def delay(self):
        return self.index_map["delay"]

    @property
    def freq(self):
        return self.attrs["freq"]


# --------------------------------------------------------------------------- ringmap side (powerspec.py)
class LocalArray(np.ndarray):
    """Single-process slice of an ``MPIArray``: ``dataset[pp]`` keeps ``.local_array`` / ``.allgather()``
    (the two calls ``draco/analysis/powerspec.py:241, 251-252`` make on such slices)."""

    @property
    def local_array(self):
        return np.asarray(self)

    def allgather(self):
        return np.asarray(self)


class _SliceDataset(Dataset):
    def __getitem__(self, key):
        got = super().__getitem__(key)
        return got.view(LocalArray) if isinstance(got, np.ndarray) else got


class RingMap(FreqContainer):
    """The datasets of ``containers.RingMap`` (draco/core/containers.py:1594-1700) the Wiener delay-transform
    tasks read: ``map``, ``weight``, ``dirty_beam_power``, ``filter`` / ``complex_filter``,
    ``freq_cov`` / ``complex_freq_cov``.  Single process (the elevation axis is not split)."""

    _dataset_spec = {
        "map": {"axes": ["beam", "pol", "freq", "ra", "el"], "dtype": np.float64},
        "weight": {"axes": ["pol", "freq", "ra", "el"], "dtype": np.float64},
        "dirty_beam_power": {"axes": ["beam", "pol", "freq", "el"], "dtype": np.float64},
        "filter": {"axes": ["pol", "freq", "freq_sum", "ra"], "dtype": np.float64},
        "complex_filter": {"axes": ["pol", "freq", "freq_sum", "ra"], "dtype": np.complex128},
        "freq_cov": {"axes": ["pol", "freq", "freq_sum", "ra"], "dtype": np.float64},
        "complex_freq_cov": {"axes": ["pol", "freq", "freq_sum", "ra"], "dtype": np.complex128},
    }

    def __init__(self, freq, ra, el, pol=("XX", "YY"), beam=1, attrs_from=None):
        super().__init__(attrs_from=attrs_from, beam=beam, pol=np.asarray(pol), freq=freq, freq_sum=freq, ra=ra, el=el)
        self.add_dataset("map")
        self.add_dataset("weight")

    def add_dataset(self, name):
        spec = self._dataset_spec[name]
        shape = [len(self.index_map[a]) for a in spec["axes"]]
        arr = np.zeros(shape, dtype=spec["dtype"])
        data = DistArray(arr, axis=spec["axes"].index("el")) if name in ("map", "weight") else arr
        self.datasets[name] = _SliceDataset(data, spec["axes"])
        return self.datasets[name]

    map = property(lambda self: self.datasets["map"])
    weight = property(lambda self: self.datasets["weight"])
    dirty_beam_power = property(lambda self: self.datasets["dirty_beam_power"])

    @property
    def filter(self):
        for name in ("filter", "complex_filter"):
            if name in self.datasets:
                return self.datasets[name]
        raise KeyError("Dataset 'filter' not initialised.")

    @property
    def freq_cov(self):
        for name in ("freq_cov", "complex_freq_cov"):
            if name in self.datasets:
                return self.datasets[name]
        raise KeyError("Dataset 'freq_cov' not initialised.")


class DelayTransformOperator(FreqContainer):
    """Wiener filter that takes each pixel from frequency to delay (draco/core/containers.py:2185-2203):
    ``filter [pol, ra, el, delay, freq]`` complex64, distributed over ``el``."""

    _dataset_spec = {
        "filter": {"axes": ["pol", "ra", "el", "delay", "freq"], "dtype": np.complex64, "distributed_axis": "el"},
    }

    def __init__(self, delay, axes_from=None, attrs_from=None, **axes):
        for ax in ("pol", "ra", "el", "freq"):
            if ax not in axes and axes_from is not None:
                axes[ax] = axes_from.index_map[ax]
        super().__init__(attrs_from=attrs_from, delay=delay, **axes)
        self.add_dataset("filter")

    @property
    def filter(self):
        return self.datasets["filter"]

    @property
    def delay(self):
        return self.index_map["delay"]
