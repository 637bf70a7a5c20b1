"""ctypes binding of ``libdgibbs.so`` (the C ABI declared in ``include/dgibbs.h``).

There is no CPU fallback: if the shared library is missing, or no CUDA device can be
opened, the calls raise.  The library is loaded from this directory only (in-tree build).
"""

from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# DGIBBS_LIB_PATH: load another build of the library (A/B timing of kernel variants, scratch/build_variant.sh)
LIB_PATH = os.environ.get("DGIBBS_LIB_PATH") or os.path.join(HERE, "libdgibbs.so")

WINDOW_TUKEY_BASE = 1 << 20


class _WindowIds(dict):
    """Window name -> ``window_id`` of include/dgibbs.h; ``"tukey-0.X"`` (draco/util/tools.py:569-581) maps to
    ``DGIBBS_WINDOW_TUKEY(X * 1e6)``."""

    @staticmethod
    def _tukey(name):
        if not (isinstance(name, str) and name.startswith("tukey")):
            return None
        try:
            r = float(name.split("-")[1])
        except (IndexError, ValueError):
            return None
        m = round(r * 1e6)
        # the device rebuilds r as m / 1e6: only fractions with at most six decimals map onto themselves
        if not (0 < m <= 1000000) or m / 1e6 != r:
            return None
        return WINDOW_TUKEY_BASE + m

    def __contains__(self, name):
        return dict.__contains__(self, name) or self._tukey(name) is not None

    def __missing__(self, name):
        wid = self._tukey(name)
        if wid is None:
            raise KeyError(name)
        return wid


WINDOW_IDS = _WindowIds({
    None: -1,
    "uniform": 0,
    "hann": 1,
    "hanning": 2,
    "hamming": 3,
    "blackman": 4,
    "nuttall": 5,
    "blackman_nuttall": 6,
    "blackman_harris": 7,
    "triangular": 8,
})
RNG_INJECTED, RNG_PHILOX, RNG_ZERO = 0, 1, 2
BRANCH_SKIPPED, BRANCH_DELAY_SPACE, BRANCH_FREQ_SPACE, BRANCH_DIAGONAL = 0, 1, 2, 3

EXPORTED = [
    "dgibbs_version",
    "dgibbs_rng_info",
    "dgibbs_sizeof",
    "dgibbs_create",
    "dgibbs_destroy",
    "dgibbs_last_error",
    "dgibbs_workspace_bytes",
    "dgibbs_run",
    "dgibbs_get_timing",
    "dgibbs_setup_inputs",
    "dgibbs_wiener",
    "dgibbs_fourier_matrix",
    "dgibbs_cross_run",
    "dgibbs_cut_data",
    "dgibbs_delay_fft",
    "dgibbs_delay_fft_noalloc",
    "dgibbs_sample_variance",
]
CROSS_MAX_DATASETS = 4

_vp = C.c_void_p


class Problem(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32),
        ("on_device", C.c_int32),
        ("B", C.c_int32),
        ("N", C.c_int32),
        ("niter", C.c_int32),
        ("window_id", C.c_int32),
        ("complex_timedomain", C.c_int32),
        ("rng_mode", C.c_int32),
        ("seed", C.c_uint64),
        ("median_frac", C.c_double),
        ("T", _vp),
        ("M", _vp),
        ("data_off", _vp),
        ("chan_off", _vp),
        ("data", _vp),
        ("Ni", _vp),
        ("fsel", _vp),
        ("S0", _vp),
        ("global_id", _vp),
        ("w1", _vp),
        ("w1_off", _vp),
        ("w2", _vp),
        ("w2_off", _vp),
        ("chi2", _vp),
        ("force_branch", C.c_int32),
        ("cluster_size", C.c_int32),
        ("data_c64", C.c_int32),
        ("reserved", C.c_int32),
    ]


class Result(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32),
        ("on_device", C.c_int32),
        ("spectrum", _vp),
        ("samples", _vp),
        ("n_done", _vp),
        ("success", _vp),
        ("branch", _vp),
        ("final_S", _vp),
    ]


class CrossProblem(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32),
        ("B", C.c_int32),
        ("nd", C.c_int32),
        ("N", C.c_int32),
        ("T", C.c_int32),
        ("M", C.c_int32),
        ("niter", C.c_int32),
        ("window_id", C.c_int32),
        ("rng_mode", C.c_int32),
        ("reserved", C.c_int32),
        ("seed", C.c_uint64),
        ("data", _vp),
        ("Ni", _vp),
        ("fsel", _vp),
        ("S0", _vp),
        ("global_id", _vp),
        ("w1", _vp),
        ("w2", _vp),
        ("wn", _vp),
        ("wg", _vp),
    ]


class Cut(C.Structure):
    """``dgibbs_cut`` of include/dgibbs.h (device-side ``_cut_data``)."""

    _fields_ = [
        ("struct_size", C.c_int32),
        ("on_device", C.c_int32),
        ("B", C.c_int32),
        ("T", C.c_int32),
        ("nfreq", C.c_int32),
        ("weight_f32", C.c_int32),
        ("remove_mean", C.c_int32),
        ("scale_freq", C.c_int32),
        ("data_c64", C.c_int32),
        ("reserved0", C.c_int32),
        ("time_frac", C.c_double),
        ("freq_frac", C.c_double),
        ("weight_boost", C.c_double),
        ("data", _vp),
        ("weight", _vp),
        ("channel_ind", _vp),
        ("data_out", _vp),
        ("Ni_out", _vp),
        ("fsel_out", _vp),
        ("T_out", _vp),
        ("M_out", _vp),
        ("alive", _vp),
        ("keep_t", _vp),
        ("keep_f", _vp),
    ]


class Timing(C.Structure):
    _fields_ = [
        ("ms_h2d", C.c_double),
        ("ms_setup", C.c_double),
        ("ms_sampler", C.c_double),
        ("ms_finish", C.c_double),
        ("launches", C.c_int64),
        ("sampler_launches", C.c_int64),
        ("flop_sampler", C.c_double),
        ("bytes_sampler", C.c_double),
        ("ms_diag_project", C.c_double),
        ("ms_sampler_diag", C.c_double),
        ("n_dense", C.c_int64),
        ("n_diag", C.c_int64),
    ]


_lib = None


def load():
    """Load (once) and return the shared library; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m draco_b200._build` "
            "(or __graft_entry__.build()).  There is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    lib.dgibbs_version.restype = C.c_int
    lib.dgibbs_rng_info.restype = C.c_char_p
    lib.dgibbs_sizeof.argtypes = [C.c_int]
    lib.dgibbs_create.argtypes = [C.c_int, C.c_uint64, C.POINTER(_vp)]
    lib.dgibbs_destroy.argtypes = [_vp]
    lib.dgibbs_last_error.argtypes = [_vp]
    lib.dgibbs_last_error.restype = C.c_char_p
    lib.dgibbs_workspace_bytes.argtypes = [_vp, C.POINTER(Problem), C.POINTER(C.c_int64)]
    lib.dgibbs_run.argtypes = [_vp, C.POINTER(Problem), C.POINTER(Result)]
    lib.dgibbs_get_timing.argtypes = [_vp, C.POINTER(Timing)]
    lib.dgibbs_setup_inputs.argtypes = [_vp, C.POINTER(Problem), C.c_int32, _vp, _vp, _vp]
    lib.dgibbs_wiener.argtypes = [_vp, C.POINTER(Problem), _vp, _vp]
    lib.dgibbs_fourier_matrix.argtypes = [_vp, C.c_int32, C.c_int32, _vp, C.c_int32, _vp]
    lib.dgibbs_cross_run.argtypes = [_vp, C.POINTER(CrossProblem), _vp, _vp, _vp]
    lib.dgibbs_cut_data.argtypes = [_vp, C.POINTER(Cut)]
    lib.dgibbs_delay_fft.argtypes = [_vp, C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.c_int32, _vp, _vp,
                                     C.POINTER(C.c_double)]
    lib.dgibbs_delay_fft_noalloc.argtypes = [_vp, C.c_int64, C.c_int32, C.c_int32, C.c_int32, _vp, _vp, C.c_int32]
    lib.dgibbs_sample_variance.argtypes = [_vp, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _vp, _vp, _vp, _vp,
                                           C.POINTER(C.c_double)]
    for name in EXPORTED:
        if name not in ("dgibbs_last_error", "dgibbs_version", "dgibbs_rng_info"):
            getattr(lib, name).restype = C.c_int
    for which, st in enumerate((Problem, Result, Timing, CrossProblem, Cut)):
        if lib.dgibbs_sizeof(which) != C.sizeof(st):
            raise RuntimeError(f"ABI mismatch: {st.__name__} is {C.sizeof(st)} bytes here, "
                               f"{lib.dgibbs_sizeof(which)} in {LIB_PATH}; rebuild the library")
    _lib = lib
    return lib


class DGibbsError(RuntimeError):
    pass


class Handle:
    """One sampler context bound to one CUDA device (1 rank <-> 1 GPU)."""

    def __init__(self, device=0, stream=0):
        self.lib = load()
        self._h = _vp()
        rc = self.lib.dgibbs_create(int(device), int(stream), C.byref(self._h))
        if rc != 0:
            raise DGibbsError(f"dgibbs_create(device={device}) failed with code {rc} (no CUDA device / bad id?)")
        self.device = int(device)

    def check(self, rc, what):
        if rc != 0:
            msg = self.lib.dgibbs_last_error(self._h)
            raise DGibbsError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    def timing(self):
        t = Timing()
        self.check(self.lib.dgibbs_get_timing(self._h, C.byref(t)), "dgibbs_get_timing")
        return {k: getattr(t, k) for k, _ in Timing._fields_}

    def close(self):
        if self._h:
            self.lib.dgibbs_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
