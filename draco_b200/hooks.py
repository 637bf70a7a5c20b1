"""``draco_b200.install()``: run existing pipeline configs on the GPU WITHOUT editing them.

A caput YAML addresses the task by dotted path (``type: draco.analysis.delay.DelayPowerSpectrumGibbs``,
delay.py:1218).  ``install()`` rebinds, on the REFERENCE's own classes, only the methods SURVEY 8b names as the
plug-in surface -- ``_evaluate`` (delay.py:1117) and ``_estimator`` (delay.py:1242) plus the private helpers they
call -- and adds the few optional config properties of this package (``device``, ``batch_baselines``, ``rng_mode``,
``device_cut``; their defaults change nothing for an existing config).  Everything else stays the reference's:
``process`` (delay.py:431-459), ``_calculate_delays``, ``_cut_data``, ``_prepare_inputs`` / ``flatten_axes`` -- so the
MPI redistribution over the flattened baseline axis (delay.py:714, 2278-2298) is caput's own -- ``_create_output``
and the container classes.  Call it once per process before the pipeline is built, e.g. from a one-line
``sitecustomize`` or at the top of the job script:

    import draco_b200; draco_b200.install()

``uninstall()`` restores the reference's methods.
"""
from __future__ import annotations

import importlib

# name of the reference class -> attributes taken from the class of the same name in draco_b200.delay
_COMMON = ("_device", "_cut_all", "_device_cut_serves", "_chunks", "device", "batch_baselines")
_SURFACE = {
    "DelayPowerSpectrumGibbs": _COMMON + ("_evaluate", "_evaluate_device_cut", "_estimator", "_device_seed", "_window",
                                          "_ncalls", "rng_mode", "device_cut"),
    "DelayPowerSpectrumNRML": _COMMON + ("_evaluate", "_estimator"),
    "DelaySpectrumWienerFilter": _COMMON + ("_evaluate", "_estimator", "device_cut"),
    "DelaySpectrumFFT": _COMMON + ("_evaluate", "_estimator", "_check_width", "device_cut"),
    # the reference's wrapper cannot run (delay.py:1328, 1330): its input / output plumbing is replaced as well
    "DelayCrossPowerSpectrumEstimator": ("_evaluate", "_prepare_inputs", "_create_output"),
    "DelaySpectrumToPowerSpectrum": ("process", "device"),
}
_saved: dict = {}
_MISSING = object()


def _lookup(cls, name):
    for klass in cls.__mro__:
        if name in vars(klass):
            return vars(klass)[name]
    raise AttributeError(f"{cls.__name__} has no attribute {name}")


def install(module=None):
    """Rebind the device evaluation on the task classes of ``module`` (default ``draco.analysis.delay``).
    Returns the list of ``Class.attribute`` names that were bound."""
    from . import delay as ours

    if module is None:
        module = importlib.import_module("draco.analysis.delay")
    elif isinstance(module, str):
        module = importlib.import_module(module)
    bound = []
    for cname, attrs in _SURFACE.items():
        ref = getattr(module, cname, None)
        new = getattr(ours, cname, None)
        if ref is None or new is None or ref is new:
            continue
        for a in attrs:
            obj = _lookup(new, a)
            key = (module.__name__, cname, a)
            if key not in _saved:
                _saved[key] = vars(ref).get(a, _MISSING)
            setattr(ref, a, obj)
            if hasattr(obj, "__set_name__"):  # config properties learn their name when a class body is executed
                obj.__set_name__(ref, a)
            bound.append(f"{cname}.{a}")
    # subclasses that only rename (historical aliases) inherit the rebinding through the MRO
    return bound


def uninstall(module=None):
    """Undo :func:`install` (all modules, or one)."""
    name = None if module is None else (module if isinstance(module, str) else module.__name__)
    for (mname, cname, a), old in list(_saved.items()):
        if name is not None and mname != name:
            continue
        mod = importlib.import_module(mname) if mname in importlib.sys.modules else None
        ref = getattr(mod, cname, None) if mod is not None else None
        if ref is not None:
            if old is _MISSING:
                if a in vars(ref):
                    delattr(ref, a)
            else:
                setattr(ref, a, old)
        del _saved[(mname, cname, a)]
