"""B200-native delay power-spectrum Gibbs estimator (drop-in for the hot path of
``draco.analysis.delay``).  The compute lives in ``libdgibbs.so`` (hand-written sm_100a
CUDA behind a C ABI, ``include/dgibbs.h``); this package is the Python host side."""

__version__ = "0.1.0"


def install(module=None):
    """Rebind the GPU evaluation on the reference's own task classes so that an existing pipeline config
    (``type: draco.analysis.delay.DelayPowerSpectrumGibbs``) runs unmodified; see :mod:`draco_b200.hooks`."""
    from . import hooks

    return hooks.install(module)


def uninstall(module=None):
    from . import hooks

    return hooks.uninstall(module)
