"""B200-native delay power-spectrum Gibbs estimator (drop-in for the hot path of
``draco.analysis.delay``).  The compute lives in ``libdgibbs.so`` (hand-written sm_100a
CUDA behind a C ABI, ``include/dgibbs.h``); this package is the Python host side."""

__version__ = "0.1.0"
