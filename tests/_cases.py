"""Seeded inputs shared by the CPU and GPU tests (same generator as the golden script)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden import CHAIN_CASES, make_case  # noqa: E402,F401

import oracle  # noqa: E402

CASES = {c[0]: c for c in CHAIN_CASES}


def recorded_chain(data, N, Ni, S0, window, fsel, niter, cplx, seed=7):
    """Oracle chain + the draws it consumed, packed as (w1, w2, chi2)."""
    rec = oracle.RecordingDraws(np.random.default_rng(seed))
    samples, ok = oracle.gibbs_chain(data, N, Ni, S0, window, fsel, niter, rec, cplx)
    return np.array(samples), ok, rec.packed()


# parameter sets of tests/golden/make_golden_cut_c64.py (complex64 / float32 _cut_data)
C64_CASES = [("default", {}), ("timefrac", dict(time_frac=0.3)), ("freqfrac", dict(freq_frac=0.85)),
             ("nomean", dict(remove_mean=False)), ("boost", dict(weight_boost=2.5, time_frac=0.1))]
