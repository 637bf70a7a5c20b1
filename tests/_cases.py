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
