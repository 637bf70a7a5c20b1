"""Synthetic visibility stacks for the BASELINE configs.

The data model follows the reference's own generator ``mock_freq_data``
(draco/util/testing.py:47-113): a band-limited (``|tau| < delaycut``) Gaussian delay
spectrum of unit power, transformed to frequency, plus white noise of known rms, with
``weight = 1 / noise**2`` and optionally zeroed channels.  Everything is seeded with
``np.random.default_rng`` so the CPU oracle and the GPU path see identical inputs.
"""

from __future__ import annotations

import dataclasses

import numpy as np


def _complex_normal(rng, shape):
    # unit total variance: re and im each N(0, 1/2)   (draco/util/random.py:7-100)
    return (rng.standard_normal(shape) + 1j * rng.standard_normal(shape)) * np.sqrt(0.5)


def band_limited_stack(nbase, ntime, nfreq, delaycut=0.2, noise=1e-2, seed=0):
    """Return ``(freq[nfreq], vis[nbase, ntime, nfreq] c128, weight[nbase, ntime, nfreq] f64)``."""
    rng = np.random.default_rng(seed)
    freq = np.linspace(800.0, 400.0, nfreq, endpoint=False)
    df = np.abs(freq[1] - freq[0])
    delays = np.fft.fftfreq(nfreq, df)
    amp = np.where(np.abs(delays) < delaycut, 1.0, 0.0) ** 0.5
    vis = np.empty((nbase, ntime, nfreq), dtype=np.complex128)
    # generate in slabs to bound the temporary
    step = max(1, (1 << 24) // max(1, ntime * nfreq))
    for b0 in range(0, nbase, step):
        b1 = min(nbase, b0 + step)
        spec = _complex_normal(rng, (b1 - b0, ntime, nfreq)) * amp
        vis[b0:b1] = np.fft.fft(spec, axis=-1) + noise * _complex_normal(rng, (b1 - b0, ntime, nfreq))
    weight = np.full((nbase, ntime, nfreq), 1.0 / noise**2)
    return freq, vis, weight


@dataclasses.dataclass
class SynthConfig:
    """One BASELINE.json configuration."""

    name: str
    nbase: int
    nfreq: int
    ntime: int
    nsamp: int
    window: str | None
    seed: int
    mask: str = "none"  # none | common20 | bands40
    weights: str = "uniform"  # uniform | perfreq
    ngroup: int = 0

    @property
    def ndelay(self):
        # skip_nyquist=True default: nfreq_total = nfreq + 1, N = 2 * nfreq  (delay.py:500-507)
        return 2 * self.nfreq


CONFIGS = {
    "cfg1": SynthConfig("cfg1", 64, 128, 32, 20, "nuttall", 1, mask="common20"),
    "cfg2": SynthConfig("cfg2", 1024, 1024, 64, 100, "nuttall", 2, weights="perfreq"),
    "cfg3": SynthConfig("cfg3", 16384, 1024, 64, 20, None, 3),
    "cfg4": SynthConfig("cfg4", 4096, 1024, 64, 200, "nuttall", 2, weights="perfreq", ngroup=64),
    "cfg5": SynthConfig("cfg5", 40000, 2048, 16, 10, "nuttall", 5, mask="bands40", weights="perfreq"),
}


def make_stack(cfg: SynthConfig, nbase=None, base_offset=0):
    """Build the ``[baseline, time, freq]`` stack of a config (optionally only
    ``nbase`` baselines starting at global index ``base_offset``: each baseline's data
    depends only on its global index, so shards of a multi-GPU run agree with the
    single-GPU run)."""
    nb = cfg.nbase if nbase is None else nbase
    noise = 1e-2
    vis = np.empty((nb, cfg.ntime, cfg.nfreq), dtype=np.complex128)
    weight = np.empty((nb, cfg.ntime, cfg.nfreq), dtype=np.float64)
    freq = None
    # one seed per block of 64 baselines keeps generation cheap and shard-independent
    blk = 64
    b = 0
    while b < nb:
        g = base_offset + b
        gblk = g // blk
        lo = g - gblk * blk
        n = min(blk - lo, nb - b)
        freq, v, w = band_limited_stack(blk, cfg.ntime, cfg.nfreq, noise=noise, seed=(cfg.seed, gblk))
        vis[b : b + n] = v[lo : lo + n]
        weight[b : b + n] = w[lo : lo + n]
        if cfg.weights == "perfreq":
            u = np.random.default_rng((cfg.seed, gblk, 1)).uniform(0.5, 2.0, (blk, 1, cfg.nfreq))
            weight[b : b + n] *= u[lo : lo + n]
        if cfg.mask == "bands40":
            # contiguous RFI-like bands, redrawn for every block of 500 baselines
            for i in range(n):
                bad = rfi_band_mask(cfg.nfreq, 0.4, seed=(cfg.seed, (g + i) // 500, 2))
                vis[b + i][:, bad] = 0.0
                weight[b + i][:, bad] = 0.0
        b += n
    if cfg.mask == "common20":
        bad = np.random.default_rng(cfg.seed).choice(cfg.nfreq, size=round(0.2 * cfg.nfreq), replace=False)
        vis[:, :, bad] = 0.0
        weight[:, :, bad] = 0.0
    return freq, vis, weight


def rfi_band_mask(nfreq, frac, seed):
    """Boolean mask flagging ~``frac`` of channels in contiguous bands."""
    rng = np.random.default_rng(seed)
    bad = np.zeros(nfreq, dtype=bool)
    target = int(round(frac * nfreq))
    while bad.sum() < target:
        width = int(rng.integers(max(2, nfreq // 128), max(3, nfreq // 16)))
        start = int(rng.integers(0, nfreq - width))
        bad[start : start + width] = True
    # trim the overshoot from the top so the flagged fraction is exact
    extra = int(bad.sum() - target)
    if extra > 0:
        idx = np.flatnonzero(bad)[-extra:]
        bad[idx] = False
    return bad
