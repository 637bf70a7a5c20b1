"""Posterior statistics of the REFERENCE's own sampler at the geometries the benchmarks time (N = 2048, T = 64).

Run in the build container only (needs ``/root/reference``; ~2 minutes on 8 cores):

    python tests/golden/make_golden_stats.py

``delay_power_spectrum_gibbs`` (draco/analysis/delay.py:1713-1904, executed in place through ``oracle/ref_loader.py``)
is run for NCHAIN seeded chains x NITER iterations on ONE baseline of
  * ``cfg3``: BASELINE config 3 -- uniform weights, no window, Nyquist skipped: the device's diagonal(+rank-1) branch
    (``diag_resident_kernel``: collapsed-moment draws, fp32 Marsaglia-Tsang chi^2);
  * ``cfg2``: BASELINE config 2 -- per-frequency weights, Nuttall window: the dense branch (``dense_sampler_kernel``).
Stored per case (``reference_stats.npz``): per-delay mean and variance over the chains of the chain-mean of log S after
burn-in, the final-iteration log S of every chain at 64 delays spread inside and outside the 0.2 us signal band, and
the seeds.  The GPU tests (tests/test_gpu_stats.py) hold the device-RNG chains to them: 3 sigma on the means and a
two-sample KS test per stored delay.
"""
from __future__ import annotations

import multiprocessing as mp
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

NITER, BURN = 40, 20
CASES = {"cfg3": 32, "cfg2": 16}  # chains per case
NSEL = 64


def case_inputs(name):
    """One baseline of the named BASELINE config after the reference's _cut_data with default parameters."""
    import oracle
    from draco_b200 import synth

    cfg = synth.CONFIGS[name]
    freq, vis, wgt = synth.make_stack(cfg, nbase=1, base_offset=0, threads=1)
    delays, chan = oracle.calculate_delays(freq)
    d, w, nzf, _ = oracle.cut_data(vis[0], wgt[0])
    return cfg, d, w, chan[nzf], len(delays), delays


def selected_delays(delays_shifted, N):
    """64 delay indices (unshifted order): half inside |tau| < 0.2 us, half outside."""
    tau = np.fft.ifftshift(delays_shifted)
    inside = np.flatnonzero(np.abs(tau) < 0.2)
    outside = np.flatnonzero(np.abs(tau) >= 0.2)
    pick = lambda idx: idx[np.linspace(0, len(idx) - 1, NSEL // 2).astype(int)]  # noqa: E731
    return np.sort(np.concatenate([pick(inside), pick(outside)]))


def _chain(args):
    os.environ["OPENBLAS_NUM_THREADS"] = "1"
    from threadpoolctl import threadpool_limits

    from oracle import ref_loader

    name, seed = args
    with threadpool_limits(1):
        ref = ref_loader.load()
        cfg, d, w, fsel, N, _ = case_inputs(name)
        spec, ok = ref.delay_power_spectrum_gibbs(d.copy(), N, w, np.full(N, 10.0), window=cfg.window, fsel=fsel,
                                                  niter=NITER, rng=np.random.default_rng(seed))
    assert ok and len(spec) == NITER
    ls = np.log(np.array(spec))
    return ls[BURN:].mean(axis=0), ls[-1]


def main():
    out = {"niter": np.array(NITER), "burn": np.array(BURN)}
    ctx = mp.get_context("spawn")
    with ctx.Pool(os.cpu_count()) as pool:
        for name, nchain in CASES.items():
            seeds = [70000 + 1000 * list(CASES).index(name) + c for c in range(nchain)]
            res = pool.map(_chain, [(name, s) for s in seeds])
            cm = np.array([r[0] for r in res])      # [chain, delay] chain means of log S
            last = np.array([r[1] for r in res])    # [chain, delay]
            _, _, _, _, N, delays = case_inputs(name)
            sel = selected_delays(delays, N)
            out[f"{name}_mean"] = cm.mean(axis=0)
            out[f"{name}_var"] = cm.var(axis=0, ddof=1)
            out[f"{name}_nchain"] = np.array(nchain)
            out[f"{name}_sel"] = sel
            out[f"{name}_last_sel"] = last[:, sel]
            out[f"{name}_seeds"] = np.array(seeds)
            print(name, "done:", nchain, "chains; mean log S in/out of band:",
                  cm.mean(axis=0)[sel[: NSEL // 2]].mean(), cm.mean(axis=0)[sel[NSEL // 2:]].mean())
    np.savez_compressed(os.path.join(HERE, "reference_stats.npz"), **out)


if __name__ == "__main__":
    main()
