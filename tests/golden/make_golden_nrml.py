"""Generate tests/golden/reference_nrml.npz by running the REFERENCE's own NRML / maximum-posterior estimator.

Build-container only (needs /root/reference): executes ``LogLikePS``, ``GaussianProcessPrior`` and
``delay_power_spectrum_maxpost`` (draco/analysis/delayopt.py:61-561, with draco/util/kernels.py) through
``oracle/ref_loader.py`` and stores objective values / gradients / Hessians at seeded points plus every
iterate of short optimisations.  The inputs are re-created from the same seeds by the tests (``make_nrml_case``).

    python tests/golden/make_golden_nrml.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

# name, N, M, T, window, flagged (Ni = 0) channels, maxiter
NRML_CASES = [
    ("m_n32", 32, 17, 12, "nuttall", 2, 25),
    ("m_n64_nowin", 64, 33, 20, None, 0, 20),
    ("m_n48_bh", 48, 20, 9, "blackman_harris", 3, 20),
]


def make_nrml_case(N, M, T, nflag, seed):
    """Seeded inputs of one baseline: band-limited delay spectrum + noise on M of the N/2+1 channels."""
    r = np.random.default_rng(seed)
    fsel = np.sort(r.choice(N // 2 + 1, size=M, replace=False))
    tau = np.fft.fftfreq(N)
    amp = np.where(np.abs(tau) < 0.12, 3.0, 0.05)
    d = (r.standard_normal((T, N)) + 1j * r.standard_normal((T, N))) * amp
    data = np.fft.fft(d, axis=-1)[:, fsel] / np.sqrt(N)
    data = data + 0.3 * (r.standard_normal((T, M)) + 1j * r.standard_normal((T, M)))
    Ni = r.uniform(0.5, 2.0, M) * 10.0
    flagged = r.choice(M, size=nflag, replace=False)
    Ni[flagged] = 0.0
    data[:, flagged] = 0.0
    return data, Ni, fsel


def main():
    from oracle import nrml_oracle, ref_loader

    ref = ref_loader.load()
    out = {}
    for k, (name, N, M, T, window, nflag, maxiter) in enumerate(NRML_CASES):
        data, Ni, fsel = make_nrml_case(N, M, T, nflag, 100 + k)
        _, X, F, Nm = nrml_oracle.maxpost_inputs(data, N, Ni, window, fsel)
        like = ref.delayopt.LogLikePS(X, F, Nm, T, exact_hessian=True, bounds=(1e-15, 1e10))
        fish = ref.delayopt.LogLikePS(X, F, Nm, T, exact_hessian=False, bounds=(1e-15, 1e10))
        prior = ref.delayopt.GaussianProcessPrior(N, width=5, alpha=1.0, kernel="matern", nu=1.5)
        r = np.random.default_rng(900 + k)
        for j in range(2):
            x = r.normal(0.0, 1.5, N)
            out[f"{name}_x{j}"] = x
            out[f"{name}_value{j}"] = np.array(like.value(x))
            out[f"{name}_grad{j}"] = like.gradient(x)
            out[f"{name}_hess{j}"] = like.hessian(x)
            out[f"{name}_fisher{j}"] = fish.hessian(x)
            out[f"{name}_prior_value{j}"] = np.array(prior.value(x))
            out[f"{name}_prior_grad{j}"] = prior.gradient(x)
        out[f"{name}_prior_Ci"] = prior.Ci
        samples, ok = ref.delayopt.delay_power_spectrum_maxpost(data, N, Ni, None, window=window, fsel=fsel,
                                                                maxiter=maxiter, tol=1e-3)
        out[f"{name}_samples"] = np.array(samples)
        out[f"{name}_success"] = np.array(ok)
        print(name, "iterates", len(samples), "success", ok)
    path = os.path.join(HERE, "reference_nrml.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
