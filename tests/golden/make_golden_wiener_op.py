"""Generate tests/golden/reference_wiener_op.npz by running the REFERENCE's own ``ConstructWienerDelayTransform``.

Build-container only (needs /root/reference): the class body, draco/analysis/powerspec.py:118-369, is executed
where it lies over the caput / container stand-ins of ``draco_b200/_compat.py`` (``import draco`` is impossible
here) and its ``process`` is run on small seeded ring maps.  The stand-in operator container is made
complex128 for the recording, so the golden operators carry the reference's full fp64 arithmetic (the real
container rounds them to complex64).  The inputs are re-created from the same seeds by the tests
(``make_wiener_case``).

    python tests/golden/make_golden_wiener_op.py
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

# name, nfreq, nra, nel, npol, complex filter/cov, task parameters
WIENER_CASES = [
    ("w_f24_real", 24, 3, 2, 2, False, {}),
    ("w_f32_cplx", 32, 4, 3, 1, True,
     {"window": "nuttall", "window_lower_freq": 620.0, "window_upper_freq": 790.0, "prior_scale": 0.05}),
    ("w_f40_hann", 40, 2, 2, 1, False, {"window": "hann", "prior_amp": 1e-4}),
    ("w_f20_bh_cplx", 20, 5, 1, 2, True, {"window": "blackman_harris", "prior_scale": 0.2}),
]


def make_wiener_case(nfreq, nra, nel, npol, cplx, seed):
    """Seeded ring-map inputs: a soft foreground filter ``K = I - 0.9 V V^H`` per right ascension, the noise
    covariance it leaves (``K diag(sigma^2) K^H``), inverse-variance weights with flagged channels and one dead
    pixel, a positive beam power.  Returns a dict of arrays in the container's axis order."""
    r = np.random.default_rng(seed)
    freq = np.linspace(800.0, 600.0, nfreq, endpoint=False)
    x = np.linspace(-1.0, 1.0, nfreq)
    dt = np.complex128 if cplx else np.float64
    filt = np.zeros((npol, nfreq, nfreq, nra), dtype=dt)
    cov = np.zeros((npol, nfreq, nfreq, nra), dtype=dt)
    for pp in range(npol):
        for rr in range(nra):
            modes = np.polynomial.legendre.legvander(x, 2) * r.uniform(0.8, 1.2, 3)
            if cplx:
                modes = modes * np.exp(1j * r.uniform(0, 2 * np.pi, 3) * x[:, None])
            V, _ = np.linalg.qr(modes)
            K = np.eye(nfreq) - 0.9 * V @ V.conj().T
            sig2 = r.uniform(0.5, 2.0, nfreq) * 1e-5
            filt[pp, :, :, rr] = K
            Cm = (K * sig2) @ K.conj().T
            cov[pp, :, :, rr] = 0.5 * (Cm + Cm.conj().T)
    weight = r.uniform(0.5, 2.0, (npol, nfreq, nra, nel)) * 1e5
    flag = r.random((npol, nfreq, nra, nel)) < 0.15
    weight[flag] = 0.0
    weight[:, :, nra - 1, nel - 1] = 0.0  # a pixel without any valid channel
    beam = r.uniform(0.5, 1.5, (1, npol, nfreq, nel))
    return {"freq": freq, "filter": filt, "freq_cov": cov, "weight": weight, "dirty_beam_power": beam}


def make_ringmap(case):
    """The ``_compat.RingMap`` stand-in holding one case."""
    from draco_b200 import _compat

    npol, nfreq, nra, nel = case["weight"].shape
    rm = _compat.RingMap(freq=case["freq"], ra=np.arange(nra) * 360.0 / nra, el=np.linspace(-1, 1, nel),
                         pol=[f"P{i}" for i in range(npol)])
    cplx = np.iscomplexobj(case["filter"])
    rm.add_dataset("complex_filter" if cplx else "filter")[:] = case["filter"]
    rm.add_dataset("complex_freq_cov" if cplx else "freq_cov")[:] = case["freq_cov"]
    rm.add_dataset("dirty_beam_power")[:] = case["dirty_beam_power"]
    rm.weight[:].local_array[:] = case["weight"]
    return rm


def load_reference_task():
    """exec powerspec.py:118-369 (class ConstructWienerDelayTransform) over the stand-ins."""
    import scipy.linalg

    from draco_b200 import _compat
    from oracle import ref_loader

    ref = ref_loader.load()

    class Operator128(_compat.DelayTransformOperator):
        _dataset_spec = {"filter": dict(_compat.DelayTransformOperator._dataset_spec["filter"], dtype=np.complex128)}

    def invert_no_zero(x):
        x = np.asarray(x)
        with np.errstate(divide="ignore", invalid="ignore"):
            return np.where(x == 0, 0.0, 1.0 / np.where(x == 0, 1.0, x))

    ns = {
        "np": np,
        "scipy": types.SimpleNamespace(linalg=scipy.linalg),
        "config": _compat.config,
        "tasklib": types.SimpleNamespace(base=types.SimpleNamespace(ContainerTask=_compat.ContainerTask)),
        "containers": types.SimpleNamespace(DelayTransformOperator=Operator128),
        "tools": types.SimpleNamespace(window_generalised=ref.window_generalised, invert_no_zero=invert_no_zero),
        "__name__": "draco_reference_powerspec",
    }
    rel = "draco/analysis/powerspec.py"
    exec(compile(ref_loader._lines(rel, 118, 369), rel, "exec"), ns)
    return ns["ConstructWienerDelayTransform"]


def load_reference_apply_task():
    """exec powerspec.py:372-458 (class ApplyWienerDelayTransform) over the stand-ins."""
    from draco_b200 import _compat
    from oracle import ref_loader

    def invert_no_zero(x):
        x = np.asarray(x)
        with np.errstate(divide="ignore", invalid="ignore"):
            return np.where(x == 0, 0.0, 1.0 / np.where(x == 0, 1.0, x))

    ns = {
        "np": np,
        "tasklib": types.SimpleNamespace(base=types.SimpleNamespace(ContainerTask=_compat.ContainerTask)),
        "containers": types.SimpleNamespace(DelayTransform=_compat.DelayTransform),
        "tools": types.SimpleNamespace(invert_no_zero=invert_no_zero),
        "__name__": "draco_reference_powerspec",
    }
    rel = "draco/analysis/powerspec.py"
    exec(compile(ref_loader._lines(rel, 372, 458), rel, "exec"), ns)
    return ns["ApplyWienerDelayTransform"]


def make_map(case, seed):
    """A seeded map [beam, pol, freq, ra, el] for the weights of ``case``."""
    r = np.random.default_rng(seed)
    npol, nfreq, nra, nel = case["weight"].shape
    return r.standard_normal((1, npol, nfreq, nra, nel)) * 1e-2 * (case["weight"] > 0)


def make_operator(case, filt64, delay, params):
    """``_compat.DelayTransformOperator`` (complex64, as the real container) holding the rounded golden operator."""
    from draco_b200 import _compat

    op = _compat.DelayTransformOperator(delay=delay, axes_from=make_ringmap(case))
    op.filter[:].local_array[:] = filt64
    for attr in ["window", "window_lower_freq", "window_upper_freq"]:
        op.attrs[attr] = params.get(attr, "uniform" if attr == "window" else None)
    return op


def main():
    Task = load_reference_task()
    out = {}
    for k, (name, nfreq, nra, nel, npol, cplx, params) in enumerate(WIENER_CASES):
        case = make_wiener_case(nfreq, nra, nel, npol, cplx, 300 + k)
        task = Task.from_config(params) if hasattr(Task, "from_config") else Task()
        op = task.process(make_ringmap(case))
        out[f"{name}_filter"] = op.filter[:].local_array.copy()
        out[f"{name}_delay"] = np.asarray(op.index_map["delay"])
        print(name, out[f"{name}_filter"].shape, "max |D|", np.abs(out[f"{name}_filter"]).max())
        # ApplyWienerDelayTransform on the operator as the real container stores it (complex64)
        rm = make_ringmap(case)
        rm.map[:].local_array[:] = make_map(case, 700 + k)
        dt = load_reference_apply_task()().process(rm, make_operator(case, out[f"{name}_filter"], out[f"{name}_delay"], params))
        out[f"{name}_apply_spectrum"] = dt.spectrum[:].local_array.copy()
        out[f"{name}_apply_weight"] = dt.weight[:].local_array.copy()
    path = os.path.join(HERE, "reference_wiener_op.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
