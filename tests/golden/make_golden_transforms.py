"""Golden vectors of the REFERENCE's plain FFT delay transform and of its delay-spectrum -> power-spectrum task.

Build-container only (needs /root/reference).

* ``delay_spectrum_fft`` (draco/analysis/delay.py:2102-2130) is executed where it lies.  Its only external call is
  ``caput.fft.fftw.ifft`` (an FFTW wrapper with numpy's ifft convention; caput is absent from this image), for which
  ``numpy.fft.ifft`` stands in -- the two differ by FFT rounding only, so the vectors are compared at 1e-12.
* ``DelaySpectrumToPowerSpectrum`` (delay.py:1061-1106): the class body is executed over the container stand-ins of
  ``draco_b200/_compat.py`` and its ``process`` run on seeded ``DelayTransform`` containers with a spectrum mask
  (one that flags nothing, random ones, a baseline whose samples are all masked).

    python tests/golden/make_golden_transforms.py
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

# name, nsample, N, window
FFT_CASES = [("fft_n64_nuttall", 5, 64, "nuttall"), ("fft_n256_none", 3, 256, None), ("fft_n100_hann", 4, 100, "hann"),
             ("fft_n8_bh", 2, 8, "blackman_harris"), ("fft_n2048_bn", 2, 2048, "blackman_nuttall"),
             ("fft_n37_hamming", 3, 37, "hamming")]
# name, B, T, N, masked
VAR_CASES = [("var_plain", 4, 9, 16, False), ("var_masked", 6, 12, 24, True), ("var_long", 2, 150, 8, True)]


def fft_input(nsample, N, seed):
    r = np.random.default_rng(seed)
    return (r.standard_normal((nsample, N)) + 1j * r.standard_normal((nsample, N))) * np.exp(r.uniform(-3, 3, (nsample, 1)))


def var_input(B, T, N, masked, seed):
    r = np.random.default_rng(seed)
    spec = (r.standard_normal((B, T, N)) + 1j * r.standard_normal((B, T, N))) * 3.0 + (1.5 - 0.5j)
    # NB without a spectrum_mask dataset the reference calls np.var(..., where=None) (delay.py:1097 with w = None), which
    # raises TypeError in numpy: only containers that carry a mask can be recorded; "plain" = a mask that flags nothing
    mask = np.zeros((B, T), dtype=bool)
    if masked:
        mask = r.random((B, T)) < 0.3
        mask[1] = True  # a baseline without any valid sample: NaN variance -> 0 and flagged
        mask[0] = False
    return spec, mask


def load_reference_fft():
    from oracle import ref_loader

    ref = ref_loader.load()
    ns = {"np": np, "tools": types.SimpleNamespace(window_generalised=ref.window_generalised),
          "fft": types.SimpleNamespace(fftw=types.SimpleNamespace(ifft=lambda x, axes=-1: np.fft.ifft(x, axis=axes)))}
    rel = "draco/analysis/delay.py"
    exec(compile(ref_loader._lines(rel, 2102, 2130), rel, "exec"), ns)
    return ns["delay_spectrum_fft"]


def load_reference_to_ps():
    from draco_b200 import _compat
    from oracle import ref_loader

    ns = {"np": np, "containers": _compat,
          "tasklib": types.SimpleNamespace(base=types.SimpleNamespace(ContainerTask=_compat.ContainerTask))}
    rel = "draco/analysis/delay.py"
    exec(compile("from __future__ import annotations\n" + ref_loader._lines(rel, 1061, 1106), rel, "exec"), ns)
    return ns["DelaySpectrumToPowerSpectrum"]


def make_transform(spec, mask):
    from draco_b200 import _compat

    B, T, N = spec.shape
    dt = _compat.DelayTransform(baseline=B, sample=T, delay=np.fft.fftshift(np.fft.fftfreq(N)))
    dt.spectrum[:].local_array[:] = spec
    if mask is not None:
        dt.add_dataset("spectrum_mask")
        dt.datasets["spectrum_mask"][:].local_array[:] = mask
    return dt


def main():
    out = {}
    ref_fft = load_reference_fft()
    for i, (name, ns_, N, window) in enumerate(FFT_CASES):
        out[name] = ref_fft(fft_input(ns_, N, 900 + i).copy(), N, window)
    task = load_reference_to_ps()
    for i, (name, B, T, N, masked) in enumerate(VAR_CASES):
        spec, mask = var_input(B, T, N, masked, 950 + i)
        with np.errstate(all="ignore"):
            ps = task().process(make_transform(spec, mask))
        out[f"{name}_spectrum"] = np.array(ps.spectrum[:].local_array)
        out[f"{name}_mask"] = np.array(ps.datasets["spectrum_mask"][:].local_array)
    path = os.path.join(HERE, "reference_transforms.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
