"""Reference-generated pin of the Wiener delay-transform operator at F = 256 channels -- the size from which
``dgwiener_construct`` takes its contractions to the int8 tensor cores (solver 4).  The GPU tests hold the device to the
oracle at this size (tests/test_gpu_wiener_op.py::test_auto_solver_takes_the_int8_contractions_from_256_channels); this
fixture holds the ORACLE to the reference's own ``ConstructWienerDelayTransform.process`` (draco/analysis/powerspec.py:118-369)
there, so the chain device -> oracle -> reference is closed at the geometry of the int8 path.

    python tests/golden/make_golden_wiener_f256.py        (needs /root/reference; writes reference_wiener_op_f256.npz)

To keep the fixture small only four delay rows of every pixel's operator and the Frobenius norm of each pixel's full
operator are stored (complex128: the recording container is widened as in make_golden_wiener_op.py).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from make_golden_wiener_op import load_reference_task, make_ringmap, make_wiener_case  # noqa: E402

# name, nfreq, nra, nel, npol, complex inputs, task parameters, seed
LARGE_CASE = ("w_f256_cplx", 256, 3, 1, 1, True, {"window": "blackman_harris", "prior_scale": 0.02}, 411)
ROWS = (0, 1, 37, 127)


def main():
    name, nfreq, nra, nel, npol, cplx, params, seed = LARGE_CASE
    Task = load_reference_task()
    case = make_wiener_case(nfreq, nra, nel, npol, cplx, seed)
    task = Task.from_config(params) if hasattr(Task, "from_config") else Task()
    op = task.process(make_ringmap(case))
    full = op.filter[:].local_array  # [pol, ra, el, delay, freq]
    delay = np.asarray(op.index_map["delay"])
    out = {
        f"{name}_rows": full[:, :, :, list(ROWS), :].copy(),
        f"{name}_norm": np.sqrt((np.abs(full) ** 2).sum(axis=(-1, -2))),
        f"{name}_delay": delay,
    }
    print(name, full.shape, "max |D|", np.abs(full).max(), "ndelay", delay.size)
    path = os.path.join(HERE, "reference_wiener_op_f256.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
