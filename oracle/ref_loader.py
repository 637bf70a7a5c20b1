"""Load the reference's own delay functions by executing them where they lie.

TEST INFRASTRUCTURE ONLY, and BUILD-CONTAINER ONLY: ``/root/reference`` does not exist
on the GPU box, so nothing in the gpu tests / smoke / bench calls this.  It is used by
``tests/golden/make_golden.py`` to generate the committed golden vectors and by
``tests/test_oracle.py`` (skipped when the reference tree is absent) to compare the
restatement in ``gibbs_oracle.py`` with the real thing.

``import draco`` is impossible here (caput / cora / the Cython module are absent,
SURVEY.md 8c), but the numerical functions are pure numpy/scipy: we ``exec`` the line
ranges ``draco/analysis/delay.py:1475-2201`` and ``draco/util/tools.py:547-601`` with a
small shim for the two caput helpers they touch.  No reference source is copied into
this repository; it is read from the reference tree at run time.
"""

from __future__ import annotations

import os
import textwrap
import types

import numpy as np
import scipy.linalg as la

REFERENCE_ROOT = os.environ.get("DRACO_REFERENCE_ROOT", "/root/reference")
_DELAY = "draco/analysis/delay.py"
_TOOLS = "draco/util/tools.py"
_RANDOM = "draco/util/random.py"


def available() -> bool:
    return os.path.exists(os.path.join(REFERENCE_ROOT, _DELAY))


def _lines(rel, first, last):
    with open(os.path.join(REFERENCE_ROOT, rel)) as fh:
        src = fh.readlines()
    return "".join(src[first - 1 : last])


def _invert_no_zero(x):
    # caput.algorithms.invert_no_zero: 1/x where x != 0 else 0
    x = np.asarray(x)
    with np.errstate(divide="ignore", invalid="ignore"):
        return np.where(x == 0, 0.0, 1.0 / np.where(x == 0, 1.0, x))


def load():
    """Return a namespace holding the reference's module-level delay functions plus
    ``cut_data(params, data, weight)`` and ``calculate_delays(params, freq)`` wrappers
    around the two task methods."""
    if not available():
        raise FileNotFoundError(f"reference tree not found under {REFERENCE_ROOT}")

    tools_ns = {"np": np}
    exec(compile(_lines(_TOOLS, 547, 601), _TOOLS, "exec"), tools_ns)
    tools = types.SimpleNamespace(
        window_generalised=tools_ns["window_generalised"],
        invert_no_zero=_invert_no_zero,
    )
    # caput.algorithms.random is external; its in-tree twin draco/util/random.py:7-166 supplies the
    # complex normal / complex Wishart helpers the cross-spectrum sampler calls (delay.py:2029-2030, 2079)
    rand_ns = {"np": np, "default_rng": np.random.default_rng}
    exec(compile(_lines(_RANDOM, 7, 166), _RANDOM, "exec"), rand_ns)
    random = types.SimpleNamespace(
        default_rng=np.random.default_rng,
        standard_complex_normal=rand_ns["standard_complex_normal"],
        complex_normal=rand_ns["complex_normal"],
        standard_complex_wishart=rand_ns["standard_complex_wishart"],
        complex_wishart=rand_ns["complex_wishart"],
    )

    class FreqContainer:  # stand-in for draco.core.containers.FreqContainer
        def __init__(self, freq):
            self.freq = np.asarray(freq)

    containers = types.SimpleNamespace(FreqContainer=FreqContainer)

    ns = {
        "np": np,
        "la": la,
        "tools": tools,
        "random": random,
        "containers": containers,
        "__name__": "draco_reference_delay",
    }
    # Annotations in the def lines are not evaluated (postponed evaluation).
    header = "from __future__ import annotations\n"
    exec(compile(header + _lines(_DELAY, 1475, 2201), _DELAY, "exec"), ns)

    # The two task methods that do arithmetic: dedent and exec as plain functions.
    meth_src = header + textwrap.dedent(_lines(_DELAY, 461, 512)) + "\n"
    meth_src += textwrap.dedent(_lines(_DELAY, 516, 597))
    exec(compile(meth_src, _DELAY, "exec"), ns)

    def cut_data(params, data, weight):
        return ns["_cut_data"](types.SimpleNamespace(**params), data, weight)

    def calculate_delays(params, freq):
        return ns["_calculate_delays"](types.SimpleNamespace(**params), FreqContainer(freq))

    # draco/analysis/delayopt.py (the NRML / maximum-posterior estimator) with draco/util/kernels.py: both pure
    # numpy/scipy.  `from ..util import kernels, tools` and the late `from .delay import fourier_matrix`
    # (delayopt.py:472) are resolved through stand-in modules holding the objects loaded above.
    import sys

    kern_ns = {"__name__": "draco_reference_kernels"}
    exec(compile(_lines("draco/util/kernels.py", 1, 10**6), "draco/util/kernels.py", "exec"), kern_ns)
    kernels = types.SimpleNamespace(**{k: v for k, v in kern_ns.items() if callable(v)})
    pkg, util, ana = (types.ModuleType(n) for n in ("_dref", "_dref.util", "_dref.analysis"))
    dmod = types.ModuleType("_dref.analysis.delay")
    dmod.fourier_matrix = ns["fourier_matrix"]
    util.kernels, util.tools = kernels, tools
    pkg.util, pkg.analysis, ana.delay = util, ana, dmod
    saved = {k: sys.modules.get(k) for k in ("_dref", "_dref.util", "_dref.analysis", "_dref.analysis.delay")}
    sys.modules.update({"_dref": pkg, "_dref.util": util, "_dref.analysis": ana, "_dref.analysis.delay": dmod})
    opt_ns = {"__name__": "_dref.analysis.delayopt", "__package__": "_dref.analysis"}
    exec(compile(_lines("draco/analysis/delayopt.py", 1, 10**6), "draco/analysis/delayopt.py", "exec"), opt_ns)

    out = types.SimpleNamespace(**{k: v for k, v in ns.items() if callable(v)})
    out.delayopt = types.SimpleNamespace(**{k: v for k, v in opt_ns.items() if callable(v)})
    out.kernels = kernels
    out.cut_data = cut_data
    out.calculate_delays = calculate_delays
    out.window_generalised = tools.window_generalised
    out.random = random
    return out
