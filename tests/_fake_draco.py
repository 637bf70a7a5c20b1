"""A stand-in for the module ``draco.analysis.delay`` (the reference package cannot be imported here: caput, cora and its
Cython extension are absent).  Its task classes own ``process`` / ``_prepare_inputs`` / ``_create_output`` /
``_cut_data`` like the reference's, and their ``_evaluate`` / ``_estimator`` stand for the reference's per-baseline CPU
loop: they raise, so a test that gets a spectrum out of them proves ``draco_b200.install()`` rebound the device path."""
import sys
import types


class CpuLoopRan(AssertionError):
    pass


def make(name="draco.analysis.delay"):
    from draco_b200 import delay as ours

    def cpu_loop(self, *a, **k):
        raise CpuLoopRan("the reference's CPU loop ran")

    mod = types.ModuleType(name)
    for cname in ("DelayPowerSpectrumGibbs", "DelaySpectrumWienerFilter", "DelaySpectrumFFT", "DelayPowerSpectrumNRML"):
        base = getattr(ours, cname)

        def process(self, *a, _base=base, **k):
            self.reference_process_calls = getattr(self, "reference_process_calls", 0) + 1
            return _base.process(self, *a, **k)

        ns = {"_evaluate": cpu_loop, "_estimator": cpu_loop, "process": process, "__module__": name,
              # not part of the reference's classes: must come from install()
              "_evaluate_device_cut": None, "_device_seed": None, "_device": None}
        setattr(mod, cname, type(cname, (base,), ns))
    mod.DelayPowerSpectrumStochasticEstimator = mod.DelayPowerSpectrumGibbs
    return mod


def inject(monkeypatch, name="draco.analysis.delay"):
    mod = make(name)
    parts = name.split(".")
    for i in range(1, len(parts)):
        pkg = ".".join(parts[:i])
        if pkg not in sys.modules:
            monkeypatch.setitem(sys.modules, pkg, types.ModuleType(pkg))
    monkeypatch.setitem(sys.modules, name, mod)
    return mod
