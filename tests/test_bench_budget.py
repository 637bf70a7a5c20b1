"""bench.py's wall-clock guard: the optional records (e2e_task / strong / diag) may never push the driver's run past
its limit; the headline `value` / `e2e` legs are not subject to it."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_optional_record_runs_only_inside_the_budget():
    b = _bench()
    assert b.budget_allows(420.0, 137.0, 780.0)
    assert not b.budget_allows(700.0, 137.0, 780.0)
    assert b.budget_allows(780.0, 0.0, 780.0)
    assert b.budget_allows(1e9, 1e9, 0.0)  # --budget-s 0 switches the guard off


def test_budget_flag_has_a_default_below_the_drivers_limit(monkeypatch):
    b = _bench()
    monkeypatch.setattr("sys.argv", ["bench.py"])
    args = b.parse()
    assert 0 < args.budget_s <= 800.0
    assert args.steps >= 1 and args.warmup >= 3


def test_int8_side_record_never_raises():
    """Without a GPU the C-ABI call fails loudly; the side record must turn that into text, not into a lost bench line."""
    import torch

    b = _bench()
    rec = b.int8_gemm_record(batch=1)
    assert isinstance(rec, dict)
    if not torch.cuda.is_available():
        assert "error" in rec and "CUDA" in rec["error"]
