"""The C-ABI boundary without a GPU: the library builds/loads, exports every symbol the
header declares, struct layouts agree, and compute entry points fail loudly (no fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from draco_b200 import _build, _lib

    _build.build()
    return _lib.load()


def _declared():
    hdr = open(os.path.join(ROOT, "include", "dgibbs.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(dgibbs_[a-z_]+)\s*\(", hdr)))


def test_header_symbols_exported(lib):
    names = _declared()
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dgibbs.h but not exported"


def test_binding_lists_every_symbol():
    from draco_b200 import _lib

    assert sorted(_lib.EXPORTED) == _declared()


def test_struct_sizes(lib):
    from draco_b200 import _lib

    assert lib.dgibbs_sizeof(0) == C.sizeof(_lib.Problem)
    assert lib.dgibbs_sizeof(1) == C.sizeof(_lib.Result)
    assert lib.dgibbs_sizeof(2) == C.sizeof(_lib.Timing)
    assert lib.dgibbs_version() == 100


def test_no_cpu_fallback(lib):
    """Without a CUDA device the product path must raise, never compute on the CPU."""
    import torch

    from draco_b200 import _lib

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_lib.DGibbsError):
        _lib.Handle(0)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "draco_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f


def test_nrml_library_exports_header_symbols(lib):
    """libdgnrml.so exports everything include/dgnrml.h declares and fails loudly without a GPU."""
    import numpy as np
    import torch

    from draco_b200 import _lib, delayopt

    hdr = open(os.path.join(ROOT, "include", "dgnrml.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(dgnrml_[a-z_]+)\s*\(", hdr)))
    nl = delayopt.load_nrml()
    assert names == sorted(delayopt.NRML_EXPORTED)
    for n in names:
        assert hasattr(nl, n), f"{n} declared in include/dgnrml.h but not exported"
    assert lib.dgibbs_sizeof(4) == C.sizeof(_lib.Cut)
    if not torch.cuda.is_available():
        with pytest.raises(_lib.DGibbsError):
            delayopt.LogLikePS(np.eye(3, dtype=complex), np.ones((3, 4), dtype=complex), np.ones(3), 5)


def test_wiener_operator_entry_points():
    """The dgwiener_* entry points of libdgnrml.so (include/dgwiener.h): exported, struct layout agreed,
    and a loud failure -- not a CPU fallback -- when there is no GPU."""
    import numpy as np
    import torch

    from draco_b200 import _lib, powerspec

    hdr = open(os.path.join(ROOT, "include", "dgwiener.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(dgwiener_[a-z_]+)\s*\(", hdr)))
    wl = powerspec.load_wiener()
    assert names == sorted(powerspec.WIENER_EXPORTED)
    for n in names:
        assert hasattr(wl, n), f"{n} declared in include/dgwiener.h but not exported"
    assert wl.dgwiener_sizeof() == C.sizeof(powerspec.WienerProblem)
    if not torch.cuda.is_available():
        F = np.ones((4, 2), dtype=complex)
        with pytest.raises(_lib.DGibbsError):
            powerspec.wiener_delay_operator(F, np.ones(2), np.ones(4), np.eye(4)[None], np.eye(4)[None],
                                            np.ones((1, 1, 4)), np.ones((1, 4)))


def test_int8_gemm_wrapper_validates_shapes_before_the_device():
    """powerspec.gemm_nt_int8 (dgwiener_gemm_ozaki): shape errors are Python errors; without a GPU the call itself fails loudly."""
    import numpy as np
    import torch

    from draco_b200 import _lib, powerspec

    with pytest.raises(ValueError):
        powerspec.gemm_nt_int8(np.zeros((100, 128)), np.zeros((64, 128)))  # M % 128
    with pytest.raises(ValueError):
        powerspec.gemm_nt_int8(np.zeros((128, 128)), np.zeros((60, 128)))  # N % 64
    with pytest.raises(ValueError):
        powerspec.gemm_nt_int8(np.zeros((128, 192)), np.zeros((64, 192)))  # K % 128
    with pytest.raises(ValueError):
        powerspec.gemm_nt_int8(np.zeros((2, 128, 128)), np.zeros((3, 64, 128)))  # batch mismatch
    if not torch.cuda.is_available():
        with pytest.raises(_lib.DGibbsError):
            powerspec.gemm_nt_int8(np.zeros((128, 128)), np.zeros((64, 128)))


def test_window_ids_follow_the_header():
    """Window names of draco/util/tools.py:547-601 -> window_id of include/dgibbs.h, 'tukey-0.X' included."""
    from draco_b200 import _lib

    hdr = open(os.path.join(ROOT, "include", "dgibbs.h")).read()
    enum = dict(re.findall(r"DGIBBS_WINDOW_([A-Z_]+)\s*=\s*(-?\d+)", hdr))
    for name, wid in _lib.WINDOW_IDS.items():
        assert int(enum["NONE" if name is None else name.upper()]) == wid
    assert "#define DGIBBS_WINDOW_TUKEY_BASE (1 << 20)" in hdr
    assert _lib.WINDOW_IDS["tukey-0.5"] == (1 << 20) + 500000
    assert _lib.WINDOW_IDS["tukey-0.25"] == (1 << 20) + 250000
    assert "tukey-0.3" in _lib.WINDOW_IDS and "tukey-1.0" in _lib.WINDOW_IDS
    for bad in ("tukey", "tukey-", "tukey-0", "tukey-1.5", "tukey-0.1234567", "kaiser", "tukey-x"):
        assert bad not in _lib.WINDOW_IDS
        with pytest.raises(KeyError):
            _lib.WINDOW_IDS[bad]
