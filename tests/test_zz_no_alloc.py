"""``dgibbs_delay_fft_noalloc``: the plain FFT delay transform for callers that cannot allocate device memory (its window
and twiddle tables live in a static __device__ array of the library).  No counterpart in the reference.
Called through the C ABI (ctypes) and held to the existing entry point
``dgibbs_delay_fft`` on consecutive calls that change (N, window) -- the shared scratch keeps the last call's tables.

Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-13  # of the row maximum: the tolerance of tests/test_gpu_transforms.py (the kernels are the same: expect 0)


def _rows(nrow, N, seed):
    import torch

    r = np.random.default_rng(seed)
    x = r.standard_normal((nrow, N)) + 1j * r.standard_normal((nrow, N))
    return torch.from_numpy(np.ascontiguousarray(x).view(np.float64).reshape(nrow, N, 2)).to("cuda:0")


def test_noalloc_entry_matches_existing_entry_on_consecutive_calls():
    import torch

    from draco_b200 import _lib, gibbs

    h = gibbs.get_handle(0)
    lib = h.lib
    # set-up only (nrows = 0): the once-per-device cudaGetSymbolAddress look-up
    assert lib.dgibbs_delay_fft_noalloc(h._h, 0, 1024, 5, 0, None, None, 1) == 0
    cases = [(96, 1024, "nuttall", False), (41, 256, "hann", True), (96, 1024, "nuttall", False), (7, 4096, None, True),
             (300, 8, "blackman_harris", False), (96, 1024, "blackman", True)]
    ins = [_rows(nrow, N, 10 + i) for i, (nrow, N, _, _) in enumerate(cases)]
    outs = [torch.empty_like(x) for x in ins]
    want = [gibbs.fft_rows(x, w, shift=s, handle=h) for x, (_, _, w, s) in zip(ins, cases)]
    torch.cuda.synchronize()
    free_before = None
    for i, ((nrow, N, w, s), x, o) in enumerate(zip(cases, ins, outs)):
        if i == 2:  # every kernel instantiation below has been loaded by now
            free_before = torch.cuda.mem_get_info(0)[0]
        o.fill_(float("nan"))
        torch.cuda.synchronize()
        rc = lib.dgibbs_delay_fft_noalloc(h._h, nrow, N, _lib.WINDOW_IDS[w], int(s), C.c_void_p(x.data_ptr()),
                                          C.c_void_p(o.data_ptr()), 1)
        assert rc == 0, lib.dgibbs_last_error(h._h)
        got, ref = o.cpu().numpy(), want[i].cpu().numpy()
        assert np.isfinite(got).all()
        assert np.max(np.abs(got - ref)) <= TOL * np.max(np.abs(ref)), (i, N, w)
        assert np.array_equal(got, ref), (i, N, w)  # same kernels, same tables: bit-identical
        if i == 2:
            assert free_before - torch.cuda.mem_get_info(0)[0] < (1 << 20)  # nothing was allocated on the device
    # asynchronous form + the Python wrapper
    o = torch.empty_like(ins[1])
    gibbs.fft_rows_noalloc(ins[1], o, "hann", shift=True, handle=h, sync=False)
    torch.cuda.synchronize()
    assert np.array_equal(o.cpu().numpy(), want[1].cpu().numpy())


def test_noalloc_entry_rejects_what_the_static_scratch_cannot_hold():
    import torch

    from draco_b200 import gibbs

    h = gibbs.get_handle(0)
    lib = h.lib
    x = torch.zeros((2, 8192, 2), dtype=torch.float64, device="cuda:0")
    o = torch.empty_like(x)
    for N in (8192, 1000, 4, 0):
        rc = lib.dgibbs_delay_fft_noalloc(h._h, 1, N, -1, 0, C.c_void_p(x.data_ptr()), C.c_void_p(o.data_ptr()), 1)
        assert rc == -2
        msg = lib.dgibbs_last_error(h._h).decode()
        assert "4096" in msg and "dgibbs_delay_fft" in msg, msg
    assert lib.dgibbs_delay_fft_noalloc(h._h, 1, 1024, -1, 0, C.c_void_p(x.data_ptr()), C.c_void_p(x.data_ptr()), 1) == -2
    assert lib.dgibbs_delay_fft_noalloc(h._h, 1, 1024, 9, 0, C.c_void_p(x.data_ptr()), C.c_void_p(o.data_ptr()), 1) == -2
    with pytest.raises(gibbs.DGibbsError):
        gibbs.fft_rows_noalloc(x, o, None, handle=h)
