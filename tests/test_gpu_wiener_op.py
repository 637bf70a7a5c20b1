"""Device Wiener delay-transform operator (include/dgwiener.h, draco_b200/powerspec.py) against the reference's
golden operators (tests/golden/reference_wiener_op.npz, from powerspec.py:118-369) and against the oracle."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(__file__)
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_golden_wiener_op import WIENER_CASES, make_map, make_operator, make_ringmap, make_wiener_case  # noqa: E402

from draco_b200 import powerspec  # noqa: E402
from oracle import wiener_op_oracle as wo  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "reference_wiener_op.npz"))


def close_c64(got, ref):
    """The container stores complex64: agreement to a few ulp of float32 relative to the operator's scale."""
    scale = np.abs(ref).max()
    np.testing.assert_allclose(got, ref.astype(np.complex64), rtol=2e-5, atol=2e-6 * scale)


@pytest.mark.parametrize("k", range(len(WIENER_CASES)))
def test_task_matches_reference_operator(golden, k):
    """ConstructWienerDelayTransform.process, same YAML parameters, against the reference's own output."""
    name, nfreq, nra, nel, npol, cplx, params = WIENER_CASES[k]
    case = make_wiener_case(nfreq, nra, nel, npol, cplx, 300 + k)
    task = powerspec.ConstructWienerDelayTransform.from_config(params)
    op = task.process(make_ringmap(case))
    ref = golden[f"{name}_filter"]
    got = op.filter[:].local_array
    assert got.dtype == np.complex64 and got.shape == ref.shape
    assert list(op.filter.attrs["axis"]) == ["pol", "ra", "el", "delay", "freq"]
    np.testing.assert_array_equal(op.index_map["delay"], golden[f"{name}_delay"])
    for attr in ("window", "window_lower_freq", "window_upper_freq"):
        assert op.attrs[attr] == params.get(attr, "uniform" if attr == "window" else None)
    close_c64(got, ref)
    assert not got[:, nra - 1, nel - 1].any()  # the pixel without a valid channel
    dead = case["weight"].transpose(0, 2, 3, 1) == 0
    assert not got[np.broadcast_to(dead[:, :, :, None, :], got.shape)].any()
    assert all(t["launches"] > 0 for t in task.last_timing)


def random_problem(F, nra, nel, cplx_filter, cplx_cov, seed, window="blackman"):
    case = make_wiener_case(F, nra, nel, 1, True, seed)
    freq = case["freq"]
    win = wo.get_window(freq, window)
    tau, ntau = wo.delay_axis(freq, win)
    S = wo.get_prior(tau, 2.8e-5, 0.02)
    filt = np.ascontiguousarray(case["filter"][0].transpose(2, 0, 1))
    cov = np.ascontiguousarray(case["freq_cov"][0].transpose(2, 0, 1))
    if not cplx_filter:
        filt = np.ascontiguousarray(filt.real)
    if not cplx_cov:
        cov = np.ascontiguousarray(cov.real)
    weight = np.ascontiguousarray(case["weight"][0].transpose(2, 1, 0))
    beam = np.ascontiguousarray(case["dirty_beam_power"][0, 0].T)
    return freq, weight, beam, filt, cov, win, S, tau, ntau


@pytest.mark.parametrize("solver", [1, 2, 3, 4])
@pytest.mark.parametrize("cplx_filter,cplx_cov", [(False, False), (True, True), (True, False), (False, True)])
def test_operator_matches_oracle(solver, cplx_filter, cplx_cov):
    """Both library factorisation paths (batched / per-matrix on several streams), the own DMMA kernels (3) and the own
    kernels with the contractions on the int8 tensor cores (4: Ozaki split, tcgen05 + TMEM); every dtype combination, F
    not a multiple of anything, several chunks of pixels (tiny workspace budget)."""
    freq, weight, beam, filt, cov, win, S, tau, ntau = random_problem(52, 7, 3, cplx_filter, cplx_cov, 77)
    ref = wo.construct_operator(freq, weight, beam, filt, cov, win, S, tau, ntau)
    t = {}
    got = powerspec.wiener_delay_operator(wo.fourier_operator(freq, tau, ntau), S, win, filt, cov, weight, beam,
                                          solver=solver, workspace_bytes=3 * 52 * 52 * 100, timing=t)
    close_c64(got, ref)
    assert t["launches"] > 0 and t["flop"] > 0


def test_auto_solver_takes_the_int8_contractions_from_256_channels():
    """solver = 0 picks the int8 tensor-core contractions (solver 4) from F = 256 and the DMMA ones (3) below; both agree
    with the oracle at complex64 resolution."""
    freq, weight, beam, filt, cov, win, S, tau, ntau = random_problem(256, 3, 1, True, False, 79)
    ref = wo.construct_operator(freq, weight, beam, filt, cov, win, S, tau, ntau)
    Fm = wo.fourier_operator(freq, tau, ntau)
    t0, t3, t4 = {}, {}, {}
    got0 = powerspec.wiener_delay_operator(Fm, S, win, filt, cov, weight, beam, timing=t0)
    got3 = powerspec.wiener_delay_operator(Fm, S, win, filt, cov, weight, beam, solver=3, timing=t3)
    got4 = powerspec.wiener_delay_operator(Fm, S, win, filt, cov, weight, beam, solver=4, timing=t4)
    for got in (got0, got3, got4):
        close_c64(got, ref)
    assert t0["launches"] == t4["launches"] > t3["launches"]  # the split kernels of the int8 path
    np.testing.assert_array_equal(got0, got4)
    small = random_problem(52, 2, 1, False, False, 80)
    ts0, ts3 = {}, {}
    args = (wo.fourier_operator(small[0], small[7], small[8]), small[6], small[5], small[3], small[4], small[1], small[2])
    a = powerspec.wiener_delay_operator(*args, timing=ts0)
    b = powerspec.wiener_delay_operator(*args, solver=3, timing=ts3)
    assert ts0["launches"] == ts3["launches"]
    np.testing.assert_array_equal(a, b)


def test_int8_contractions_with_a_short_last_exponent_block():
    """F = 160: the padded real width is 384 = one exponent block of four 64-column stages and one of two; two row tiles,
    six column tiles, lower-only tiles of A (csrc/ozaki_gemm.cuh)."""
    freq, weight, beam, filt, cov, win, S, tau, ntau = random_problem(160, 4, 2, True, True, 81, window="uniform")
    ref = wo.construct_operator(freq, weight, beam, filt, cov, win, S, tau, ntau)
    Fm = wo.fourier_operator(freq, tau, ntau)
    got4 = powerspec.wiener_delay_operator(Fm, S, win, filt, cov, weight, beam, solver=4)
    got3 = powerspec.wiener_delay_operator(Fm, S, win, filt, cov, weight, beam, solver=3)
    close_c64(got4, ref)
    close_c64(got3, ref)
    # both paths compute in (at least) fp64: they differ from each other far less than either differs from complex64 rounding
    assert np.abs(got4.astype(np.complex128) - got3).max() <= 4e-7 * np.abs(ref).max()


def test_strided_output_and_larger_band():
    """F = 160 (per-matrix path by default), the output written into a [pol, ra, el, delay, freq] slab."""
    freq, weight, beam, filt, cov, win, S, tau, ntau = random_problem(160, 5, 2, False, False, 78, window="uniform")
    ref = wo.construct_operator(freq, weight, beam, filt, cov, win, S, tau, ntau)
    slab = np.full((2, 5, 2, tau.size, 160), 7.0, dtype=np.complex64)
    powerspec.wiener_delay_operator(wo.fourier_operator(freq, tau, ntau), S, win, filt, cov, weight, beam, out=slab[1])
    close_c64(slab[1], ref)
    assert (slab[0] == 7.0).all()


def test_not_positive_definite_raises():
    """scipy.linalg.cho_factor raises LinAlgError in the reference (powerspec.py:296); so does the device path."""
    from test_wiener_op_oracle import make_indefinite

    name, nfreq, nra, nel, npol, cplx, params = WIENER_CASES[0]
    case = make_wiener_case(nfreq, nra, nel, npol, cplx, 300)
    make_indefinite(case)
    with pytest.raises(np.linalg.LinAlgError):
        powerspec.ConstructWienerDelayTransform.from_config(params).process(make_ringmap(case))


def test_empty_and_bad_arguments():
    freq, weight, beam, filt, cov, win, S, tau, ntau = random_problem(16, 2, 1, False, False, 5)
    Fm = wo.fourier_operator(freq, tau, ntau)
    out = powerspec.wiener_delay_operator(Fm, S, win, filt[:0], cov[:0], weight[:, :0], beam)
    assert out.shape == (0, 1, tau.size, 16)
    with pytest.raises(ValueError):
        powerspec.wiener_delay_operator(Fm, S, win, filt, cov[:, :-1], weight, beam)
    with pytest.raises(ValueError):
        powerspec.wiener_delay_operator(Fm, S[:-1], win, filt, cov, weight, beam)


# ------------------------------------------------------------------------------------------------ apply
@pytest.mark.parametrize("k", range(len(WIENER_CASES)))
def test_apply_task_matches_reference(golden, k):
    """ApplyWienerDelayTransform.process against the reference's own output (powerspec.py:372-458)."""
    name, nfreq, nra, nel, npol, cplx, params = WIENER_CASES[k]
    case = make_wiener_case(nfreq, nra, nel, npol, cplx, 300 + k)
    rm = make_ringmap(case)
    rm.map[:].local_array[:] = make_map(case, 700 + k)
    op = make_operator(case, golden[f"{name}_filter"], golden[f"{name}_delay"], params)
    out = powerspec.ApplyWienerDelayTransform().process(rm, op)
    ref = golden[f"{name}_apply_spectrum"]
    got = out.spectrum[:].local_array
    assert got.shape == ref.shape == (npol * nel, nra, golden[f"{name}_delay"].size) and got.dtype == np.complex128
    np.testing.assert_allclose(got, ref, rtol=1e-10, atol=1e-13 * np.abs(ref).max())
    gw, rw = out.weight[:].local_array, golden[f"{name}_apply_weight"]
    assert gw.dtype == np.float32
    np.testing.assert_allclose(gw, rw, rtol=2e-6)
    assert (gw == 0).sum() == (rw == 0).sum()
    assert list(out.attrs["baseline_axes"]) == ["pol", "el"]
    assert out.attrs["window_los"] == params.get("window", "uniform")
    np.testing.assert_array_equal(out.attrs["freq"], case["freq"])
    np.testing.assert_array_equal(out.index_map["el"], rm.index_map["el"])


@pytest.mark.parametrize("F", [33, 96, 258])
def test_apply_matches_oracle(F):
    """Odd and even channel counts (scalar and 16-byte loads), several chunks of a host-resident operator, a strided
    operator slab, zero weights."""
    r = np.random.default_rng(F)
    nra, nel, nd = 6, 3, F // 2 + 1
    slab = (r.standard_normal((2, nra, nel, nd, F)) + 1j * r.standard_normal((2, nra, nel, nd, F))).astype(np.complex64)
    op = slab[1]
    m = r.standard_normal((F, nra, nel))
    w = r.uniform(0.5, 2.0, (F, nra, nel))
    w[r.random(w.shape) < 0.2] = 0.0
    w[:, 2, 1] = 0.0
    ref_s, ref_w = wo.apply_operator(op, m, w)
    t = {}
    got_s, got_w = powerspec.apply_wiener_operator(op, m, w, workspace_bytes=2 * nel * nd * F * 8, timing=t)
    np.testing.assert_allclose(got_s, ref_s, rtol=1e-10, atol=1e-13 * np.abs(ref_s).max())
    np.testing.assert_allclose(got_w, ref_w, rtol=2e-6)
    assert not got_w[1, 2].any() and t["launches"] == 3


def test_construct_then_apply_on_device():
    """The operator written into a CUDA tensor (never copied to the host) gives the same spectra as the host path."""
    import torch

    freq, weight, beam, filt, cov, win, S, tau, ntau = random_problem(48, 5, 2, True, False, 91)
    Fm = wo.fourier_operator(freq, tau, ntau)
    host = powerspec.wiener_delay_operator(Fm, S, win, filt, cov, weight, beam)
    dev = torch.zeros((5, 2, tau.size, 48), dtype=torch.complex64, device="cuda:0")
    powerspec.wiener_delay_operator(Fm, S, win, filt, cov, weight, beam, out=dev)
    np.testing.assert_array_equal(dev.cpu().numpy(), host)
    r = np.random.default_rng(3)
    m = r.standard_normal((48, 5, 2))
    w = np.ascontiguousarray(weight.transpose(2, 1, 0))
    s0, w0 = powerspec.apply_wiener_operator(host, m, w)
    s1, w1 = powerspec.apply_wiener_operator(dev, m, w)
    np.testing.assert_array_equal(s0, s1)
    np.testing.assert_array_equal(w0, w1)
