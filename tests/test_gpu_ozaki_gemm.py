"""The int8 tensor-core GEMM (csrc/ozaki_gemm.cuh: tcgen05.mma kind::i8, TMEM accumulators) through the C ABI
(dgwiener_gemm_ozaki) against the host emulation of its arithmetic: integer work, so the bar is BIT FOR BIT."""
import numpy as np
import pytest

from draco_b200 import _lib, powerspec
from oracle import ozaki_emulation as oe

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K,batch", [(128, 64, 128, 1), (128, 64, 384, 1), (256, 192, 512, 2), (128, 128, 1280, 1)])
def test_device_product_equals_the_emulation_bit_for_bit(M, N, K, batch):
    """K = 128: one exponent block of two stages; 384: four + two stages; 1280: five blocks through the two-stage ring;
    several row / column tiles and a batch; rows spanning 24 decades, exact zeros, powers of two, an all-zero row."""
    rng = np.random.default_rng(M + N + K)
    A = np.stack([oe.wide_range_matrix(rng, M, K) for _ in range(batch)])
    B = np.stack([oe.wide_range_matrix(rng, N, K) for _ in range(batch)])
    t = {}
    got = powerspec.gemm_nt_int8(A, B, timing=t)
    assert got.shape == (batch, M, N) and t["ms_device"] > 0
    for b in range(batch):
        want = oe.gemm_nt(A[b], B[b])
        np.testing.assert_array_equal(got[b], want)
        assert not got[b, 3].any()  # the all-zero row of A


def test_device_product_is_of_fp64_quality():
    rng = np.random.default_rng(9)
    A, B = rng.standard_normal((256, 1024)), rng.standard_normal((128, 1024))
    got = powerspec.gemm_nt_int8(A, B)
    ex = oe.exact_nt(A, B)
    bound = np.abs(A) @ np.abs(B).T
    assert (np.abs(got - ex) / bound).max() <= 4e-16  # numpy's DGEMM on the same operands: ~5e-16
    np.testing.assert_array_equal(got, powerspec.gemm_nt_int8(A[None], B[None])[0])


def test_non_finite_operands_give_nan_not_zero():
    """A NaN / infinity in an operand must not be swallowed by the split (fmax skips NaNs, NaN -> int8 is 0): every output
    that touches the row block is NaN; everything else is bit for bit the emulation's."""
    rng = np.random.default_rng(12)
    A, B = rng.standard_normal((128, 512)), rng.standard_normal((64, 512))
    A[5, 300] = np.nan
    B[7, 10] = np.inf
    A[9, 0] = 2.0**990
    A[11] = 2.0**-1000 * rng.standard_normal(512)
    got = powerspec.gemm_nt_int8(A, B)
    assert np.isnan(got[5]).all() and np.isnan(got[:, 7]).all() and np.isnan(got[9]).all()
    np.testing.assert_array_equal(got, oe.gemm_nt(A, B))  # NaNs in the same places, the rest equal


def test_bad_shapes_are_refused():
    with pytest.raises(ValueError):
        powerspec.gemm_nt_int8(np.zeros((100, 128)), np.zeros((64, 128)))
    with pytest.raises(ValueError):
        powerspec.gemm_nt_int8(np.zeros((128, 128)), np.zeros((64, 256)))
    lib = powerspec.load_wiener()
    z = np.zeros(8)
    assert lib.dgwiener_gemm_ozaki(0, z.ctypes.data, z.ctypes.data, z.ctypes.data, 128, 64, 100, 1, None) == -5
    assert lib.dgwiener_gemm_ozaki(0, None, None, None, 0, 64, 128, 1, None) == 0  # empty product
    assert isinstance(_lib.DGibbsError("x"), Exception)
