"""Host emulation of the int8 tensor-core GEMM (oracle/ozaki_emulation.py): what the split arithmetic costs in accuracy.
The device kernel is held to this emulation bit for bit in tests/test_gpu_ozaki_gemm.py; here the emulation itself is held
to extended precision."""
import numpy as np
import pytest

from oracle import ozaki_emulation as oe


def test_split_reconstructs_the_operand_to_55_bits_below_the_row_maximum():
    rng = np.random.default_rng(1)
    X = oe.wide_range_matrix(rng, 32, 256)
    e, q = oe.split_block(X, 8)
    assert q.shape == (8, 32, 256) and np.abs(q).max() <= 64 and (q == np.rint(q)).all()
    rec = sum(q[s] * 2.0 ** -(6 + 7 * s) for s in range(8)) * np.exp2(e.astype(float))[:, None]
    rowmax = np.abs(X).max(axis=1, keepdims=True)
    assert (np.abs(rec - X) <= rowmax * 2.0**-55 + 0.0).all()
    assert (e[np.abs(X).max(axis=1) == 0] == 0).all() and not q[:, 3].any()  # the all-zero row
    # a row whose maximum is an exact power of two keeps |m| <= 1 with the smaller exponent
    P = np.zeros((1, 128))
    P[0, 5] = 4.0
    e, q = oe.split_block(P, 8)
    assert e[0] == 2 and q[0, 0, 5] == 64 and not q[1:].any()


@pytest.mark.parametrize("K", [128, 256, 384, 1024])
def test_emulated_product_is_of_fp64_quality(K):
    """Error of the 8-slice product against 80-bit arithmetic, relative to the product of the row maxima (the split is a
    fixed-point format under each row's block maximum, so that -- not |A||B|^T -- is its natural scale)."""
    rng = np.random.default_rng(K)
    A, B = rng.standard_normal((128, K)), rng.standard_normal((64, K))
    ex = oe.exact_nt(A, B)
    bound = np.abs(A) @ np.abs(B).T
    err_oz = (np.abs(oe.gemm_nt(A, B) - ex) / bound).max()
    err_dgemm = (np.abs(A @ B.T - ex) / bound).max()
    assert err_oz <= 4e-16 and err_oz <= 2 * err_dgemm  # measured: 0.7e-16 against 4e-16 for numpy's DGEMM
    A, B = oe.wide_range_matrix(rng, 128, K), oe.wide_range_matrix(rng, 64, K)
    ex = oe.exact_nt(A, B)
    scale = np.abs(A).max(axis=1)[:, None] * np.abs(B).max(axis=1)[None, :]
    ok = scale > 0
    assert (np.abs(oe.gemm_nt(A, B) - ex)[ok] / scale[ok]).max() <= 1e-14  # measured 1e-15 at K = 384


def test_fewer_slices_lose_seven_bits_each():
    rng = np.random.default_rng(5)
    A, B = rng.standard_normal((128, 256)), rng.standard_normal((64, 256))
    ex = oe.exact_nt(A, B)
    errs = [float(np.abs(oe.gemm_nt(A, B, nslice=s) - ex).max()) for s in (6, 7, 8)]
    assert errs[0] > 30 * errs[1] > 30 * 30 * errs[2] / 40  # ~128 x per slice until the fp64 rounding of the sum shows
    assert errs[2] < 1e-13


def test_exponent_blocks_follow_the_device_layout():
    assert oe.block_edges(128) == [(0, 128)]
    assert oe.block_edges(384) == [(0, 256), (256, 384)]
    assert oe.block_edges(1024) == [(0, 256), (256, 512), (512, 768), (768, 1024)]


def test_non_finite_operands_stay_loud_and_tiny_blocks_count_as_zero():
    """fmax() skips NaNs and a NaN converts to the int8 value 0: without the sentinel exponent a NaN would silently count as
    zero.  A row block with a NaN / infinity / magnitude >= 2^960 makes every output that touches it NaN."""
    rng = np.random.default_rng(11)
    A, B = rng.standard_normal((128, 512)), rng.standard_normal((64, 512))
    A[5, 300] = np.nan
    B[7, 10] = np.inf
    A[9, 0] = 2.0**990
    A[11] = 2.0**-1000 * rng.standard_normal(512)
    out = oe.gemm_nt(A, B)
    assert np.isnan(out[5]).all() and np.isnan(out[:, 7]).all() and np.isnan(out[9]).all()
    rest = np.ones(out.shape, dtype=bool)
    rest[[5, 9]] = False
    rest[:, 7] = False
    assert np.isfinite(out[rest]).all() and not out[11][rest[11]].any()
    clean = oe.gemm_nt(np.delete(A, [5, 9, 11], axis=0)[:128 - 3], np.delete(B, 7, axis=0))
    np.testing.assert_array_equal(np.delete(np.delete(out, [5, 9, 11], axis=0), 7, axis=1), clean)
    e, q = oe.split_block(A[:, :256], 8)
    assert e[9] == oe.E_NONFINITE and not q[:, 9].any() and e[11] == 0 and not q[:, 11].any() and e[5] < 8
    e, q = oe.split_block(A[:, 256:], 8)  # the NaN of row 5 sits in the second exponent block
    assert e[5] == oe.E_NONFINITE and not q[:, 5].any() and e[9] < 8
