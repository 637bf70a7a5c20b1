"""Host emulation of the int8 tensor-core GEMM of ``draco_b200/csrc/ozaki_gemm.cuh`` -- TEST INFRASTRUCTURE, like the
rest of ``oracle/``: only ``tests/`` and ``scratch/`` import it; the product never does.

This is not a restatement of reference code (the reference forms these products with numpy's ``@``,
``draco/analysis/powerspec.py:288-304``); it restates the DEVICE kernel's integer arithmetic exactly, so that the kernel can
be held to it bit for bit (``tests/test_gpu_ozaki_gemm.py``), and it measures what that arithmetic costs in accuracy
against extended precision (``tests/test_ozaki_emulation.py``, ``scratch/ozaki_study.py``).

Arithmetic (``oz_split_kernel`` / ``oz_gemm_kernel``).  Per row and per exponent block of 4 stages x 64 columns (the last
block of a row may hold 2 stages): ``x = 2^e m``, ``|m| <= 1``, ``e`` from ``frexp`` of the block's row maximum;
``m = sum_s q_s 2^-(6 + 7 s)`` with ``q_0 = rint(64 m)`` and ``q_s = rint(128 (previous residual))``, all in ``[-64, 64]``.
``D_d = sum_{s + t = d} Q_s R_t^T`` over the block's columns in int32; ``hi = ((D0 128 + D1) 128 + D2) 128 + D3`` and ``lo``
likewise from ``D4 ..`` in int64; block value ``hi 2^-33 + lo 2^-(12 + 7 (S - 1))`` times ``2^(e_i + f_j)``; blocks added in
ascending order in fp64.
"""
from __future__ import annotations

import numpy as np

STAGE = 64
BLOCK_STAGES = 4
# a row block with a NaN, an infinity or a magnitude >= 2^960 carries this exponent and no slices: every output that touches it
# is NaN (a non-finite operand stays loud); a block below 2^-960 counts as zero
E_NONFINITE = 1 << 20
E_LIMIT = 960


def split_block(X, nslice=8):
    """(e, q): ``X[i, k] ~ 2^e[i] sum_s q[s, i, k] 2^-(6 + 7 s)``; q integer-valued float64 in [-64, 64]."""
    bad = ~np.isfinite(X).all(axis=1)
    with np.errstate(invalid="ignore"):
        amax = np.fmax.reduce(np.abs(X), axis=1, initial=0.0)  # fmax: NaNs are skipped, as on the device
    f, x = np.frexp(np.where(np.isfinite(amax), amax, 0.0))  # amax = f 2^x, f in [0.5, 1)
    e = np.where(amax > 0, np.where(f == 0.5, x - 1, x), 0).astype(np.int64)
    bad |= e > E_LIMIT
    zero = bad | (e < -E_LIMIT)
    e = np.where(bad, E_NONFINITE, np.where(zero, 0, e))
    with np.errstate(invalid="ignore", over="ignore"):
        r = np.where(zero[:, None], 0.0, X * np.exp2(6.0 - np.where(zero, 0, e))[:, None])  # exact
    q = np.empty((nslice,) + X.shape)
    for s in range(nslice):
        q[s] = np.rint(r)
        r = (r - q[s]) * 128.0  # exact
    assert np.abs(q).max(initial=0.0) <= 64
    return e, q


def block_edges(K, block_columns=STAGE * BLOCK_STAGES):
    """Column ranges of the exponent blocks of a K-column operand (K a multiple of 128)."""
    assert K % 128 == 0
    return [(k0, min(K, k0 + block_columns)) for k0 in range(0, K, block_columns)]


def accumulators(qa, qb, nslice):
    """``D[d] = sum_{s + t = d} qa[s] qb[t]^T`` (exact: the partial sums are integers below 2^24 in float64)."""
    D = np.zeros((nslice, qa.shape[1], qb.shape[1]))
    for s in range(nslice):
        for t in range(nslice - s):
            D[s + t] += qa[s] @ qb[t].T
    assert np.abs(D).max(initial=0.0) < 2**31
    return D.astype(np.int64)


def gemm_nt(A, B, nslice=8, block_columns=STAGE * BLOCK_STAGES, return_accumulators=False):
    """``A B^T`` exactly as the device forms it (bit for bit for ``nslice`` in 5..8)."""
    A, B = np.asarray(A, dtype=np.float64), np.asarray(B, dtype=np.float64)
    out = np.zeros((A.shape[0], B.shape[0]))
    accs = []
    lo_scale = 2.0 ** -(12 + 7 * (nslice - 1))
    for k0, k1 in block_edges(A.shape[1], block_columns):
        ea, qa = split_block(A[:, k0:k1], nslice)
        eb, qb = split_block(B[:, k0:k1], nslice)
        D = accumulators(qa, qb, nslice)
        hi = D[0].copy()
        for d in range(1, 4):
            hi = hi * 128 + D[d]
        lo = np.zeros_like(hi)
        for d in range(4, nslice):
            lo = lo * 128 + D[d]
        comb = hi.astype(np.float64) * 2.0**-33 + lo.astype(np.float64) * lo_scale
        es = ea[:, None] + eb[None, :]
        nonfinite = es >= E_NONFINITE // 2
        out += np.where(nonfinite, np.nan, comb * np.exp2(np.where(nonfinite, 0, es).astype(np.float64)))
        accs.append(D)
    return (out, accs) if return_accumulators else out


def exact_nt(A, B):
    """``A B^T`` in 80-bit extended precision (64-bit mantissa: 2^-11 of an fp64 ulp)."""
    Al, Bl = np.asarray(A).astype(np.longdouble), np.asarray(B).astype(np.longdouble)
    out = np.empty((Al.shape[0], Bl.shape[0]), dtype=np.longdouble)
    for i in range(Al.shape[0]):
        out[i] = (Al[i][None, :] * Bl).sum(axis=1)
    return out


def wide_range_matrix(rng, rows, cols, row_decades=12, entry_decades=6):
    """Rows whose scales span ``row_decades`` and whose entries span ``entry_decades``: what a Cholesky panel of a
    windowed system looks like to the split; a few exact zeros, powers of two and one all-zero row included."""
    X = rng.uniform(-1, 1, (rows, cols)) * np.exp2(rng.integers(-int(3.32 * entry_decades), 1, (rows, cols)))
    X *= np.exp2(rng.integers(-int(1.66 * row_decades), int(1.66 * row_decades) + 1, rows))[:, None]
    X[rng.integers(0, rows, 8), rng.integers(0, cols, 8)] = 0.0
    X[rng.integers(0, rows, 4), rng.integers(0, cols, 4)] = 2.0 ** rng.integers(-10, 10, 4)
    if rows > 3:
        X[3] = 0.0
    return X
