"""Ozaki-split fp64 GEMM for the trailing updates of the dense sampler: accuracy study on the host (exact emulation).

VERDICT r01 item 9 (stretch): the only way past the 37 TFLOP/s DMMA roof is to run `acc -= A B^T` on the int8 tensor cores
(tcgen05.mma kind::i8, int32 accumulators in TMEM) with A and B split into signed 7-bit slices.  Before any of that
enters the product it has to keep the cfg2 golden chains inside rtol 1e-6, so this script measures on the HOST what the
split arithmetic does to a real panel and to the whole solve, for cond(Ci) ~ 4e8 (cfg2: Nuttall window, per-frequency
weights).  The int8 x int8 -> int32 arithmetic is emulated exactly (integers < 2^24 in float64 BLAS), so the numbers here
are the numbers the GPU kernel gives (scratch/ozaki_probe.cu checks its accumulators bit for bit against
`ozaki_panel_expected` written here).

    python scratch/ozaki_study.py [--slices 7 8 9] [--full]      # log: profiles/r04a_ozaki_study.log
    python scratch/ozaki_study.py --write-panel gpurun_out/ozaki_panel.bin   # input of scratch/ozaki_probe.cu

Split.  Row i of a panel: x_ik = 2^{e_i} m_ik, |m| <= 1, e_i from the row maximum over the K range of the update;
m = sum_s q_s 2^{-(6 + 7 s)} with q_0 = rint(m 2^6) in [-64, 64] and q_s = rint(residual 2^{6 + 7 s}) in [-64, 64]:
S slices carry 6 + 7 (S - 1) bits below the row maximum.  Product: sum_k x_ik y_jk =
2^{e_i + f_j} sum_{s + t < S} 2^{-(12 + 7 (s + t))} sum_k q_s,ik r_t,jk, the inner sums exact in int32 for K <= 2^16.
S (S + 1) / 2 int8 GEMMs per fp64 GEMM (S = 8: 36).
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np
import scipy.linalg as la

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

TILE = 128


def cfg2_system(seed_baseline=0, burn=2):
    """Ci = F^T N^-1 F + diag(1/S) of one cfg2 baseline with S after `burn` Gibbs iterations (the state the chains live in)."""
    import oracle
    from draco_b200 import synth

    cfg = synth.CONFIGS["cfg2"]
    freq, vis, wgt = synth.make_stack(cfg, nbase=1, base_offset=seed_baseline, threads=1)
    _, chan = oracle.calculate_delays(freq)
    d, w, nzf, _ = oracle.cut_data(vis[0], wgt[0])
    N = cfg.ndelay
    z, FTNih, FTNiF = oracle.whiten_inputs(d, N, w, chan[nzf], cfg.window, False)
    rng = np.random.default_rng(7)
    S = np.full(N, 10.0)
    for _ in range(burn):
        x = oracle.gibbs_oracle._signal_draw_delay_space(S, z, FTNih, FTNiF, N, rng, False)
        S = oracle.gibbs_oracle._spectrum_draw(x, rng, False)
    Ci = FTNiF.copy()
    Ci[np.diag_indices_from(Ci)] += 1.0 / S
    T = z.shape[1]
    rhs = rng.standard_normal((N, T)) * (1.0 / S)[:, None] ** 0.5 + FTNih @ (z + rng.standard_normal(z.shape))
    return Ci, rhs


def split(X, nslice):
    """(e, q): X[i, k] ~ 2^e[i] sum_s q[s, i, k] 2^-(6 + 7 s), q integer-valued float64 in [-64, 64]."""
    amax = np.abs(X).max(axis=1)
    f, x = np.frexp(amax)  # amax = f 2^x, f in [0.5, 1): the exponent rule of split_kernel in scratch/ozaki_probe.cu
    e = np.where(amax > 0, np.where(f == 0.5, x - 1, x), 0).astype(np.float64)
    m = X * np.exp2(-e)[:, None]
    assert np.abs(m).max() <= 1.0
    q = np.empty((nslice,) + X.shape)
    r = m * 64.0
    for s in range(nslice):
        q[s] = np.rint(r)
        r = (r - q[s]) * 128.0  # exact: r - q is a multiple of the ulp of r and < 0.5 in magnitude
    assert np.abs(q).max() <= 64
    return e, q


def ozaki_accumulators(qa, qb, nslice):
    """D[d] = sum_{s + t = d} qa[s] qb[t]^T for d < nslice, exact integers (what the int32 accumulators in TMEM hold)."""
    D = np.zeros((nslice, qa.shape[1], qb.shape[1]))
    for s in range(nslice):
        for t in range(nslice - s):
            D[s + t] += qa[s] @ qb[t].T
    assert np.abs(D).max() < 2**31
    return D


def ozaki_gemm_nt(A, B, nslice, kblock=None):
    """A B^T through the split; kblock: exponents per K block of that width (None: one exponent per row)."""
    if kblock is None:
        ea, qa = split(A, nslice)
        eb, qb = split(B, nslice)
        D = ozaki_accumulators(qa, qb, nslice)
        acc = np.zeros(D.shape[1:])
        for d in range(nslice - 1, -1, -1):  # small terms first
            acc += D[d] * 2.0 ** -(12 + 7 * d)
        return acc * np.exp2(ea)[:, None] * np.exp2(eb)[None, :]
    out = np.zeros((A.shape[0], B.shape[0]))
    for k0 in range(0, A.shape[1], kblock):
        out += ozaki_gemm_nt(A[:, k0:k0 + kblock], B[:, k0:k0 + kblock], nslice)
    return out


def exact_nt(A, B):
    """A B^T in 80-bit extended precision (64-bit mantissa: 2^-11 of an fp64 ulp)."""
    Al, Bl = A.astype(np.longdouble), B.astype(np.longdouble)
    out = np.empty((A.shape[0], B.shape[0]), dtype=np.longdouble)
    for i in range(A.shape[0]):
        out[i] = (Al[i][None, :] * Bl).sum(axis=1)
    return out


def panel_study(L, Ci, slices, log):
    n = L.shape[0]
    i0, j0 = n - TILE, n - 2 * TILE
    A, B = L[i0:i0 + TILE, :j0], L[j0:j0 + TILE, :j0]
    K = j0
    ex = exact_nt(A, B)
    ref_scale = float(np.abs(ex).max())
    bound = (np.abs(A) @ np.abs(B).T)  # |A||B|^T: the componentwise scale of any summation error
    dgemm = A @ B.T
    err = lambda X: np.abs(X.astype(np.longdouble) - ex).astype(np.float64)  # noqa: E731
    log(f"panel update of tile (15, 14): A = L[{i0}:{i0 + TILE}, :{K}], B = L[{j0}:{j0 + TILE}, :{K}], K = {K}")
    log(f"  max |A B^T| = {ref_scale:.3e}; dynamic range of the rows of A: max/min row maxima = "
        f"{np.abs(A).max(axis=1).max() / np.abs(A).max(axis=1).min():.2e}; entries span "
        f"{np.abs(A).max():.2e} .. {np.abs(A)[A != 0].min():.2e}")
    e = err(dgemm)
    log(f"  fp64 DGEMM              : max abs err {e.max():.3e} = {e.max() / ref_scale:.2e} of max|AB^T|, max err / (|A||B|^T) = {(e / bound).max():.2e}")
    for S in slices:
        for kb in (None, 128):
            t0 = time.perf_counter()
            oz = ozaki_gemm_nt(A, B, S, kb)
            e = err(oz)
            log(f"  Ozaki S = {S} ({S * (S + 1) // 2:2d} int8 GEMMs), exponents per {'row' if kb is None else f'row x {kb}-column block'}: "
                f"max abs err {e.max():.3e} = {e.max() / ref_scale:.2e} of max|AB^T|, max err / (|A||B|^T) = {(e / bound).max():.2e}"
                f"   [{time.perf_counter() - t0:.1f} s]")


def blocked_cholesky(Ci, gemm_nt):
    """Left-looking blocked Cholesky (128-wide block columns, the device kernel's order) with the trailing updates
    `acc -= A B^T` through `gemm_nt`; diagonal tiles and triangular solves in fp64 LAPACK."""
    n = Ci.shape[0]
    nb = n // TILE
    L = np.zeros_like(Ci)
    for j in range(nb):
        c0 = j * TILE
        for i in range(j, nb):
            r0 = i * TILE
            acc = Ci[r0:r0 + TILE, c0:c0 + TILE].copy()
            if j > 0:
                acc -= gemm_nt(L[r0:r0 + TILE, :c0], L[c0:c0 + TILE, :c0])
            if i == j:
                L[r0:r0 + TILE, c0:c0 + TILE] = la.cholesky(acc, lower=True)
            else:
                L[r0:r0 + TILE, c0:c0 + TILE] = la.solve_triangular(L[c0:c0 + TILE, c0:c0 + TILE], acc.T, lower=True).T
    return L


def solve_study(Ci, rhs, slices, log):
    """What the chain sees: x = Ci^-1 rhs and the row variances var_t(x) that feed the spectrum draw."""
    fac = la.cho_factor(Ci, lower=False)
    x_ref = la.cho_solve(fac, rhs)
    v_ref = x_ref.var(axis=1)

    def report(name, L):
        x = la.solve_triangular(L.T, la.solve_triangular(L, rhs, lower=True), lower=False)
        v = x.var(axis=1)
        log(f"  {name:<58s}: max rel err of x {np.abs(x - x_ref).max() / np.abs(x_ref).max():.2e}, "
            f"max rel err of var_t(x) {np.abs(v / v_ref - 1).max():.2e}")

    log(f"full solve, n = {Ci.shape[0]}, cond(Ci) = {np.linalg.cond(Ci):.2e}; reference = LAPACK dpotrf/dpotrs (upper)")
    report("LAPACK lower (rounding floor between two fp64 orders)", la.cholesky(Ci, lower=True))
    report("blocked left-looking, fp64 updates (the device's order)", blocked_cholesky(Ci, lambda A, B: A @ B.T))
    for S in slices:
        t0 = time.perf_counter()
        L = blocked_cholesky(Ci, lambda A, B: ozaki_gemm_nt(A, B, S))
        report(f"blocked, Ozaki S = {S} updates, exponent per row [{time.perf_counter() - t0:.0f} s]", L)
    for S in slices:
        t0 = time.perf_counter()
        L = blocked_cholesky(Ci, lambda A, B: ozaki_gemm_nt(A, B, S, 128))
        report(f"blocked, Ozaki S = {S}, exponent per row x 128 columns [{time.perf_counter() - t0:.0f} s]", L)


def write_panel(path, nslice=8):
    """Binary input of scratch/ozaki_probe.cu: header (int32 M, N, K, S), A [M][K], B [N][K] fp64 row-major -- the row
    panels of L that update tile (15, 14) of a cfg2 system (the probe computes a 128 x 64 tile: the first 64 rows of B)."""
    Ci, _ = cfg2_system()
    L = la.cholesky(Ci, lower=True)
    n = L.shape[0]
    i0, j0 = n - TILE, n - 2 * TILE
    A, B = np.ascontiguousarray(L[i0:i0 + TILE, :j0]), np.ascontiguousarray(L[j0:j0 + 64, :j0])
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        np.array([A.shape[0], B.shape[0], A.shape[1], nslice], dtype=np.int32).tofile(f)
        A.tofile(f)
        B.tofile(f)
    print(f"wrote {path}: M = {A.shape[0]}, N = {B.shape[0]}, K = {A.shape[1]}, S = {nslice}, {os.path.getsize(path)} bytes")


def read_panel(path):
    with open(path, "rb") as f:
        M, N, K, S = np.fromfile(f, dtype=np.int32, count=4)
        A = np.fromfile(f, dtype=np.float64, count=M * K).reshape(M, K)
        B = np.fromfile(f, dtype=np.float64, count=N * K).reshape(N, K)
    return A, B, int(S)


def expected_device_result(A, B, nslice=8, kblock=128):
    """(D [nkb][S][M][N] int32, OUT [M][N]) exactly as the probe's kernels form them: per K block the int32 accumulators,
    hi = ((D0 128 + D1) 128 + D2) 128 + D3 and lo from D4..D7 in int64, hi 2^-33 + lo 2^-61 in fp64, times 2^(e_i + f_j),
    K blocks added in ascending order."""
    assert nslice == 8
    nkb = A.shape[1] // kblock
    D = np.empty((nkb, nslice, A.shape[0], B.shape[0]), dtype=np.int64)
    out = np.zeros((A.shape[0], B.shape[0]))
    for kb in range(nkb):
        sl = slice(kb * kblock, (kb + 1) * kblock)
        ea, qa = split(A[:, sl], nslice)
        eb, qb = split(B[:, sl], nslice)
        D[kb] = ozaki_accumulators(qa, qb, nslice).astype(np.int64)
        hi = ((D[kb, 0] * 128 + D[kb, 1]) * 128 + D[kb, 2]) * 128 + D[kb, 3]
        lo = ((D[kb, 4] * 128 + D[kb, 5]) * 128 + D[kb, 6]) * 128 + D[kb, 7]
        comb = hi.astype(np.float64) * 2.0**-33 + lo.astype(np.float64) * 2.0**-61
        out += comb * np.exp2(ea[:, None] + eb[None, :])
    return D.astype(np.int32), out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--slices", type=int, nargs="+", default=[7, 8, 9])
    ap.add_argument("--full", action="store_true", help="also run the whole blocked factorisation + solve (minutes)")
    ap.add_argument("--write-panel", default=None)
    ap.add_argument("--log", default=None)
    args = ap.parse_args()
    if args.write_panel:
        return write_panel(args.write_panel)
    lines = []

    def log(s):
        print(s, flush=True)
        lines.append(s)

    Ci, rhs = cfg2_system()
    L = la.cholesky(Ci, lower=True)
    panel_study(L, Ci, args.slices, log)
    if args.full:
        solve_study(Ci, rhs, args.slices, log)
    if args.log:
        with open(args.log, "w") as f:
            f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
