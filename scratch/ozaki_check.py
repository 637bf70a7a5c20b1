"""Check the outputs of scratch/ozaki_probe.cu against the exact host emulation (scratch/ozaki_study.py).

    python scratch/ozaki_check.py <panel.bin> <out prefix> [<out prefix> ...]

Per prefix: the int32 accumulators of every K block bit for bit, the fp64 result bit for bit against the emulation of the
tcgen05 engine's hi / lo combination (the mma.sync engine folds the diagonals one by one: rounding-level agreement), and
both against A B^T in 80-bit arithmetic next to the device's own DMMA result.
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ozaki_study as oz  # noqa: E402


def main():
    panel, prefixes = sys.argv[1], sys.argv[2:]
    A, B, S = oz.read_panel(panel)
    M, N, K = A.shape[0], B.shape[0], A.shape[1]
    D_exp, out_exp = oz.expected_device_result(A, B, S)
    exact = oz.exact_nt(A, B)
    scale = float(np.abs(exact).max())
    err = lambda X: float(np.abs(X.astype(np.longdouble) - exact).max())  # noqa: E731
    for pre in prefixes:
        rec = {"prefix": os.path.basename(pre), "M": M, "N": N, "K": K, "slices": S}
        if not os.path.exists(pre + "_out.bin"):
            rec["status"] = "no output (the run failed)"
            print(json.dumps(rec))
            continue
        D = np.fromfile(pre + "_D.bin", dtype=np.int32).reshape(D_exp.shape)
        out = np.fromfile(pre + "_out.bin", dtype=np.float64).reshape(M, N)
        dmma = np.fromfile(pre + "_dmma.bin", dtype=np.float64).reshape(M, N)
        bad = D != D_exp
        rec["accumulators_bit_exact"] = bool(not bad.any())
        rec["accumulator_mismatches"] = int(bad.sum())
        if bad.any():
            kb, d, i, j = [int(v[0]) for v in np.nonzero(bad)]
            rec["first_mismatch"] = {"kb": kb, "d": d, "row": i, "col": j, "got": int(D[kb, d, i, j]), "want": int(D_exp[kb, d, i, j])}
            # which rows / columns / diagonals are wrong tells a layout error from an accumulate-flag error
            rec["bad_by_diagonal"] = [int(bad[:, d].sum()) for d in range(S)]
            rec["bad_rows_first_block"] = int(bad[0].any(axis=(0, 2)).sum())
            rec["bad_cols_first_block"] = int(bad[0].any(axis=(0, 1)).sum())
        rec["result_bit_exact_vs_emulation"] = bool((out == out_exp).all())
        rec["result_max_abs_diff_vs_emulation"] = float(np.abs(out - out_exp).max())
        rec["max_abs_exact"] = scale
        rec["err_ozaki_device_vs_exact"] = err(out)
        rec["err_ozaki_emulation_vs_exact"] = err(out_exp)
        rec["err_device_dmma_vs_exact"] = err(dmma)
        rec["err_numpy_dgemm_vs_exact"] = err(A @ B.T)
        rec["rel"] = {k: rec[k] / scale for k in ("err_ozaki_device_vs_exact", "err_device_dmma_vs_exact", "err_numpy_dgemm_vs_exact")}
        print(json.dumps(rec))


if __name__ == "__main__":
    main()
