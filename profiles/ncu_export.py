"""Reduce an .ncu-rep (read here, no GPU needed) to the `metric,unit,value` CSVs kept under profiles/:
    python profiles/ncu_export.py gpurun_out/r02p_dense.ncu-rep profiles/r02p_dense_sampler_ncu_raw.csv "header comment"
One block per captured launch; only the metric groups that matter (time, DRAM, L2, pipes, issue/stall, shared memory,
launch geometry)."""
import csv
import subprocess
import sys

KEEP = ("gpu__time_duration", "dram__bytes", "dram__cycles_active", "dram__throughput", "lts__t_sector_hit_rate", "lts__t_bytes.sum",
        "sm__pipe_tensor", "sm__inst_executed_pipe_tensor", "sm__inst_executed_pipe_lsu", "sm__inst_executed_pipe_fp64", "sm__inst_executed.sum",
        "sm__ops_path_tensor", "smsp__issue_active", "smsp__average_warps_issue_stalled", "smsp__warps_active", "sm__warps_active",
        "sm__cycles_elapsed.max", "sm__cycles_active.avg", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared", "l1tex__data_pipe_lsu_wavefronts_mem_shared",
        "launch__", "sm__throughput", "gpu__compute_memory_throughput", "smsp__inst_executed.sum")


def main():
    rep, out, note = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        f.write(f"# {note}\n")
        f.write("metric,unit,value\n")
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            f.write(f"Kernel Name,,\"{d.get('Kernel Name')}\"\n")
            for k, u in zip(hdr, units):
                if k.startswith(KEEP) and "pcsamp" not in k:
                    f.write(f"{k},{u},{d[k]}\n")


if __name__ == "__main__":
    main()
