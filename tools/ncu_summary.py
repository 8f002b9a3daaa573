#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here, no GPU needed) into a small CSV for profiles/.

    python tools/ncu_summary.py gpurun_out/prof_x.ncu-rep profiles/r1_x_ncu_summary.csv
"""
import csv
import io
import subprocess
import sys

KEEP = (
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__cycles_elapsed.max", "smsp__pcsamp_sample_count",
)


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["launch", "kernel", "metric", "unit", "value"])
        for li, vals in enumerate(rows[2:]):
            rec = dict(zip(hdr, vals))
            name = rec.get("Kernel Name", "?")
            for h, u, v in zip(hdr, units, vals):
                if h in KEEP or (h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")):
                    w.writerow([li, name, h, u, v])
    print("wrote", out)


if __name__ == "__main__":
    main()
