#!/usr/bin/env python3
"""Distil one `ncu --set full` capture (read here, no GPU needed) into the per-launch facts bench.py reports,
keyed by kernel, and tie them to the build they were taken from (source hash):

    python tools/ncu_facts.py <kernel-key> <report.ncu-rep> --units N [--sha16 H] [--summary profiles/x.csv]

  kernel-key   simplex_batched_kernel | simplex_single_kernel | gauss_jordan_batched_kernel | verify_bounds_kernel
  --units      what ONE captured launch processed: pivots (simplex kernels) or systems / LPs (the others)
  --sha16      bench.source_sha16() of the tree the capture ran on (default: gpurun_out/<report>.sha, else now)

Writes/updates profiles/ncu_facts.json.  bench.py reports `traffic` & co. only while the hash still matches."""
import argparse
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
        "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "": 1.0, "%": 1.0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kernel")
    ap.add_argument("report")
    ap.add_argument("--units", type=float, required=True)
    ap.add_argument("--sha16", default=None)
    ap.add_argument("--summary", default=None)
    ap.add_argument("--launch", type=int, default=0, help="which captured launch of the report")
    args = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", args.report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2 + args.launch]
    rec = {h: (u, v) for h, u, v in zip(hdr, units, vals)}

    def num(name, scale_units=True):
        if name not in rec:
            return None
        u, v = rec[name]
        try:
            x = float(v.replace(",", ""))
        except ValueError:
            return None
        return x * UNIT.get(u, 1.0) if scale_units else x

    sha = args.sha16
    if sha is None:  # the kernel's own files (bench.KERNEL_SOURCES), as they are NOW: run this right after the capture
        import bench
        sha = bench.source_sha16(args.kernel)
    dram = (num("dram__bytes_read.sum") or 0.0) + (num("dram__bytes_write.sum") or 0.0)
    wave = num("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", False)
    stalls = {}
    tot = 0.0
    for h, (u, v) in rec.items():
        if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued"):
            try:
                stalls[h[len("smsp__pcsamp_warps_issue_stalled_"):]] = float(v.replace(",", ""))
                tot += float(v.replace(",", ""))
            except ValueError:
                pass
    top = sorted(stalls.items(), key=lambda kv: -kv[1])[:5]
    out = {
        "file": args.summary or os.path.relpath(args.report, ROOT), "src_sha16": sha,
        "kernel_name": rec.get("Kernel Name", ("", "?"))[1], "units_per_launch": args.units,
        "duration_ms": num("gpu__time_duration.sum"), "dram_bytes": dram,
        "registers": num("launch__registers_per_thread", False),
        "smem_wavefronts": wave,
        "smem_pipe_pct": num("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", False),
        "fp64_pipe_pct": num("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", False),
        "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active", False),
        "dram_pct_of_peak": num("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", False),
        "warp_instructions": num("smsp__inst_executed.sum", False),
        "stall_top5_pct": {k: round(100.0 * v / tot, 1) for k, v in top} if tot else None,
    }
    if args.kernel.startswith("simplex_batched") or args.kernel.startswith("simplex_single"):
        out["pivots"] = args.units
        out["dram_bytes_per_pivot"] = dram / args.units if args.units else None
        if wave:
            out["smem_wavefronts_per_pivot"] = wave / args.units
    else:
        out["dram_bytes_per_system"] = dram / args.units if args.units else None
    path = os.path.join(ROOT, "profiles", "ncu_facts.json")
    allf = {}
    if os.path.exists(path):
        with open(path) as f:
            allf = json.load(f)
    allf[args.kernel] = out
    with open(path, "w") as f:
        json.dump(allf, f, indent=1, sort_keys=True)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
