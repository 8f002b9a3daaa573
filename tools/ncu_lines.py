#!/usr/bin/env python3
"""Per-source-line view of an ncu --set full --import-source capture (read here): instructions, stall samples and
shared-memory wavefronts per source line, sorted by samples.
    python tools/ncu_lines.py report.ncu-rep [--top 40] [--per N]      (--per: divide counts by N units)"""
import argparse
import csv
import io
import subprocess


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--per", type=float, default=1.0)
    ap.add_argument("--sort", default="samples", choices=["samples", "inst", "wav"])
    args = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", args.report, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    cur, hdr, agg = None, None, {}
    for r in csv.reader(io.StringIO(raw)):
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif r[0].isdigit() and len(r) > 2 and r[2] == "-" and hdr:
            d = dict(zip(hdr, r))
            try:
                agg[(cur, int(r[0]))] = (int(d["Instructions Executed"]), int(d["# Samples"]),
                                         int(d.get("L1 Wavefronts Shared", "0") or 0), r[1].strip()[:100])
            except ValueError:
                pass
    ti = sum(v[0] for v in agg.values())
    ts = sum(v[1] for v in agg.values()) or 1
    tw = sum(v[2] for v in agg.values())
    print("total: %.4g instructions, %d samples, %.4g shared wavefronts (per unit: %.1f inst, %.1f wavefronts)"
          % (ti, ts, tw, ti / args.per, tw / args.per))
    key = {"samples": 1, "inst": 0, "wav": 2}[args.sort]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][key])[:args.top]:
        print("%-26s %5d  inst %10.1f  samples %5.1f%%  wavefronts %9.1f | %s"
              % (k[0][:26], k[1], v[0] / args.per, 100.0 * v[1] / ts, v[2] / args.per, v[3]))


if __name__ == "__main__":
    main()
