#!/usr/bin/env python3
"""Per-kernel SASS census of liblp_b200.so (runs here, no GPU): the instruction mnemonics that back the design
claims -- TMA bulk copies + mbarriers in the grid-wide sweep (UBLKCP / SYNCS), warp-level integer reductions in
the argmins (REDUX), no FMA on the data path (DFMA only inside the IEEE division / reciprocal sequences: a handful
per MUFU.RCP64H), 128-bit shared/global accesses, no tensor-core instruction anywhere.

    python tools/sass_summary.py > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

WATCH = ["UBLKCP", "SYNCS", "CREDUX", "DFMA", "MUFU.RCP64H", "DMUL", "DADD", "DSETP", "LDS.128", "STS.128", "LDS.64",
         "STS.64", "LDG.E.128", "STG.E.128", "LD.E.128", "ST.E.128", "BAR.SYNC", "BAR.ARV", "SHFL", "ATOM", "HMMA",
         "UTCMMA", "UTMALDG", "LDL", "STL"]


def main():
    import bench
    lib = os.path.join(ROOT, "linear_programming_b200", "liblp_b200.so")
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
    regs = {}
    cur = None
    for ln in res.splitlines():
        m = re.search(r"Function ([^:]+):", ln)
        if m:
            cur = m.group(1)
        m = re.search(r"REG:(\d+)", ln)
        if m and cur:
            regs[cur] = (int(m.group(1)), (re.search(r"SHARED:(\d+)", ln) or [0, 0])[1], (re.search(r"LOCAL:(\d+)", ln) or [0, 0])[1])
    counts = collections.OrderedDict()
    name = None
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            name = m.group(1)
            counts[name] = collections.Counter()
            continue
        if name is None:
            continue
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m:
            op = m.group(1)
            counts[name]["_total"] += 1
            for w in WATCH:
                if op == w or op.startswith(w + ".") or (w.endswith("128") and op.startswith(w)):
                    counts[name][w] += 1
    print("# SASS census of linear_programming_b200/liblp_b200.so  (sm_100a; sources sha16 %s)" % bench.source_sha16())
    print("# produced by tools/sass_summary.py; one block per kernel: registers / static shared / local bytes, then counts")
    demangle = subprocess.run(["c++filt"] + list(counts), capture_output=True, text=True).stdout.splitlines()
    for (mangled, cnt), nice in zip(counts.items(), demangle):
        short = re.sub(r"\(.*", "", nice)
        r = regs.get(mangled, ("?", "?", "?"))
        print("\n%s\n  registers %s, static smem %s B, local %s B, %d instructions" % (short, r[0], r[1], r[2], cnt["_total"]))
        line = ["%s %d" % (w, cnt[w]) for w in WATCH if cnt[w]]
        print("  " + ", ".join(line))
        if cnt["MUFU.RCP64H"]:
            print("  DFMA per MUFU.RCP64H: %.1f (the IEEE division / reciprocal expansions; the tableau update itself is DMUL + DADD)"
                  % (cnt["DFMA"] / cnt["MUFU.RCP64H"]))
        elif cnt["DFMA"]:
            print("  DFMA without a reciprocal: CHECK")
    tensor = sum(c["HMMA"] + c["UTCMMA"] for c in counts.values())
    print("\n# tensor-core instructions in the whole library: %d (rank-1 AXPY at 0.125 flop/B: deliberately unused)" % tensor)


if __name__ == "__main__":
    main()
