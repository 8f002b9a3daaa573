#!/usr/bin/env python3
"""Time the UNMODIFIED reference on its own course fixtures (BASELINE.md measurement plan, item 2).

Runs in the build container only (it needs /root/reference): `./simplex.py < file` as the reference
is used (a fresh interpreter per problem, week-2/diet/simplex.py:720-751), wall time per file, best of
`--reps`.  The result is committed as tests/golden/reference_course_times.json so that
tools/bench_course.py can print it beside the GPU numbers on the box, where the reference does not exist.
"""
import argparse
import glob
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("LP_REFERENCE_ROOT", "/root/reference")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--timeout", type=float, default=600.0)
    args = ap.parse_args()
    simplex = os.path.join(REF, "week-2", "diet", "simplex.py")
    files = sorted(glob.glob(os.path.join(REF, "week-2", "ad_allocation", "tests", "*")) +
                   glob.glob(os.path.join(REF, "week-2", "diet", "tests", "*", "*")) +
                   glob.glob(os.path.join(REF, "week-2", "diet", "tests", "[0-9]*")))
    files = [f for f in files if os.path.isfile(f) and not f.endswith(".a")]
    out = {"python": sys.version.split()[0], "cores": os.cpu_count(), "cmd": "python3 simplex.py < file",
           "note": "wall time of one process per problem, best of %d; measured in the build container" % args.reps,
           "cases": {}}
    for f in files:
        name = os.path.relpath(f, REF)
        data = open(f, "rb").read()
        best = None
        rc = None
        for _ in range(args.reps):
            t0 = time.perf_counter()
            try:
                p = subprocess.run([sys.executable, simplex], input=data, capture_output=True, timeout=args.timeout)
                rc = p.returncode
            except subprocess.TimeoutExpired:
                rc = "timeout"
                break
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        n, m = (int(t) for t in data.split()[:2])
        out["cases"][name] = {"n": n, "m": m, "seconds": best, "returncode": rc}
        print(name, n, m, best, rc, flush=True)
    # interpreter start-up alone, for scale
    t0 = time.perf_counter()
    subprocess.run([sys.executable, "-c", "pass"])
    out["interpreter_startup_seconds"] = time.perf_counter() - t0
    dst = os.path.join(ROOT, "tests", "golden", "reference_course_times.json")
    json.dump(out, open(dst, "w"), indent=1, sort_keys=True)
    print("wrote", dst)


if __name__ == "__main__":
    main()
