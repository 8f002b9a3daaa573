#!/usr/bin/env python3
"""Time the UNMODIFIED Python reference, in-process, on samples of the bench workloads
(BASELINE.md section 4, items 1 and 4).  Build container only (needs /root/reference); the result is
committed as tests/golden/reference_batch_times.json and quoted in DESIGN.md / BASELINE.md beside the
numbers measured on the GPU box, where the reference does not exist.

  * Config 4: the first --d1 LPs of D1 and the first --d2 LPs of D2 (the very inputs bench.py uses),
    one after the other on one core: seconds and pivots per LP, parity with the C oracle re-checked;
  * Config 5: D1-style single LPs at 128x256 and 256x512: ms per pivot (the reference allocates an
    nvars x nvars list of lists, simplex.py:252, so 16384x32768 is out of reach: extrapolate ~ n*(n+m)).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

import oracle
from oracle import ref_harness
from linear_programming_b200 import workloads


def time_lp(A, b, c, timeout):
    t0 = time.perf_counter()
    r = ref_harness.run_reference(A.tolist(), b.tolist(), c.tolist(), timeout_s=timeout)
    dt = time.perf_counter() - t0
    o = oracle.simplex(A, b, c)
    same = (r["status"] == o["status"] and r["pivots"] == o["pivots"])
    return dt, len(r["pivots"]), r["status"], same


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--d1", type=int, default=16)
    ap.add_argument("--d2", type=int, default=2)
    args = ap.parse_args()
    out = {"python": sys.version.split()[0], "cores_used": 1, "host_cores": os.cpu_count(),
           "note": "unmodified week-2/diet/simplex.py imported in-process, one LP after the other on one core; "
                   "measured in the build container"}
    for dist, cnt in (("D1", args.d1), ("D2", args.d2)):
        A, b, c = workloads.simplex_batch(dist, cnt, 100, 100, seed=2)
        secs, piv, ok = [], [], True
        for p in range(cnt):
            dt, k, st, same = time_lp(A[p], b[p], c[p], 600.0)
            secs.append(dt); piv.append(k); ok = ok and same
            print(dist, p, "%.2f s" % dt, k, st, same, flush=True)
        out["config4_" + dist] = {"lps": cnt, "seconds_total": sum(secs), "pivots_total": sum(piv),
                                  "lps_per_s_per_core": cnt / sum(secs), "pivots_per_s_per_core": sum(piv) / sum(secs),
                                  "ms_per_pivot": 1e3 * sum(secs) / max(1, sum(piv)), "oracle_parity": ok}
    for n, m in ((128, 256), (256, 512)):
        A, b, c = workloads.single_lp(n, m, seed=5)
        dt, k, st, same = time_lp(A, b, c, 1800.0)
        print("single", n, m, "%.2f s" % dt, k, st, same, flush=True)
        out["config5_%dx%d" % (n, m)] = {"seconds": dt, "pivots": k, "ms_per_pivot": 1e3 * dt / max(1, k),
                                         "status": st, "oracle_parity": same}
    dst = os.path.join(ROOT, "tests", "golden", "reference_batch_times.json")
    json.dump(out, open(dst, "w"), indent=1, sort_keys=True)
    print("wrote", dst)


if __name__ == "__main__":
    main()
