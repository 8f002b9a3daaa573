#!/usr/bin/env python3
"""One-off fuzz of the CPU oracle (oracle/simplex_oracle.c) against the UNMODIFIED reference imported in this
container (oracle/ref_harness.py): fresh seeded random LPs, every outcome class, compared on status, the full pivot
sequence and the bits of x and of the optimum.  Writes a summary to tests/golden/oracle_vs_reference_fuzz.json.

    python tools/fuzz_oracle_vs_reference.py [--small 4000] [--mid 48] [--seed 2026]
"""
import argparse
import json
import multiprocessing
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from oracle import ref_harness as rh  # noqa: E402


def make(job):
    kind, seed = job
    rng = np.random.default_rng([seed, 404])
    if kind == "small":   # the stress.py distribution and a bit beyond, 8 % zero b
        n, m = (int(v) for v in rng.integers(1, 21, size=2))
        A = rng.integers(-100, 101, size=(n, m)).astype(np.float64)
        b = rng.integers(-1000000, 1000001, size=n).astype(np.float64)
        b[rng.random(n) < 0.08] = 0.0
        c = rng.integers(-100, 101, size=m).astype(np.float64)
    elif kind == "float":  # non-integer data (no sub-tolerance entries: |A| >= 1e-3)
        n, m = (int(v) for v in rng.integers(2, 15, size=2))
        A = np.round(rng.uniform(-50, 50, size=(n, m)), 3)
        A[np.abs(A) < 1e-3] = 1.0
        b = np.round(rng.uniform(-1000, 5000, size=n), 2)
        c = np.round(rng.uniform(-50, 50, size=m), 3)
    else:                  # mid size, all three bench distributions
        n, m = (int(v) for v in rng.integers(40, 91, size=2))
        lo = 0 if seed % 3 == 0 else -100
        A = rng.integers(lo, 101, size=(n, m)).astype(np.float64)
        b = rng.integers(1, 1000001, size=n).astype(np.float64)
        if seed % 3 == 2:
            b[rng.random(n) < 0.09] *= -0.1
            b = np.round(b)
            b[b == 0] = 1.0
        c = rng.integers(1 if lo == 0 else -100, 101, size=m).astype(np.float64)
    return A, b, c


def one(job):
    A, b, c = make(job)
    r = rh.run_reference(A, b, c, timeout_s=600.0)
    if r["status"] is None or r["status"] == 99:
        return job, "skipped", None, 0
    o = oracle.simplex(A, b, c, max_pivots=200000)
    ok = (o["status"] == r["status"] and [list(p) for p in o["pivots"]] == [list(p) for p in r["pivots"]])
    if ok and r["status"] == 0:
        ok = ([float(v).hex() for v in o["x"]] == [float(v).hex() for v in r["x"]]
              and float(o["objective"]).hex() == float(r["optimum"]).hex())
    return job, ("match" if ok else "MISMATCH"), r["status"], len(r["pivots"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--small", type=int, default=4000)
    ap.add_argument("--floats", type=int, default=1500)
    ap.add_argument("--mid", type=int, default=48)
    ap.add_argument("--seed", type=int, default=2026)
    args = ap.parse_args()
    oracle.build()
    jobs = [("mid", args.seed + i) for i in range(args.mid)] + [("small", args.seed + i) for i in range(args.small)] + \
           [("float", args.seed + i) for i in range(args.floats)]
    t0 = time.time()
    with multiprocessing.Pool(os.cpu_count()) as pool:
        res = pool.map(one, jobs, chunksize=8)
    out = {"generator": "tools/fuzz_oracle_vs_reference.py", "seed": args.seed, "seconds": round(time.time() - t0, 1),
           "python": "%d.%d.%d" % sys.version_info[:3], "by_kind": {}}
    for kind in ("small", "float", "mid"):
        rows = [r for r in res if r[0][0] == kind]
        st = {}
        for r in rows:
            if r[1] != "skipped":
                st[str(r[2])] = st.get(str(r[2]), 0) + 1
        out["by_kind"][kind] = {"lps": len(rows), "match": sum(r[1] == "match" for r in rows),
                                "mismatch": [list(r[0]) for r in rows if r[1] == "MISMATCH"],
                                "skipped_reference_timeout": sum(r[1] == "skipped" for r in rows),
                                "pivots": int(sum(r[3] for r in rows)), "by_reference_status": st}
    path = os.path.join(ROOT, "tests", "golden", "oracle_vs_reference_fuzz.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
