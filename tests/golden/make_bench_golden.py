#!/usr/bin/env python3
"""Golden vectors AT THE HEADLINE SHAPE, produced by the UNMODIFIED reference (week-2/diet/simplex.py, imported
through oracle/ref_harness.py; run in the build container, which has /root/reference):

    python tests/golden/make_bench_golden.py        -> tests/golden/bench_shape_golden.json

The cases are the FIRST problems of the very batches bench.py solves (BASELINE.json configs[3]: random feasible
100 x 100 LPs, linear_programming_b200/workloads.py simplex_batch(dist, count, 100, 100, seed=2)): 96 of distribution
D1 (the timed batch), 6 of D2 (mixed signs: 1 300-1 700 pivots each, 25-40 s of reference time per LP) and 6 of D3
(negative b: phase 1 runs).  The inputs are NOT stored -- the seeded generator reproduces them, a CRC of each pins
that -- only what parity is defined on: status, the full (entering, leaving) sequence, x and the optimum as hex floats.
"""
import json
import multiprocessing
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_harness as rh  # noqa: E402
from linear_programming_b200 import workloads  # noqa: E402

N = M = 100
COUNTS = (("D1", 96), ("D2", 6), ("D3", 6))


def input_crc(A, b, c):
    return zlib.crc32(np.ascontiguousarray(c).tobytes(),
                      zlib.crc32(np.ascontiguousarray(b).tobytes(), zlib.crc32(np.ascontiguousarray(A).tobytes())))


def one(job):
    dist, idx = job
    A, b, c = workloads.simplex_batch(dist, idx + 1, N, M, seed=2)
    A, b, c = A[idx], b[idx], c[idx]
    r = rh.run_reference(A, b, c, timeout_s=900.0)
    assert r["status"] is not None, "reference did not terminate on %s/%d" % (dist, idx)
    return {"dist": dist, "index": idx, "input_crc": input_crc(A, b, c), "status": r["status"],
            "npivots": len(r["pivots"]), "pivots": [list(p) for p in r["pivots"]],
            "x": [float(v).hex() for v in r["x"]] if r["x"] is not None else None,
            "optimum": float(r["optimum"]).hex() if r["optimum"] is not None else None}


def one_single(shape):
    """configs[4]'s generator (workloads.single_lp, seed 5) at sizes the reference can hold: wide tableaus."""
    n, m = shape
    A, b, c = workloads.single_lp(n, m, seed=5)
    r = rh.run_reference(A, b, c, timeout_s=900.0)
    assert r["status"] is not None
    return {"dist": "single_lp", "index": 0, "n": n, "m": m, "input_crc": input_crc(A, b, c), "status": r["status"],
            "npivots": len(r["pivots"]), "pivots": [list(p) for p in r["pivots"]],
            "x": [float(v).hex() for v in r["x"]] if r["x"] is not None else None,
            "optimum": float(r["optimum"]).hex() if r["optimum"] is not None else None}


def main():
    jobs = [(d, i) for d, cnt in COUNTS for i in range(cnt)]
    jobs.sort(key=lambda j: j[0] == "D1")  # the slow ones first
    with multiprocessing.Pool(os.cpu_count()) as pool:
        cases = pool.map(one, jobs, chunksize=1)
    cases.sort(key=lambda c: (c["dist"], c["index"]))
    with multiprocessing.Pool(2) as pool:
        singles = pool.map(one_single, [(128, 256), (256, 512)])
    out = {"meta": {"generator": "tests/golden/make_bench_golden.py", "reference": "week-2/diet/simplex.py (unmodified)",
                    "python": "%d.%d.%d" % sys.version_info[:3], "n": N, "m": M, "seed": 2,
                    "inputs": "linear_programming_b200.workloads.simplex_batch(dist, count, 100, 100, seed=2)[index]"},
           "cases": cases, "single_lp_cases": singles}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "bench_shape_golden.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote", path, len(cases), "cases,", sum(c["npivots"] for c in cases), "pivots")


if __name__ == "__main__":
    main()
