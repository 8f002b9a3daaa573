#!/usr/bin/env python3
"""Golden vectors for inputs OUTSIDE the bit-exact domain: LPs whose constraint matrix holds non-zero entries of
magnitude <= tolerance (1e-9).  Produced by the UNMODIFIED reference (oracle/ref_harness.py imports
/root/reference/week-2/diet/simplex.py); run in the build container, the output is committed:

    python tests/golden/make_subtol_golden.py > tests/golden/subtol_golden.json

Why they are separate: the engine's condensed tableau treats "coefficient is None"
and "coefficient is 0.0" alike.  That is exact for integer input and for float input without sub-tolerance entries.
With such an entry the reference's substitute() (simplex.py:203-205) returns the row unchanged, so the stale
coefficient stays in a now-basic column and comes back when that variable leaves the basis; the engine zeroes it.
Status and pivot sequence still agree; x may differ in the last digits -- inside north_star's 1e-9 relative bar."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_harness  # noqa: E402


def main():
    rng = np.random.default_rng([2026, 11])
    cases = []
    tries = 0
    while len(cases) < 120 and tries < 4000:
        tries += 1
        n, m = (int(v) for v in rng.integers(2, 9, size=2))
        A = rng.integers(-100, 101, size=(n, m)).astype(np.float64)
        tiny = rng.random((n, m)) < 0.25
        A[tiny] = rng.choice([-1.0, 1.0], size=int(tiny.sum())) * rng.uniform(1e-11, 1e-9, size=int(tiny.sum()))
        b = rng.integers(1, 1000, size=n).astype(np.float64)
        if rng.random() < 0.3:
            b[rng.integers(0, n)] = -float(rng.integers(1, 500))
        c = rng.integers(-100, 101, size=m).astype(np.float64)
        r = ref_harness.run_reference(A, b, c, timeout_s=20.0)
        if r["status"] is None or r["status"] == 99:
            continue
        cases.append({"n": n, "m": m, "A": [float(v).hex() for v in A.reshape(-1)], "b": b.tolist(), "c": c.tolist(),
                      "status": r["status"], "pivots": [list(p) for p in r["pivots"]],
                      "x": None if r["x"] is None else [float(v).hex() for v in r["x"]],
                      "optimum": None if r["optimum"] is None else float(r["optimum"]).hex()})
    json.dump({"meta": {"generator": "tests/golden/make_subtol_golden.py", "python": cases and r["python"],
                        "reference": "week-2/diet/simplex.py (unmodified)", "tolerance": 1e-9}, "cases": cases},
              sys.stdout)


if __name__ == "__main__":
    main()
