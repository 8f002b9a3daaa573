#!/usr/bin/env python3
"""Generate the committed golden vectors from the UNMODIFIED reference.

Run in the build container (needs /root/reference):

    python tests/golden/make_golden.py

Writes
  tests/golden/simplex_golden.json       status / pivot sequence / x bits / optimum bits of
                                         week-2/diet/simplex.py on every fixture of the
                                         reference, its three unit-test LPs and seeded random LPs
  tests/golden/gauss_jordan_golden.json  outputs of the compiled reference energy_values.cc
                                         (oracle/_ref, built by oracle/Makefile) on its fixtures
                                         and on seeded random systems

Floats are stored as C99 hex strings (float.hex()) so parity tests compare bit patterns.
Interpreter that produced the set-iteration order is recorded ("python").
"""
import glob
import json
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from oracle import ref_harness as rh  # noqa: E402
from linear_programming_b200 import workloads  # noqa: E402

REF = rh.REFERENCE_ROOT
HERE = os.path.dirname(os.path.abspath(__file__))


def hx(v):
    return float(v).hex()


def ints(rows):
    return [[int(v) for v in r] for r in rows]


def record(name, A, b, c, source, extra=None, timeout_s=300.0):
    r = rh.run_reference(A, b, c, timeout_s=timeout_s)
    assert r["status"] is not None, "reference did not terminate on " + name
    case = {
        "name": name, "source": source, "n": len(b), "m": len(c),
        "A": ints(A), "b": [int(v) for v in b], "c": [int(v) for v in c],
        "status": r["status"], "npivots": len(r["pivots"]),
        "pivots": [list(p) for p in r["pivots"]],
        "x": [hx(v) for v in r["x"]] if r["x"] is not None else None,
        "optimum": hx(r["optimum"]) if r["optimum"] is not None else None,
    }
    if extra:
        case.update(extra)
    return case


def simplex_cases():
    cases = []
    # 1. every stdin fixture the reference holds for this path
    files = sorted(glob.glob(REF + "/week-2/diet/tests/**/*", recursive=True)
                   + glob.glob(REF + "/week-2/diet/fail/*")
                   + glob.glob(REF + "/week-2/ad_allocation/tests/*"))
    for f in files:
        if f.endswith(".a") or os.path.isdir(f):
            continue
        rel = os.path.relpath(f, REF)
        text = open(f).read()
        try:
            A, b, c = rh.parse_reference(text)
        except ValueError:
            print("skip (malformed for the reference's own parser):", rel)
            continue
        extra = {"stdin": text}
        if os.path.exists(f + ".a"):
            extra["expected_answer_file"] = open(f + ".a").read()
        cases.append(record(rel, A, b, c, "fixture", extra))
    # 2. the three known-answer LPs of week-2/diet/test_simplex.py:16-50
    unit = [
        ("test_simplex.py::test_01", [[-1, -1], [1, 0], [0, 1]], [-1, 2, 2], [-1, 2], 4.0),
        ("test_simplex.py::test_feasible_start", [[1, 1, 3], [2, 2, 5], [4, 1, 2]], [30, 24, 36], [3, 1, 2], 28.0),
        ("test_simplex.py::test_infeasible_start", [[2, -1], [1, -5]], [2, -4], [2, -1], 2.0),
    ]
    for name, A, b, c, opt in unit:
        case = record(name, A, b, c, "unit", {"expected_optimum": hx(opt)})
        assert float.fromhex(case["optimum"]) == opt
        cases.append(case)
    # 3. seeded random LPs: stress.py distribution (+zeros in b), sparse (all crash classes),
    #    mid-size, tall (set table wraps), and the first problems of the bench distributions
    for i, (A, b, c) in enumerate(workloads.stress_batch(160, seed=0)):
        cases.append(record("stress/%03d" % i, A, b, c, "stress.py:23-26 distribution, seed 0"))
    rng = random.Random(11)
    for i in range(160):
        n, m = rng.randint(1, 7), rng.randint(1, 7)
        A = [[rng.choice([0, 0, 0, 1, -1, 2]) for _ in range(m)] for _ in range(n)]
        b = [rng.choice([1, 2, 3, 5, -1, 0, -3]) for _ in range(n)]
        c = [rng.choice([1, 2, -1, 0, 3]) for _ in range(m)]
        cases.append(record("sparse/%03d" % i, A, b, c, "sparse small ints, random.Random(11)"))
    rng = random.Random(12)
    for i in range(40):
        n, m = rng.randint(5, 40), rng.randint(5, 40)
        z = rng.random() < 0.5
        A = [[(0 if (z and rng.random() < 0.3) else rng.randint(-100, 100)) for _ in range(m)] for _ in range(n)]
        b = [rng.randint(-1000, 10**6) for _ in range(n)]
        c = [rng.randint(-100, 100) for _ in range(m)]
        cases.append(record("mid/%03d" % i, A, b, c, "mid-size, random.Random(12)"))
    rng = random.Random(13)
    for i in range(30):
        n, m = rng.randint(20, 60), rng.randint(2, 6)
        A = [[rng.randint(-20, 100) for _ in range(m)] for _ in range(n)]
        b = [rng.randint(-50, 10**4) for _ in range(n)]
        c = [rng.randint(-10, 100) for _ in range(m)]
        cases.append(record("tall/%03d" % i, A, b, c, "tall (set table wraps), random.Random(13)"))
    for dist, cnt in (("D1", 6), ("D2", 2), ("D3", 2)):
        A, b, c = workloads.simplex_batch(dist, cnt, 100, 100, seed=2)
        for i in range(cnt):
            cases.append(record("config4/%s/%d" % (dist, i), A[i], b[i], c[i],
                                "workloads.simplex_batch(%r, n=100, m=100, seed=2)[%d]" % (dist, i),
                                timeout_s=900.0))
            print("config4", dist, i, cases[-1]["status"], cases[-1]["npivots"], flush=True)
    return cases


def gj_cases():
    assert oracle.ref_gauss_jordan_available(), "run `make -C oracle ref` first"
    cases = []
    for f in sorted(glob.glob(REF + "/week-2/energy_values/tests/*")):
        if f.endswith(".a"):
            continue
        toks = open(f).read().split()
        n = int(toks[0])
        vals = [float(t) for t in toks[1:]]
        a = [vals[i * (n + 1):i * (n + 1) + n] for i in range(n)]
        b = [vals[i * (n + 1) + n] for i in range(n)]
        x = oracle.ref_gauss_jordan(np.array(a).reshape(n, n), np.array(b)) if n else []
        case = {"name": os.path.relpath(f, REF), "n": n, "a": [[hx(v) for v in r] for r in a],
                "b": [hx(v) for v in b], "x": [hx(v) for v in x], "stdin": open(f).read()}
        if os.path.exists(f + ".a"):
            case["expected_answer_file"] = open(f + ".a").read()
        cases.append(case)
    # seeded random systems from the bench generator (incl. planted zero pivots -> row swap)
    for n, cnt, frac in ((1, 2, 0.0), (2, 4, 0.5), (3, 4, 0.5), (7, 6, 0.5), (16, 6, 0.5), (33, 4, 0.5), (64, 3, 0.34)):
        a, b = workloads.gauss_jordan_batch(cnt, n=n, seed=1, swap_frac=frac)
        for i in range(cnt):
            x = oracle.ref_gauss_jordan(a[i], b[i])
            cases.append({"name": "config3/n%d/%d" % (n, i), "n": n,
                          "generator": "workloads.gauss_jordan_batch(%d, n=%d, seed=1, swap_frac=%r)[%d]" % (cnt, n, frac, i),
                          "x": [hx(v) for v in x]})
    # a singular system: the reference divides by zero and prints inf/nan -- pin that too
    a = np.array([[0.0, 1.0], [0.0, 2.0]])
    b = np.array([1.0, 2.0])
    x = oracle.ref_gauss_jordan(a, b)
    cases.append({"name": "singular/2x2", "n": 2, "a": [[hx(v) for v in r] for r in a],
                  "b": [hx(v) for v in b], "x": [hx(v) for v in x]})
    return cases


def main():
    oracle.build()
    sc = simplex_cases()
    meta = {"python": "%d.%d.%d" % sys.version_info[:3], "reference": "week-2/diet/simplex.py",
            "tolerance": 1e-9, "status_codes": {"0": "bounded", "-1": "no solution", "1": "infinity",
                                                "3": "AttributeError", "4": "IndexError", "5": "TypeError"}}
    with open(os.path.join(HERE, "simplex_golden.json"), "w") as f:
        json.dump({"meta": meta, "cases": sc}, f, separators=(",", ":"))
    hist = {}
    for c in sc:
        hist[c["status"]] = hist.get(c["status"], 0) + 1
    print("simplex cases:", len(sc), hist)
    gc = gj_cases()
    with open(os.path.join(HERE, "gauss_jordan_golden.json"), "w") as f:
        json.dump({"meta": {"reference": "week-2/energy_values/energy_values.cc",
                            "flags": "-pipe -O2 -std=c++0x"}, "cases": gc}, f, separators=(",", ":"))
    print("gauss-jordan cases:", len(gc))


if __name__ == "__main__":
    main()
