#!/usr/bin/env python3
"""GPU-batched differential fuzzer -- the reference's stress loop (week-2/diet/stress.py:21-87)
at 10^5..10^6 LPs per run.  Test infrastructure: lives under tests/
because it uses the CPU checker.

    python tests/fuzz_stress.py [--count 100000] [--seed 0] [--nmax 7] [--scipy 200]

Per run: LPs from the stress.py distribution (n, m ~ U{1..nmax}, A ~ U{-100..99},
b ~ U{-10^6..10^6-1} with 7 % zeros, c ~ U{-100..99}) are grouped by shape, each group is ONE
batched engine call, and every LP is compared with the checker: status, pivot count, pivot
sequence, x and objective bit for bit.  Then the reference's own post-conditions:
  * verify_bounds (simplex.py:663-671): A x <= b + 1e-3 on every "Bounded" answer, batched ON THE GPU
    (lpb200_verify_bounds_batch, itself compared with the checker's restatement on a sample);
    violations are COUNTED, not failed -- the reference's degenerate-skip ratio test returns
    infeasible x on some inputs and parity means reproducing that;
  * scipy cross-check (stress.py:48-83) on a sample: status agreement and |c.x - c.x_ref| <= 1e-3
    are REPORTED beside the reference's known failure classes.
Prints one JSON line; exit code 1 iff any engine-vs-checker mismatch.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def run(count=100000, seed=0, nmax=7, scipy_sample=200, max_pivots=2000):
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    import oracle
    t0 = time.perf_counter()
    probs = workloads.stress_batch(count, seed=seed, nmax=nmax)
    groups = {}
    for idx, (A, b, c) in enumerate(probs):
        groups.setdefault(A.shape, []).append(idx)
    t_gen = time.perf_counter() - t0
    status = np.zeros(count, dtype=np.int32)
    mism = []
    viol = 0
    bounded = 0
    t_gpu = t_cpu = 0.0
    xs = [None] * count
    for (n, m), idxs in sorted(groups.items()):
        A = np.stack([probs[i][0] for i in idxs])
        b = np.stack([probs[i][1] for i in idxs])
        c = np.stack([probs[i][2] for i in idxs])
        t1 = time.perf_counter()
        res = lp.solve_batch(A, b, c, max_pivots=max_pivots, log_cap=64)
        t_gpu += time.perf_counter() - t1
        t1 = time.perf_counter()
        st, x, obj, npiv, _ = oracle.simplex_batch(A, b, c, max_pivots=max_pivots)
        t_cpu += time.perf_counter() - t1
        bad = (res.status != st) | (res.npivots != npiv) | (res.objective.view(np.int64) != obj.view(np.int64)) | \
              (res.x.view(np.int64) != x.view(np.int64)).any(axis=1)
        for k in np.nonzero(bad)[0]:
            mism.append(int(idxs[k]))
        # pivot sequence of a sample of each group against the single-LP checker (it logs pivots)
        for k in range(0, len(idxs), max(1, len(idxs) // 8)):
            o = oracle.simplex(A[k], b[k], c[k], max_pivots=max_pivots)
            if res.pivots(k) != o["pivots"][:64]:
                mism.append(int(idxs[k]))
        status[idxs] = res.status
        ok = res.status == 0
        bounded += int(ok.sum())
        if ok.any():  # verify_bounds (simplex.py:663-671), batched on the GPU: first asserting row per LP
            t1 = time.perf_counter()
            first_bad = lp.simplex.verify_bounds_batch(A, b, res.x, res.status)
            t_gpu += time.perf_counter() - t1
            viol += int((first_bad >= 0).sum())
            # the kernel against the checker's restatement on a sample of the group
            for k in np.nonzero(ok)[0][:: max(1, int(ok.sum()) // 16)]:
                if int(first_bad[k]) != oracle.verify_bounds(A[k], b[k], res.x[k]):
                    mism.append(int(idxs[k]))
        for k, i in enumerate(idxs):
            xs[i] = res.x[k]
    out = {"count": count, "seed": seed, "nmax": nmax, "shapes": len(groups),
           "status_counts": {str(k): int(v) for k, v in zip(*np.unique(status, return_counts=True))},
           "engine_vs_checker_mismatches": len(set(mism)), "first_mismatches": sorted(set(mism))[:5],
           "bounded": bounded, "verify_bounds_violations_reproduced_from_reference": viol,
           "seconds": {"generate": round(t_gen, 3), "engine_calls": round(t_gpu, 3), "checker": round(t_cpu, 3)}}
    if scipy_sample:
        try:
            from scipy.optimize import linprog
            agree = disagree = objbad = 0
            smap = {0: 0, 2: -1, 3: 1}  # scipy status -> anst (optimal / infeasible / unbounded)
            for i in range(min(scipy_sample, count)):
                A, b, c = probs[i]
                if status[i] not in (-1, 0, 1):
                    continue
                r = linprog(-c, A_ub=A, b_ub=b, bounds=[(0, None)] * len(c), method="highs")
                ref = smap.get(r.status)
                if ref is None:
                    continue
                if ref == int(status[i]):
                    agree += 1
                    if ref == 0 and abs(float(c @ xs[i]) - float(c @ r.x)) > 1e-3:
                        objbad += 1
                else:
                    disagree += 1
            out["scipy_highs"] = {"checked": agree + disagree, "status_agree": agree,
                                  "status_disagree_reference_bugs": disagree, "objective_off_by_1e-3": objbad}
        except Exception as e:  # scipy is optional
            out["scipy_highs"] = {"skipped": repr(e)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--count", type=int, default=100000)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--nmax", type=int, default=7)
    ap.add_argument("--scipy", type=int, default=200)
    args = ap.parse_args()
    out = run(args.count, args.seed, args.nmax, args.scipy)
    print(json.dumps(out))
    sys.exit(1 if out["engine_vs_checker_mismatches"] else 0)


if __name__ == "__main__":
    main()
