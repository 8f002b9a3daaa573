#!/usr/bin/env python3
"""Small, self-checking invocations of EVERY kernel variant of liblp_b200.so, meant to be run under
compute-sanitizer (tools/gpu/sanitize.sh: memcheck, racecheck, synccheck) -- SURVEY.md section 5
"race detection".  Sizes are tiny because the tools slow a kernel down 10-100x; every answer is still
compared with the CPU oracle bit for bit, so a sanitizer pass is also a parity pass.

    python tools/gpu/sanitize_cases.py [case ...]        (no argument: all cases)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def _batch(variant, rule, dist, count, n, m, seed=3):
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    import oracle
    if variant:
        os.environ["LPB_BATCH_KERNEL"] = variant
    else:
        os.environ.pop("LPB_BATCH_KERNEL", None)
    A, b, c = workloads.simplex_batch(dist, count, n, m, seed=seed)
    res = lp.solve_batch(A, b, c, max_pivots=4000, log_cap=64, rule=rule, mode="batch")
    st, x, obj, npiv, _ = oracle.simplex_batch(A, b, c, max_pivots=4000, rule=rule)
    assert np.array_equal(res.status, st), (variant, rule, "status")
    assert np.array_equal(res.npivots, npiv), (variant, rule, "npivots")
    assert np.array_equal(res.x.view(np.int64), x.view(np.int64)), (variant, rule, "x")
    os.environ.pop("LPB_BATCH_KERNEL", None)
    return int(npiv.sum())


def case_smem8():
    return _batch("smem", "reference", "D3", 6, 40, 36) + _batch("smem", "reference", "D2", 3, 100, 100)


def case_mid4():
    return _batch("mid", "reference", "D3", 6, 50, 60)


def case_small1():
    return _batch("small", "reference", "D3", 6, 34, 40)


def case_tiny1():
    return _batch("tiny", "reference", "D3", 12, 7, 9)


def case_reg15():
    return _batch("reg", "reference", "D3", 4, 60, 50)


def case_bland8():
    return _batch("smem", "bland", "D3", 4, 40, 36) + _batch(None, "bland", "D3", 8, 9, 7)


def case_dantzig8():
    return _batch("smem", "dantzig", "D3", 4, 40, 36) + _batch(None, "dantzig", "D3", 8, 9, 7)


def case_devex8():
    return _batch("smem", "devex", "D3", 4, 40, 36) + _batch(None, "devex", "D3", 8, 9, 7)


def case_single():
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    import oracle
    tot = 0
    for rule in ("reference", "devex"):
        A, b, c = workloads.simplex_batch("D3", 1, 70, 600, seed=5)
        res = lp.solve_batch(A, b, c, mode="single", max_pivots=60, log_cap=60, rule=rule)
        o = oracle.simplex(A[0], b[0], c[0], max_pivots=60, rule=rule)
        assert int(res.status[0]) == o["status"] and res.pivots(0) == o["pivots"], rule
        tot += o["npivots"]
    return tot


def case_sharded():
    from linear_programming_b200 import workloads
    from linear_programming_b200.sharded import solve_sharded_emulated
    import oracle
    tot = 0
    for nranks, rule in ((8, "reference"), (3, "dantzig")):
        A, b, c = workloads.simplex_batch("D3", 1, 48, 300, seed=6)
        o = oracle.simplex_state(A[0], b[0], c[0], max_pivots=40, rule=rule, threads=1)
        res = solve_sharded_emulated(A[0], b[0], c[0], nranks, max_pivots=40, log_cap=40, rule=rule, state=True)
        for g in range(nranks):
            assert int(res.status[g]) == o["status"] and res.pivots(g) == o["pivots"], (nranks, rule, g)
            assert np.array_equal(res.const_by_var[g], o["const_by_var"])
        tot += o["npivots"]
    return tot


def case_gj():
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    import oracle
    for n, env in ((20, None), (96, None), (200, "0"), (200, "1"), (64, None)):
        if env is None:
            os.environ.pop("LPB_DEBUG_GJ_PAIRS", None)
        else:
            os.environ["LPB_DEBUG_GJ_PAIRS"] = env
        a, b = workloads.gauss_jordan_batch(3, n=n, seed=2, swap_frac=0.6)
        for piv in ("reference", "partial"):
            got = lp.solve_equation_batch(a, b, pivoting=piv)
            want, _ = oracle.gauss_jordan_batch(a, b, pivoting=piv)
            assert np.array_equal(got.view(np.int64), want.view(np.int64)), (n, env, piv)
    os.environ.pop("LPB_DEBUG_GJ_PAIRS", None)
    return 0


def case_verify():
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    import oracle
    A, b, c = workloads.simplex_batch("D3", 6, 30, 26, seed=8)
    res = lp.solve_batch(A, b, c)
    bad = lp.verify_bounds_batch(A, b, res.x, res.status)
    for i in range(6):
        want = oracle.verify_bounds(A[i], b[i], res.x[i]) if int(res.status[i]) == 0 else -1
        assert int(bad[i]) == want
    return 0


CASES = {k[5:]: v for k, v in list(globals().items()) if k.startswith("case_")}


def main():
    names = sys.argv[1:] or list(CASES)
    for nm in names:
        piv = CASES[nm]()
        print("sanitize case %-10s ok (%d pivots, answers bit-exact vs oracle)" % (nm, piv), flush=True)


if __name__ == "__main__":
    main()
