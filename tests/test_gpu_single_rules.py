"""GPU tests of the corrected pivoting rules (LPB200_RULE_BLAND / LPB200_RULE_DANTZIG) on the GRID-WIDE
single-LP kernel and its column-sharded form (lpb200_simplex_single_ex, lpb200_simplex_sharded_emulated_ex).

These rules are not reference behaviour: the CPU restatement
(oracle/simplex_oracle.c, `rule`) is checked against scipy HiGHS in tests/test_oracle_bland.py; here the
kernel has to reproduce that restatement bit for bit -- status, pivot sequence, x, optimum -- exactly as
the batch kernel's corrected modes do, so one LP gets the same answer whichever kernel serves it."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RULES = ["bland", "dantzig", "devex"]


def _same(res, g, o, cap):
    assert int(res.status[g]) == o["status"]
    assert int(res.npivots[g]) == o["npivots"]
    assert res.pivots(g) == o["pivots"][:cap]
    assert np.array_equal(res.x[g].view(np.int64), o["x"].view(np.int64))
    assert float(res.objective[g]).hex() == float(o["objective"]).hex()


def _stress_cases(oracle_lib, rule, base, scan=20000, seed=13):
    """The first `base` LPs of the stress distribution plus every LP among the first `scan` on which the
    restatement performs an x_aux drive-out (about 1 in 4000; found with the C oracle in a second or two)."""
    from linear_programming_b200 import workloads
    picked, driveouts = [], 0
    for i, (A, b, c) in enumerate(workloads.stress_batch(scan, seed=seed)):
        oracle_lib.driveouts(reset=True)
        oracle_lib.simplex(A, b, c, max_pivots=5000, rule=rule)
        hit = oracle_lib.driveouts(reset=True) > 0
        driveouts += hit
        if i < base or hit:
            picked.append((A, b, c))
    assert driveouts >= 3
    return picked


@pytest.mark.parametrize("rule", RULES)
def test_stress_distribution_single_mode_vs_oracle(oracle_lib, rule):
    """The stress.py distribution with zeros in b: phase 1, infeasible / unbounded outcomes and the
    x_aux drive-out (one extra key exchange + forced pivot), one cooperative launch per LP."""
    import linear_programming_b200 as lp
    groups = {}
    for A, b, c in _stress_cases(oracle_lib, rule, 1000):
        groups.setdefault(A.shape, []).append((A, b, c))
    seen = set()
    for (n, m), lst in groups.items():
        A = np.stack([q[0] for q in lst]); b = np.stack([q[1] for q in lst]); c = np.stack([q[2] for q in lst])
        res = lp.solve_batch(A, b, c, log_cap=64, max_pivots=5000, rule=rule, mode="single")
        for p in range(len(lst)):
            o = oracle_lib.simplex(A[p], b[p], c[p], max_pivots=5000, rule=rule)
            _same(res, p, o, 64)
            seen.add(o["status"])
    assert seen == {0, -1, 1}


@pytest.mark.parametrize("rule", RULES)
@pytest.mark.parametrize("dist,n,m", [("D1", 150, 700), ("D2", 96, 300), ("D3", 120, 260), ("D3", 300, 40)])
def test_mid_size_single_mode_vs_oracle(oracle_lib, rule, dist, n, m):
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch(dist, 1, n, m, seed=9)
    cap = 1500
    o = oracle_lib.simplex(A[0], b[0], c[0], max_pivots=cap, rule=rule)
    res = lp.solve_batch(A, b, c, max_pivots=cap, log_cap=cap, rule=rule, mode="single")
    _same(res, 0, o, cap)


@pytest.mark.parametrize("rule", RULES)
@pytest.mark.parametrize("nranks,dist,n,m", [(2, "D1", 150, 700), (3, "D2", 96, 300), (8, "D3", 120, 260),
                                             (8, "D2", 60, 2100)])
def test_mid_size_sharded_ranks_vs_oracle(oracle_lib, rule, nranks, dist, n, m):
    """Dantzig's (coefficient bits, variable id) key travels as two packets per rank; every rank must
    agree with the restatement."""
    from linear_programming_b200 import workloads
    from linear_programming_b200.sharded import solve_sharded_emulated
    A, b, c = workloads.simplex_batch(dist, 1, n, m, seed=9)
    cap = 1500
    o = oracle_lib.simplex(A[0], b[0], c[0], max_pivots=cap, rule=rule)
    res = solve_sharded_emulated(A[0], b[0], c[0], nranks, max_pivots=cap, log_cap=cap, rule=rule)
    for g in range(nranks):
        _same(res, g, o, cap)


@pytest.mark.parametrize("rule", RULES)
@pytest.mark.parametrize("nranks", [2, 5])
def test_stress_distribution_sharded_ranks_vs_oracle(oracle_lib, rule, nranks):
    """Small LPs over more ranks than some of them have columns: empty ranks, the auxiliary column on the
    last rank, drive-outs through the exchange."""
    from linear_programming_b200.sharded import solve_sharded_emulated
    seen = set()
    for A, b, c in _stress_cases(oracle_lib, rule, 200):
        o = oracle_lib.simplex(A, b, c, max_pivots=5000, rule=rule)
        res = solve_sharded_emulated(A, b, c, nranks, max_pivots=5000, log_cap=64, rule=rule)
        for g in range(nranks):
            _same(res, g, o, 64)
        seen.add(o["status"])
    assert seen == {0, -1, 1}


def test_dantzig_bland_fallback_on_degenerate_lps(oracle_lib):
    """b = 0: every pivot has ratio 0, so after 8 of them Bland's choice takes over (and ends the loop)."""
    import linear_programming_b200 as lp
    rng = np.random.default_rng(3)
    longest = 0
    for _ in range(40):
        A = rng.integers(-5, 6, size=(1, 12, 12)).astype(np.float64)
        c = rng.integers(-5, 6, size=(1, 12)).astype(np.float64)
        b = np.zeros((1, 12))
        o = oracle_lib.simplex(A[0], b[0], c[0], rule="dantzig", max_pivots=5000)
        res = lp.solve_batch(A, b, c, rule="dantzig", mode="single", max_pivots=5000, log_cap=128)
        _same(res, 0, o, 128)
        longest = max(longest, o["npivots"])
    assert longest > 12  # several cases ran into the fallback


@pytest.mark.parametrize("rule", RULES)
def test_same_answer_from_both_kernels(rule):
    """One LP, the batch kernel and the grid-wide kernel: identical pivot sequence and x bits."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch("D3", 6, 50, 70, seed=31)
    r1 = lp.solve_batch(A, b, c, rule=rule, mode="batch", max_pivots=20000, log_cap=256)
    r2 = lp.solve_batch(A, b, c, rule=rule, mode="single", max_pivots=20000, log_cap=256)
    assert np.array_equal(r1.status, r2.status) and np.array_equal(r1.npivots, r2.npivots)
    assert np.array_equal(r1.pivot_log, r2.pivot_log)
    assert np.array_equal(r1.x.view(np.int64), r2.x.view(np.int64))
    assert np.array_equal(r1.objective.view(np.int64), r2.objective.view(np.int64))


def test_single_mode_rejects_an_unknown_rule_code():
    from linear_programming_b200 import _lib
    L = _lib.lib()
    z = np.zeros(4)
    st = np.zeros(1, dtype=np.int32)
    rc = L.lpb200_simplex_single_ex(1, 1, z.ctypes.data, z.ctypes.data, z.ctypes.data, 1e-9, 10, 7,
                                    st.ctypes.data, z.ctypes.data, z.ctypes.data, st.ctypes.data, None, 0, None)
    assert rc == -1 and b"rule" in L.lpb200_last_error()
