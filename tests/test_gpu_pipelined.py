"""The pipelined pivot loop of the grid-wide / column-sharded kernel (simplex_single.cu: ping-pong tableau, a control
group of CTAs that selects pivot k+1 while the others sweep pivot k, one grid barrier per pivot) against the oracle:
status, pivot sequence, x, objective AND the tableau it stops on (objective row, constant column) -- for every size of
the control group, for pivot budgets that end the loop on either tableau buffer, for both rules that use the loop
(the reference's, simplex.py:387-436, and Bland), for LPs that go through phase 1 (where the loop runs twice, around
the step-by-step forced pivot), on one rank and on emulated column-sharded ranks; and against the step-by-step loop
(LPB_DEBUG_SINGLE_PIPELINE=0), which the other rules still use."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _solve(A, b, c, cap, rule, nranks):
    from linear_programming_b200.sharded import solve_sharded_emulated
    return solve_sharded_emulated(A, b, c, nranks, max_pivots=cap, log_cap=max(cap, 1), rule=rule, state=True)


def _check(res, want, nranks, nv, name):
    for g in range(nranks):
        assert int(res.status[g]) == want["status"], name
        assert res.pivots(g) == want["pivots"], name
        assert np.array_equal(res.x[g].view(np.int64), want["x"].view(np.int64)), name
        assert np.float64(res.objective[g]).view(np.int64) == np.float64(want["objective"]).view(np.int64), name
        assert np.array_equal(res.const_by_var[g], want["const_by_var"]), name
        assert res.obj_by_var[g, nv] == want["obj_const"], name
    assert np.array_equal(res.obj_by_var[:, :nv].sum(axis=0), want["obj_by_var"]), name


@pytest.mark.parametrize("nctl", ["1", "3", "16"])
@pytest.mark.parametrize("dist,n,m", [("D1", 130, 1100), ("D2", 90, 700), ("D3", 150, 420)])
def test_control_group_sizes_and_both_buffers(oracle_lib, monkeypatch, nctl, dist, n, m):
    from linear_programming_b200 import workloads
    monkeypatch.setenv("LPB_DEBUG_SINGLE_NCTL", nctl)
    A, b, c = (v[0] for v in workloads.simplex_batch(dist, 1, n, m, seed=17))
    nv = n + m + 1
    for rule in ("reference", "bland"):
        for cap in (1, 2, 7, 40, 41, 100000 if n * m < 70000 else 90):
            want = oracle_lib.simplex_state(A, b, c, max_pivots=cap, rule=rule, threads=2)
            for nranks in (1, 3):
                _check(_solve(A, b, c, cap, rule, nranks), want, nranks, nv,
                       "%s %dx%d rule %s cap %d ranks %d nctl %s" % (dist, n, m, rule, cap, nranks, nctl))


def test_pipelined_equals_step_by_step(oracle_lib, monkeypatch):
    """Same LP through both loops: identical bits, whichever buffer the pipelined one ends on."""
    from linear_programming_b200 import workloads
    A, b, c = (v[0] for v in workloads.simplex_batch("D3", 1, 200, 900, seed=19))
    nv = 200 + 900 + 1
    for cap in (33, 34):
        runs = []
        for mode in ("1", "0"):
            monkeypatch.setenv("LPB_DEBUG_SINGLE_PIPELINE", mode)
            runs.append(_solve(A, b, c, cap, "reference", 8))
        a, o = runs
        assert np.array_equal(a.status, o.status) and np.array_equal(a.pivot_log, o.pivot_log)
        assert np.array_equal(a.obj_by_var.view(np.int64) & 0x7fffffffffffffff, o.obj_by_var.view(np.int64) & 0x7fffffffffffffff)
        assert np.array_equal(a.const_by_var, o.const_by_var)
        want = oracle_lib.simplex_state(A, b, c, max_pivots=cap, threads=2)
        _check(a, want, 8, nv, "pipelined cap %d" % cap)
        _check(o, want, 8, nv, "step by step cap %d" % cap)


def test_every_outcome_class_through_the_pipelined_loop(simplex_golden, monkeypatch):
    """Golden cases of the reference (all six outcome classes, phase 1 included) with at least a few pivots, on the
    grid-wide kernel with the full grid forced, so that the loop really has control CTAs and sweepers."""
    import linear_programming_b200 as lp
    monkeypatch.setenv("LPB_DEBUG_SINGLE_CPR", "24")  # small LPs get 1-5 CTAs by default: too few for a control group
    cases = [c for c in simplex_golden["cases"] if 3 <= c["npivots"] <= 150]
    by_status = {}
    for c in cases:
        by_status.setdefault(c["status"], []).append(c)
    assert set(by_status) >= {-1, 0, 1, 5}
    n_checked = 0
    for st, lst in sorted(by_status.items()):
        for case in sorted(lst, key=lambda c: -c["npivots"])[:8]:
            A = np.array(case["A"], dtype=np.float64).reshape(case["n"], case["m"])
            b, c = np.array(case["b"], dtype=np.float64), np.array(case["c"], dtype=np.float64)
            cap = case["npivots"] + 3
            res = lp.solve_batch(A[None], b[None], c[None], mode="single", max_pivots=cap, log_cap=cap)
            assert int(res.status[0]) == case["status"], case["name"]
            assert [list(q) for q in res.pivots(0)] == case["pivots"], case["name"]
            if case["status"] == 0:
                assert [float(v).hex() for v in res.x[0]] == case["x"], case["name"]
                assert float(res.objective[0]).hex() == case["optimum"], case["name"]
            n_checked += 1
    assert n_checked >= 20
