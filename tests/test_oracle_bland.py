"""The corrected mode (rule="bland") has no reference behaviour to pin: the
CPU restatement in oracle/simplex_oracle.c is checked against scipy's HiGHS instead -- same
outcome class, same optimum to 1e-6 relative, x feasible -- and the GPU kernel is then checked
bit for bit against that restatement (tests/test_gpu_simplex.py)."""
import numpy as np
import pytest

scipy_opt = pytest.importorskip("scipy.optimize")


def _highs(A, b, c):
    n, m = A.shape
    kw = dict(A_ub=A, b_ub=b, bounds=[(0, None)] * m, method="highs")
    r = scipy_opt.linprog(-c, **kw)
    if r.status == 2:
        # HiGHS' presolve reports some unbounded problems as infeasible; ask again without it
        r = scipy_opt.linprog(-c, options={"presolve": False}, **kw)
    return {0: 0, 2: -1, 3: 1}.get(r.status), (-r.fun if r.status == 0 else None)


def _check(A, b, c, o):
    want, opt = _highs(A, b, c)
    assert want is not None
    assert o["status"] == want
    if want == 0:
        assert abs(o["objective"] - opt) <= 1e-6 * max(1.0, abs(opt))
        x = o["x"]
        assert x.min() >= -1e-9
        assert (A @ x - b).max() <= 1e-6 * max(1.0, float(np.abs(b).max()))
        assert abs(float(c @ x) - o["objective"]) <= 1e-6 * max(1.0, abs(opt))


@pytest.mark.parametrize("rule", ["bland", "dantzig", "devex"])
def test_bland_oracle_matches_highs_on_the_stress_distribution(oracle_lib, rule):
    """stress.py:23-26 distribution with zeros forced into b: every input on which the reference
    crashes (AttributeError / IndexError / TypeError) or cycles gets a proper answer here."""
    from linear_programming_b200 import workloads
    seen = set()
    for A, b, c in workloads.stress_batch(1500, seed=11):
        o = oracle_lib.simplex(A, b, c, rule=rule, max_pivots=5000)
        assert o["status"] in (0, -1, 1)
        _check(A, b, c, o)
        seen.add(o["status"])
    assert seen == {0, -1, 1}


@pytest.mark.parametrize("rule", ["bland", "dantzig", "devex"])
@pytest.mark.parametrize("dist,n,m,count", [("D3", 20, 20, 120), ("D2", 30, 40, 40), ("D3", 40, 25, 80)])
def test_bland_oracle_matches_highs_mid_size(oracle_lib, dist, n, m, count, rule):
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch(dist, count, n, m, seed=7)
    st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, rule=rule, max_pivots=200000)
    for p in range(count):
        _check(A[p], b[p], c[p], {"status": int(st[p]), "x": x[p], "objective": float(obj[p])})


@pytest.mark.parametrize("rule", ["bland", "dantzig", "devex"])
def test_bland_answers_the_reference_crash_fixtures(oracle_lib, simplex_golden, rule):
    """Every golden case on which the reference raises (status 3/4/5) is a well-posed LP."""
    crashed = [k for k in simplex_golden["cases"] if k["status"] in (3, 4, 5)]
    assert len(crashed) >= 10
    for case in crashed[:150]:
        A = np.array(case["A"], dtype=np.float64).reshape(case["n"], case["m"])
        b = np.array(case["b"], dtype=np.float64)
        c = np.array(case["c"], dtype=np.float64)
        _check(A, b, c, oracle_lib.simplex(A, b, c, rule=rule, max_pivots=20000))


def test_dantzig_needs_far_fewer_pivots_and_uses_the_bland_fallback(oracle_lib):
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch("D2", 12, 60, 60, seed=8)
    _, _, ob, nb, _ = oracle_lib.simplex_batch(A, b, c, rule="bland", max_pivots=200000)
    st, _, od, nd, _ = oracle_lib.simplex_batch(A, b, c, rule="dantzig", max_pivots=200000)
    assert nd.sum() * 2 < nb.sum()
    bounded = st == 0
    assert np.allclose(ob[bounded], od[bounded], rtol=1e-9)
    # a fully degenerate LP (b = 0): every pivot has ratio 0, the fallback must end it
    rng = np.random.default_rng(3)
    for _ in range(50):
        A1 = rng.integers(-5, 6, size=(6, 6)).astype(np.float64)
        o = oracle_lib.simplex(A1, np.zeros(6), rng.integers(-5, 6, size=6).astype(np.float64), rule="dantzig",
                               max_pivots=5000)
        assert o["status"] in (0, 1)


def test_reference_rule_is_untouched(oracle_lib, simplex_golden):
    """rule='reference' through the new entry point is the pinned oracle (spot check)."""
    for case in simplex_golden["cases"][:60]:
        A = np.array(case["A"], dtype=np.float64).reshape(case["n"], case["m"])
        o = oracle_lib.simplex(A, case["b"], case["c"], rule="reference", max_pivots=100000)
        assert o["status"] == case["status"] and o["npivots"] == case["npivots"]
