"""The threaded form of the CPU oracle (oracle/simplex_oracle.c lp_oracle_simplex_state) that makes a
config-5-size LP checkable: the rows of one pivot (Constraint.substitute, simplex.py:185-214, applied by
SlackForm.pivot :348-362) are independent of each other, so sweeping them on any number of host threads must
give the bits of the scalar loop -- status, pivot sequence, x, and the exported tableau."""
import numpy as np
import pytest

from linear_programming_b200 import workloads


def _same_state(a, b):
    assert a["status"] == b["status"] and a["pivots"] == b["pivots"]
    assert np.array_equal(a["x"].view(np.int64), b["x"].view(np.int64))
    assert np.float64(a["objective"]).view(np.int64) == np.float64(b["objective"]).view(np.int64)
    assert np.array_equal(a["obj_by_var"].view(np.int64), b["obj_by_var"].view(np.int64))
    assert np.array_equal(a["const_by_var"].view(np.int64), b["const_by_var"].view(np.int64))
    assert np.float64(a["obj_const"]).view(np.int64) == np.float64(b["obj_const"]).view(np.int64)


@pytest.mark.parametrize("dist,n,m,cap", [("D1", 300, 1200, 60), ("D2", 260, 520, 90), ("D3", 310, 290, 120)])
def test_thread_count_does_not_change_a_bit(oracle_lib, dist, n, m, cap):
    if dist == "D1":
        A, b, c = workloads.single_lp(n, m, seed=5)
    else:
        A, b, c = (v[0] for v in workloads.simplex_batch(dist, 1, n, m, seed=11))
    scalar = oracle_lib.simplex(A, b, c, max_pivots=cap)
    base = oracle_lib.simplex_state(A, b, c, max_pivots=cap, threads=1)
    assert scalar["status"] == base["status"] and scalar["pivots"] == base["pivots"]
    assert np.array_equal(scalar["x"].view(np.int64), base["x"].view(np.int64))
    for threads in (2, 3, 7, 0):
        _same_state(oracle_lib.simplex_state(A, b, c, max_pivots=cap, threads=threads), base)


def test_exported_state_is_the_answer_at_the_optimum(oracle_lib):
    """At status 0 the export must agree with what Simplex.solve returns (simplex.py:473-491): x_j = constant of
    x_j's row when basic, optimum = the objective row's constant, and no positive objective coefficient is left."""
    A, b, c = workloads.simplex_batch("D3", 6, 35, 28, seed=3)
    seen = 0
    for i in range(6):
        r = oracle_lib.simplex_state(A[i], b[i], c[i], threads=2)
        if r["status"] != 0:
            continue
        seen += 1
        n, m = 35, 28
        assert np.array_equal(r["const_by_var"][n:n + m], r["x"])
        assert r["obj_const"] == r["objective"]
        assert not (r["obj_by_var"] > 1e-9).any()
    assert seen >= 2


def test_attribute_error_class_exports_zeros(oracle_lib):
    A = np.array([[1.0, 2.0], [3.0, 1.0]])
    r = oracle_lib.simplex_state(A, np.zeros(2), np.array([1.0, 1.0]))
    assert r["status"] == 3 and not r["obj_by_var"].any() and not r["const_by_var"].any()
