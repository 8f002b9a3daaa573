"""Stream mode of the stdin/stdout shims.

CPU part: parsing, grouping by shape, ordering and formatting, with the engine call
substituted by the checker (the product itself never imports oracle/).  GPU part: the real
engine on the concatenated fixtures of the reference, against the reference's own answers."""
import io
import sys

import numpy as np
import pytest


def _assertion_fires(case):
    """Would the reference's own post-condition (simplex.py:745 -> verify_bounds :663-671) kill the process?"""
    import oracle
    if case["status"] != 0:
        return False
    A = np.array(case["A"], dtype=np.float64).reshape(case["n"], case["m"])
    x = np.array([float.fromhex(h) for h in case["x"]])
    return oracle.verify_bounds(A, np.array(case["b"], dtype=np.float64), x) >= 0


def _expected_stdout(case, verify=False):
    from linear_programming_b200.cli import format_answer
    names = {3: "AttributeError\n", 4: "IndexError\n", 5: "TypeError\n"}
    st = case["status"]
    if st in names:
        return names[st]
    if verify and _assertion_fires(case):
        return "AssertionError\n"
    x = [float.fromhex(h) for h in case["x"]] if case["x"] is not None else None
    return format_answer(st, x)


def _fixture_cases(simplex_golden, limit=None):
    cases = [c for c in simplex_golden["cases"] if c.get("stdin")]
    return cases[:limit] if limit else cases


def test_iter_lp_problems_roundtrip_and_errors():
    from linear_programming_b200.cli import iter_lp_problems
    text = "2 3\n1 2 3\n4 5 6\n7 8\n9 10 11\n" + "1 1\n-5\n3\n2\n"
    got = list(iter_lp_problems(text, integer=True))
    assert [(g[0], g[1]) for g in got] == [(2, 3), (1, 1)]
    assert got[0][2].tolist() == [[1, 2, 3], [4, 5, 6]] and got[0][3].tolist() == [7, 8]
    assert got[1][4].tolist() == [2]
    with pytest.raises(ValueError):
        list(iter_lp_problems("2 2\n1 2\n3 4\n5\n", integer=True))       # truncated
    with pytest.raises(ValueError):
        list(iter_lp_problems("1 1\n1.5\n2\n3\n", integer=True))          # int() like simplex.py:712-717
    assert list(iter_lp_problems("1 1\n1.5\n2\n3\n", integer=False))[0][2][0, 0] == 1.5   # cin >> double
    assert list(iter_lp_problems("  \n", integer=True)) == []


def test_lp_stream_orders_and_formats_like_the_reference(monkeypatch, oracle_lib, simplex_golden):
    from linear_programming_b200 import cli

    def checker(A, b, c):
        st, x, _, _, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=100000)
        return st, x
    monkeypatch.setattr(cli, "_solve_lp_batch", checker)
    cases = _fixture_cases(simplex_golden)
    assert len(cases) >= 30
    # interleave shapes so that grouping has to restore the input order
    text = "\n".join(c["stdin"] for c in cases)
    assert cli.solve_lp_stream(text, integer=True) == "".join(_expected_stdout(c) for c in cases)


def test_equation_stream_on_cpu(monkeypatch, oracle_lib, gj_golden):
    from linear_programming_b200 import cli, gauss_jordan
    monkeypatch.setattr(cli, "_solve_eq_batch", lambda a, b: oracle_lib.gauss_jordan_batch(a, b)[0])
    cases = [c for c in gj_golden["cases"] if c.get("stdin") is not None]
    text = "\n".join(c["stdin"] for c in cases)
    want = "".join(gauss_jordan.format_column([float.fromhex(h) for h in c["x"]]) for c in cases)
    assert cli.solve_equation_stream(text) == want
    assert cli.solve_equation_stream("0\n") == ""     # size 0 prints nothing (tests/01)


@pytest.mark.gpu
def test_stream_cli_on_gpu_matches_reference_answers(monkeypatch, capsys, simplex_golden, gj_golden):
    from linear_programming_b200 import cli, gauss_jordan
    cases = _fixture_cases(simplex_golden)
    monkeypatch.setattr(sys, "stdin", io.StringIO("\n".join(c["stdin"] for c in cases)))
    assert cli.main(["simplex", "--stream"]) == 0
    assert capsys.readouterr().out == "".join(_expected_stdout(c, verify=True) for c in cases)
    # the same stream through the C++-stub reader (doubles; the stubs' main() asserts nothing)
    monkeypatch.setattr(sys, "stdin", io.StringIO("\n".join(c["stdin"] for c in cases)))
    assert cli.main(["ad_allocation", "--stream"]) == 0
    assert capsys.readouterr().out == "".join(_expected_stdout(c) for c in cases)
    gcases = [c for c in gj_golden["cases"] if c.get("stdin") is not None]
    monkeypatch.setattr(sys, "stdin", io.StringIO("\n".join(c["stdin"] for c in gcases)))
    assert cli.main(["energy_values", "--stream"]) == 0
    want = "".join(gauss_jordan.format_column([float.fromhex(h) for h in c["x"]]) for c in gcases)
    assert capsys.readouterr().out == want


@pytest.mark.gpu
def test_single_problem_cli_is_byte_compatible(monkeypatch, capsys, simplex_golden):
    from linear_programming_b200 import cli
    for case in _fixture_cases(simplex_golden, limit=12):
        if case["status"] not in (-1, 0, 1):
            continue
        monkeypatch.setattr(sys, "stdin", io.StringIO(case["stdin"]))
        if _assertion_fires(case):  # the answer is printed, then the process dies like the reference's (:745)
            with pytest.raises(AssertionError):
                cli.main(["simplex"])
        else:
            assert cli.main(["simplex"]) == 0
        assert capsys.readouterr().out == _expected_stdout(case)


def test_rule_flag_reaches_the_batch_call(monkeypatch, oracle_lib):
    """`--rule=bland|dantzig` (corrected modes) is handed to the batch solve; the default stays the
    reference's behaviour, on an input where the two differ (all-zero b: AttributeError class)."""
    from linear_programming_b200 import cli
    seen = []

    def checker(A, b, c, rule="reference"):
        seen.append(rule)
        st, x, _, _, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=100000, rule=rule)
        return st, x
    monkeypatch.setattr(cli, "_solve_lp_batch", checker)
    monkeypatch.setattr(cli, "_verify_lp_batch", lambda A, b, x, status: np.full(len(status), -1, dtype=np.int32))
    text = "2 2\n1 1\n1 -1\n0 0\n1 1\n"          # b == 0: the reference raises (simplex.py:626)
    assert cli.solve_lp_stream(text, integer=True) == "AttributeError\n"
    out = cli.solve_lp_stream(text, integer=True, rule="bland")
    assert out.startswith("Bounded solution\n") or out == "Infinity\n"
    assert cli.solve_lp_stream(text, integer=True, rule="dantzig") == out
    assert seen == ["reference", "bland", "dantzig"]
    monkeypatch.setattr(sys, "stdin", io.StringIO(text))
    buf = io.StringIO()
    monkeypatch.setattr(sys, "stdout", buf)
    assert cli.main(["simplex", "--stream", "--rule=dantzig"]) == 0
    assert buf.getvalue() == out


def _text_of(A, b, c):
    n, m = A.shape
    rows = ["%d %d" % (n, m)] + [" ".join(str(int(v)) for v in r) for r in A]
    return "\n".join(rows + [" ".join(str(int(v)) for v in b), " ".join(str(int(v)) for v in c)]) + "\n"


def test_stream_runs_the_reference_post_condition_and_max_pivots_reaches_the_engine(monkeypatch, oracle_lib):
    """(1) the reference's __main__ asserts verify_bounds after every "Bounded solution"
    (simplex.py:745 -> :663-671) -- on answers its buggy ratio test gets wrong the process dies with AssertionError;
    the `simplex` shim does the same, and in --stream mode prints "AssertionError" for that problem.
    (2) --max-pivots=N reaches the batch call (the reference has no cap; the engine's default is documented)."""
    from linear_programming_b200 import cli, workloads
    seen = {}

    def checker(A, b, c, rule="reference", max_pivots=None):
        seen["max_pivots"] = max_pivots
        st, x, _, _, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=max_pivots or 100000, rule=rule)
        return st, x

    def verifier(A, b, x, status):
        return np.array([oracle_lib.verify_bounds(A[i], b[i], x[i]) if int(status[i]) == 0 else -1
                         for i in range(len(status))], dtype=np.int32)
    monkeypatch.setattr(cli, "_solve_lp_batch", checker)
    monkeypatch.setattr(cli, "_verify_lp_batch", verifier)
    # find, in the stress.py distribution, answers on which the reference's own assertion fires
    bad, good = None, None
    for A, b, c in workloads.stress_batch(4000, seed=3):
        r = oracle_lib.simplex(A, b, c, max_pivots=10000)
        if r["status"] != 0:
            continue
        if oracle_lib.verify_bounds(A, b, r["x"]) >= 0:
            bad = bad or (A, b, c)
        else:
            good = good or (A, b, c, r["x"])
        if bad and good:
            break
    assert bad is not None and good is not None
    text = _text_of(*bad) + _text_of(*good[:3])
    out = cli.solve_lp_stream(text, integer=True, verify=True)
    assert out == "AssertionError\n" + cli.format_answer(0, [float(v) for v in good[3]])
    assert cli.solve_lp_stream(text, integer=True).startswith("Bounded solution\n")   # the C++ stubs assert nothing
    # --max-pivots through main()
    monkeypatch.setattr(sys, "stdin", io.StringIO(_text_of(*good[:3])))
    buf = io.StringIO()
    monkeypatch.setattr(sys, "stdout", buf)
    assert cli.main(["simplex", "--stream", "--max-pivots=777"]) == 0
    assert seen["max_pivots"] == 777
    monkeypatch.setattr(sys, "stdin", io.StringIO(_text_of(*good[:3])))
    monkeypatch.setattr(sys, "stdout", io.StringIO())
    assert cli.main(["simplex", "--stream", "--max-pivots=1"]) == 0 and seen["max_pivots"] == 1
    assert sys.stdout.getvalue() in ("IterationLimit\n", buf.getvalue())
