"""stdin -> stdout shims byte-compatible with the reference programs.

    python -m linear_programming_b200.cli simplex        < problem   (week-2/diet/simplex.py:720-751)
    python -m linear_programming_b200.cli ad_allocation  < problem   (ad_allocation.cpp:21-56)
    python -m linear_programming_b200.cli diet           < problem   (diet.cc:17-58)
    python -m linear_programming_b200.cli energy_values  < system    (energy_values.cc:207-212)

`simplex` parses with int() like the reference (:712-717) and prints ' '.join('%.18f');
the C++ shims accept floats (cin >> double) and print "%.18f" separated by spaces.
Reference crash classes propagate as the same Python exception (non-zero exit), as in the
reference process.

Stream mode -- add ``--stream``: stdin holds MANY problems of the same
program, simply concatenated (whitespace separated, exactly the single-problem format
repeated).  They are grouped by shape, each group goes to the GPU as ONE batched call
(lpb200_simplex_batch / lpb200_gauss_jordan_batch), and the answers are printed in input
order, each exactly as the single-problem program prints it.  An input on which the
reference process dies prints the exception class name on a line of its own
("AttributeError" / "IndexError" / "TypeError", or "IterationLimit") and the stream goes on,
so `stress.py`-style drivers (week-2/diet/stress.py:30) can pipe a whole run through one
process instead of one subprocess per LP.

``--rule=bland`` (LP programs only) answers with the corrected mode instead of the reference's
behaviour (include/lp_b200.h LPB200_RULE_BLAND): no crash classes, Bland's anti-cycling rule;
``--rule=dantzig`` / ``--rule=devex`` the same with largest-coefficient / Devex pricing.

``--max-pivots=N`` (LP programs) raises or lowers the one deliberate deviation from the reference: the
engine stops an LP after max(10000, 50(n+m)) pivots (IterationLimitError / "IterationLimit") where the
reference, which has no cap, would keep pivoting -- its set-order entering rule needs thousands of pivots
on mixed-sign data and can cycle.

The `simplex` program also runs the reference's post-condition (simplex.py:745: verify_bounds after every
"Bounded solution"): when it fails the answer has been printed and the process then dies with
AssertionError like the reference does; in --stream mode such a problem prints "AssertionError" on a line
of its own instead of its answer (a per-process exit code has no meaning there).
"""
from __future__ import annotations

import sys

import numpy as np

from . import gauss_jordan
from . import simplex as _simplex
from .simplex import Simplex

_CRASH_NAMES = {
    _simplex.STATUS_REF_ATTRIBUTE_ERROR: "AttributeError",   # simplex.py:626
    _simplex.STATUS_REF_INDEX_ERROR: "IndexError",           # simplex.py:541 -> :183
    _simplex.STATUS_REF_TYPE_ERROR: "TypeError",             # simplex.py:409
    _simplex.STATUS_ITERATION_LIMIT: "IterationLimit",       # engine-only: the reference would hang
}


def format_answer(anst, ansx) -> str:
    """simplex.py:742-744 / ad_allocation.cpp:41-54."""
    out = Simplex.answer_type_str(anst) + "\n"
    if anst == 0:
        out += " ".join("%.18f" % v for v in ansx) + "\n"
    return out


def parse_lp_floats(text: str):
    """The C++ stubs' reader (ad_allocation.cpp:22-37): whitespace-separated numbers."""
    toks = text.split()
    n, m = int(toks[0]), int(toks[1])
    vals = [float(t) for t in toks[2:2 + n * m + n + m]]
    A = [vals[i * m:(i + 1) * m] for i in range(n)]
    b = vals[n * m:n * m + n]
    c = vals[n * m + n:n * m + n + m]
    return n, m, A, b, c


def iter_lp_problems(text: str, integer: bool):
    """Concatenated LP problems -> (n, m, A[n,m], b[n], c[m]) float64 arrays, in input order.
    integer=True applies the reference parser's int() (simplex.py:712-717: a non-integer token
    raises ValueError); integer=False reads doubles like the C++ stubs."""
    toks = text.split()
    conv = int if integer else float
    pos = 0
    while pos < len(toks):
        if pos + 2 > len(toks):
            raise ValueError("truncated problem header at token %d" % pos)
        n, m = int(toks[pos]), int(toks[pos + 1])
        need = n * m + n + m
        body = toks[pos + 2:pos + 2 + need]
        if n < 1 or m < 1 or len(body) < need:
            raise ValueError("truncated or malformed problem at token %d (n=%d, m=%d)" % (pos, n, m))
        vals = np.array([conv(t) for t in body], dtype=np.float64)
        yield n, m, vals[:n * m].reshape(n, m), vals[n * m:n * m + n], vals[n * m + n:]
        pos += 2 + need


def iter_equations(text: str):
    """Concatenated energy_values systems (energy_values.cc:66-80) -> (a[size,size], b[size])."""
    toks = text.split()
    pos = 0
    while pos < len(toks):
        size = int(toks[pos])
        need = size * (size + 1)
        body = toks[pos + 1:pos + 1 + need]
        if size < 0 or len(body) < need:
            raise ValueError("truncated system at token %d (size=%d)" % (pos, size))
        vals = np.array([float(t) for t in body], dtype=np.float64).reshape(size, size + 1)
        yield np.ascontiguousarray(vals[:, :size]), np.ascontiguousarray(vals[:, size])
        pos += 1 + need


def _solve_lp_batch(A, b, c, rule="reference", max_pivots=None):
    """One shape group on the engine; separated out so CPU tests can substitute the checker."""
    r = _simplex.solve_batch(A, b, c, rule=rule, max_pivots=max_pivots)
    return r.status, r.x


def _verify_lp_batch(A, b, x, status):
    """Simplex.verify_bounds (simplex.py:663-671) on one shape group: first failing row per LP, -1 = passes."""
    return _simplex.verify_bounds_batch(A, b, x, status)


def _solve_eq_batch(a, b):
    return gauss_jordan.solve_equation_batch(a, b)


def solve_lp_stream(text: str, integer: bool, rule: str = "reference", max_pivots=None, verify: bool = False) -> str:
    problems = list(iter_lp_problems(text, integer))
    groups = {}
    for idx, (n, m, _, _, _) in enumerate(problems):
        groups.setdefault((n, m), []).append(idx)
    out = [None] * len(problems)
    for (n, m), idxs in groups.items():
        A = np.stack([problems[i][2] for i in idxs])
        b = np.stack([problems[i][3] for i in idxs])
        c = np.stack([problems[i][4] for i in idxs])
        kw = {}
        if rule != "reference":
            kw["rule"] = rule
        if max_pivots is not None:
            kw["max_pivots"] = max_pivots
        status, x = _solve_lp_batch(A, b, c, **kw)
        bad = _verify_lp_batch(A, b, x, status) if verify else None
        for k, i in enumerate(idxs):
            st = int(status[k])
            if bad is not None and st == 0 and int(bad[k]) >= 0:
                out[i] = "AssertionError\n"   # simplex.py:745 -> :671
            elif st in (-1, 0, 1):
                out[i] = format_answer(st, [float(v) for v in x[k]])
            else:
                out[i] = _CRASH_NAMES.get(st, "Status%d" % st) + "\n"
    return "".join(out)


def solve_equation_stream(text: str) -> str:
    systems = list(iter_equations(text))
    groups = {}
    for idx, (a, _) in enumerate(systems):
        groups.setdefault(a.shape[0], []).append(idx)
    out = [None] * len(systems)
    for size, idxs in groups.items():
        if size == 0:
            for i in idxs:
                out[i] = ""   # size 0 prints nothing (week-2/energy_values/tests/01)
            continue
        a = np.stack([systems[i][0] for i in idxs])
        b = np.stack([systems[i][1] for i in idxs])
        x = _solve_eq_batch(a, b)
        for k, i in enumerate(idxs):
            out[i] = gauss_jordan.format_column(x[k])
    return "".join(out)


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    stream = "--stream" in argv
    rule = "reference"
    max_pivots = None
    for t in argv:
        if t.startswith("--rule="):
            rule = t.split("=", 1)[1]
        if t.startswith("--max-pivots="):
            max_pivots = int(t.split("=", 1)[1])
    argv = [t for t in argv if t != "--stream" and not t.startswith("--rule=") and not t.startswith("--max-pivots=")]
    prog = argv[0] if argv else "simplex"
    if prog not in ("simplex", "ad_allocation", "diet", "energy_values"):
        sys.stderr.write("unknown program %r\n" % prog)
        return 2
    if stream:
        text = sys.stdin.read()
        if prog == "energy_values":
            sys.stdout.write(solve_equation_stream(text))
        else:
            sys.stdout.write(solve_lp_stream(text, integer=(prog == "simplex"), rule=rule, max_pivots=max_pivots,
                                             verify=(prog == "simplex")))
        return 0
    if prog == "simplex":
        s = Simplex.parse(sys.stdin)
        s.rule = rule
        s.max_pivots = max_pivots
        anst, ansx = s.solve()
        sys.stdout.write(format_answer(anst, ansx))
        if anst == 0:  # simplex.py:745: the reference asserts its own answer after printing it
            sys.stdout.flush()
            s.verify_bounds()
    elif prog in ("ad_allocation", "diet"):
        n, m, A, b, c = parse_lp_floats(sys.stdin.read())
        s = Simplex(A, b, c, n, m)
        s.rule = rule
        s.max_pivots = max_pivots
        anst, ansx = s.solve()
        sys.stdout.write(format_answer(anst, ansx))
    else:
        a, b = gauss_jordan.read_equation(sys.stdin.read())
        sys.stdout.write(gauss_jordan.format_column(gauss_jordan.solve_equation(a, b)))
    return 0


if __name__ == "__main__":
    sys.exit(main())
