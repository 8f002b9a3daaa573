"""stdin -> stdout shims byte-compatible with the reference programs.

    python -m linear_programming_b200.cli simplex        < problem   (week-2/diet/simplex.py:720-751)
    python -m linear_programming_b200.cli ad_allocation  < problem   (ad_allocation.cpp:21-56)
    python -m linear_programming_b200.cli diet           < problem   (diet.cc:17-58)
    python -m linear_programming_b200.cli energy_values  < system    (energy_values.cc:207-212)

`simplex` parses with int() like the reference (:712-717) and prints ' '.join('%.18f');
the C++ shims accept floats (cin >> double) and print "%.18f" separated by spaces.
Reference crash classes propagate as the same Python exception (non-zero exit), as in the
reference process.
"""
from __future__ import annotations

import sys

from . import gauss_jordan
from .simplex import Simplex


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


def main(argv=None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    prog = argv[0] if argv else "simplex"
    if prog == "simplex":
        s = Simplex.parse(sys.stdin)
        anst, ansx = s.solve()
        sys.stdout.write(format_answer(anst, ansx))
    elif prog in ("ad_allocation", "diet"):
        n, m, A, b, c = parse_lp_floats(sys.stdin.read())
        anst, ansx = Simplex(A, b, c, n, m).solve()
        sys.stdout.write(format_answer(anst, ansx))
    elif prog == "energy_values":
        a, b = gauss_jordan.read_equation(sys.stdin.read())
        sys.stdout.write(gauss_jordan.format_column(gauss_jordan.solve_equation(a, b)))
    else:
        sys.stderr.write("unknown program %r\n" % prog)
        return 2
    return 0


if __name__ == "__main__":
    sys.exit(main())
