"""Instrumented import of the UNMODIFIED reference solver (test infrastructure only).

This module is used in the build container (where /root/reference is mounted)
to (a) pin the C restatement in ``oracle/simplex_oracle.c`` against the real
reference and (b) generate the golden vectors committed under ``tests/golden``
(see ``tests/golden/make_golden.py``).  It is never imported by the product
package, by ``bench.py`` or by the ``-m gpu`` tests: the reference tree does not
exist on the GPU box.

Reference entry points exercised (file:line in /root/reference):
  * ``Simplex(A, b, c, n, m).solve()``   week-2/diet/simplex.py:496-505, 586-630
  * ``SlackForm.pivot(entering, leaving)`` week-2/diet/simplex.py:348-385
    (wrapped to record the pivot sequence; behaviour untouched)
  * ``Simplex.parse(stream)``            week-2/diet/simplex.py:707-718
"""
from __future__ import annotations

import importlib.util
import io
import os
import signal
import sys

REFERENCE_ROOT = os.environ.get("LP_REFERENCE_ROOT", "/root/reference")
SIMPLEX_PY = os.path.join(REFERENCE_ROOT, "week-2", "diet", "simplex.py")

# Outcome codes shared with the oracle and the CUDA engine (include/lp_b200.h).
STATUS_BOUNDED = 0
STATUS_NO_SOLUTION = -1
STATUS_INFINITY = 1
STATUS_ITERATION_LIMIT = 2      # engine/oracle only: the reference has no cap and hangs
STATUS_REF_ATTRIBUTE_ERROR = 3  # simplex.py:626  (every b_i is 0 / 0 < min b <= tol)
STATUS_REF_INDEX_ERROR = 4      # simplex.py:541 -> :183 (x_aux still basic after phase 1)
STATUS_REF_TYPE_ERROR = 5       # simplex.py:409  (-1 * None in the ratio test)

_EXC_TO_STATUS = {
    "AttributeError": STATUS_REF_ATTRIBUTE_ERROR,
    "IndexError": STATUS_REF_INDEX_ERROR,
    "TypeError": STATUS_REF_TYPE_ERROR,
}


def available() -> bool:
    return os.path.isfile(SIMPLEX_PY)


_ref = None


def load_reference():
    """Import week-2/diet/simplex.py unmodified (its __main__ guard keeps it inert)."""
    global _ref
    if _ref is None:
        spec = importlib.util.spec_from_file_location("_lp_reference_simplex", SIMPLEX_PY)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _ref = mod
    return _ref


class _Timeout(Exception):
    pass


def _alarm(signum, frame):
    raise _Timeout()


def run_reference(A, b, c, timeout_s: float = 60.0, tolerance: float | None = None):
    """Solve one LP with the reference and record what parity is defined on.

    Returns dict(status, x, optimum, pivots=[(entering, leaving), ...], python=...).
    ``status`` is the reference's anst (-1/0/1) or one of the crash codes above;
    ``None`` status means the reference did not terminate within ``timeout_s``.
    """
    ref = load_reference()
    n, m = len(b), len(c)
    pivots = []
    orig_pivot = ref.SlackForm.pivot

    def recording_pivot(self, entering_idx, leaving_idx):
        pivots.append((int(entering_idx), int(leaving_idx)))
        return orig_pivot(self, entering_idx, leaving_idx)

    old_tol = ref.global_tolerance
    if tolerance is not None:
        ref.global_tolerance = float(tolerance)
    ref.SlackForm.pivot = recording_pivot
    old_handler = signal.signal(signal.SIGALRM, _alarm)
    signal.setitimer(signal.ITIMER_REAL, timeout_s)
    out = {"status": None, "x": None, "optimum": None, "pivots": pivots,
           "python": "%d.%d.%d" % sys.version_info[:3]}
    try:
        s = ref.Simplex([list(r) for r in A], list(b), list(c), n, m)
        try:
            anst, ansx = s.solve()
            out["status"] = int(anst)
            if anst == 0:
                out["x"] = [float(v) for v in ansx]
                out["optimum"] = float(s.optimum)
        except _Timeout:
            out["status"] = None
        except Exception as e:  # the reference's crash classes
            out["status"] = _EXC_TO_STATUS.get(type(e).__name__, 99)
            out["exception"] = type(e).__name__
    except _Timeout:
        out["status"] = None
    finally:
        signal.setitimer(signal.ITIMER_REAL, 0)
        signal.signal(signal.SIGALRM, old_handler)
        ref.SlackForm.pivot = orig_pivot
        ref.global_tolerance = old_tol
    return out


def parse_reference(text: str):
    """Run the reference's own parser (simplex.py:707-718) on a problem text."""
    ref = load_reference()
    s = ref.Simplex.parse(stream=io.StringIO(text))
    return s.A, s.b, s.c
