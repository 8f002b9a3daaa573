"""CPU oracle for the hot path -- TEST INFRASTRUCTURE ONLY.

ctypes front-end of ``oracle/_build/liblp_oracle.so`` (our scalar C restatement of
/root/reference/week-2/diet/simplex.py and week-2/energy_values/energy_values.cc) and,
when it was built, of ``oracle/_ref/libref_energy_values.so`` (the UNMODIFIED reference
Gauss-Jordan compiled from its own source).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package -- as the checker or the timed CPU
baseline, never as the product.  ``linear_programming_b200`` does not import it.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "liblp_oracle.so")
REF_GJ_LIB_PATH = os.path.join(_HERE, "_ref", "libref_energy_values.so")
REF_GJ_BIN_PATH = os.path.join(_HERE, "_ref", "energy_values")

STATUS_BOUNDED = 0
STATUS_NO_SOLUTION = -1
STATUS_INFINITY = 1
STATUS_ITERATION_LIMIT = 2
STATUS_REF_ATTRIBUTE_ERROR = 3
STATUS_REF_INDEX_ERROR = 4
STATUS_REF_TYPE_ERROR = 5

_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int32)


def build(reference_root: str = "/root/reference", force: bool = False) -> None:
    """Compile the C restatement, and the reference shim when the reference tree is present."""
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".h"))]
    stale = (not os.path.exists(LIB_PATH)) or any(
        os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "all"], stdout=subprocess.DEVNULL)
    if os.path.isfile(os.path.join(reference_root, "week-2", "energy_values", "energy_values.cc")):
        if force or not (os.path.exists(REF_GJ_LIB_PATH) and os.path.exists(REF_GJ_BIN_PATH)):
            subprocess.check_call(["make", "-C", _HERE, "ref", "REFERENCE_ROOT=" + reference_root],
                                  stdout=subprocess.DEVNULL)


_lib = None
_ref_gj = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = ctypes.CDLL(LIB_PATH)
        L.lp_oracle_simplex.restype = ctypes.c_int
        L.lp_oracle_simplex.argtypes = [ctypes.c_int, ctypes.c_int, _dp, _dp, _dp, ctypes.c_double,
                                        ctypes.c_int, _dp, _dp, _ip, _ip, ctypes.c_int]
        L.lp_oracle_simplex_batch.restype = ctypes.c_int
        L.lp_oracle_simplex_batch.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, _dp, _dp, _dp,
                                              ctypes.c_double, ctypes.c_int, _ip, _dp, _dp, _ip,
                                              ctypes.c_int]
        L.lp_oracle_simplex_rule.restype = ctypes.c_int
        L.lp_oracle_simplex_rule.argtypes = [ctypes.c_int, ctypes.c_int, _dp, _dp, _dp, ctypes.c_double,
                                             ctypes.c_int, ctypes.c_int, _dp, _dp, _ip, _ip, ctypes.c_int]
        L.lp_oracle_simplex_state.restype = ctypes.c_int
        L.lp_oracle_simplex_state.argtypes = [ctypes.c_int, ctypes.c_int, _dp, _dp, _dp, ctypes.c_double,
                                              ctypes.c_int, ctypes.c_int, ctypes.c_int, _dp, _dp, _ip, _ip,
                                              ctypes.c_int, _dp, _dp]
        L.lp_oracle_simplex_batch_rule.restype = ctypes.c_int
        L.lp_oracle_simplex_batch_rule.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, _dp, _dp, _dp,
                                                   ctypes.c_double, ctypes.c_int, ctypes.c_int, _ip, _dp, _dp,
                                                   _ip, ctypes.c_int]
        L.lp_oracle_gauss_jordan.restype = None
        L.lp_oracle_gauss_jordan.argtypes = [ctypes.c_int, _dp, _dp, _dp, ctypes.c_int]
        L.lp_oracle_gauss_jordan_batch.restype = ctypes.c_int
        L.lp_oracle_gauss_jordan_batch.argtypes = [ctypes.c_int, ctypes.c_int, _dp, _dp, _dp,
                                                   ctypes.c_int, ctypes.c_int]
        L.lp_oracle_gauss_jordan_pivot.restype = None
        L.lp_oracle_gauss_jordan_pivot.argtypes = [ctypes.c_int, _dp, _dp, _dp, ctypes.c_int, ctypes.c_int]
        L.lp_oracle_gauss_jordan_batch_pivot.restype = ctypes.c_int
        L.lp_oracle_gauss_jordan_batch_pivot.argtypes = [ctypes.c_int, ctypes.c_int, _dp, _dp, _dp,
                                                         ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.lp_oracle_verify_bounds.restype = ctypes.c_int
        L.lp_oracle_verify_bounds.argtypes = [ctypes.c_int, ctypes.c_int, _dp, _dp, _dp, ctypes.c_double]
        L.lp_oracle_pyset_replay.restype = ctypes.c_int
        L.lp_oracle_pyset_replay.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_int, _ip, _ip, _ip]
        _lib = L
    return _lib


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _p(a, t=_dp):
    return a.ctypes.data_as(t)


RULES = {"reference": 0, "bland": 1, "dantzig": 2, "devex": 3}


def simplex(A, b, c, tol: float = 1e-9, max_pivots: int = 1_000_000, log_cap: int | None = None,
            rule: str = "reference"):
    """One LP through the C restatement.  Returns dict(status, x, objective, npivots, pivots).
    rule="bland" is the corrected (non-reference) mode, see simplex_oracle.c."""
    A, b, c = _d(A), _d(b), _d(c)
    n, m = b.shape[0], c.shape[0]
    assert A.shape == (n, m), (A.shape, n, m)
    if log_cap is None:
        log_cap = min(max_pivots, 1 << 16)
    x = np.zeros(m)
    obj = np.zeros(1)
    npiv = np.zeros(1, dtype=np.int32)
    log = np.zeros((max(log_cap, 1), 2), dtype=np.int32)
    st = lib().lp_oracle_simplex_rule(n, m, _p(A), _p(b), _p(c), tol, max_pivots, RULES[rule], _p(x), _p(obj),
                                      _p(npiv, _ip), _p(log, _ip), log_cap)
    k = int(npiv[0])
    return {"status": int(st), "x": x, "objective": float(obj[0]), "npivots": k,
            "pivots": [tuple(int(v) for v in r) for r in log[:min(k, log_cap)]]}


def simplex_state(A, b, c, tol: float = 1e-9, max_pivots: int = 1_000_000, log_cap: int | None = None,
                  rule: str = "reference", threads: int = 0):
    """One LP with the rows of every pivot swept on `threads` host threads (0 = all cores; the bits do not
    depend on the count) and the tableau it stopped on exported by variable id.  Returns the dict of
    simplex() plus obj_by_var[n+m+1], obj_const and const_by_var[n+m+1] (see simplex_oracle.c).  With
    max_pivots = K: the state after the first K pivots -- how a config-5-size LP is checked."""
    A, b, c = _d(A), _d(b), _d(c)
    n, m = b.shape[0], c.shape[0]
    assert A.shape == (n, m), (A.shape, n, m)
    if log_cap is None:
        log_cap = min(max_pivots, 1 << 16)
    nv = n + m + 1
    x = np.zeros(m)
    obj = np.zeros(1)
    npiv = np.zeros(1, dtype=np.int32)
    log = np.zeros((max(log_cap, 1), 2), dtype=np.int32)
    ov = np.zeros(nv + 1)
    cv = np.zeros(nv)
    st = lib().lp_oracle_simplex_state(n, m, _p(A), _p(b), _p(c), tol, max_pivots, RULES[rule], threads, _p(x),
                                       _p(obj), _p(npiv, _ip), _p(log, _ip), log_cap, _p(ov), _p(cv))
    k = int(npiv[0])
    return {"status": int(st), "x": x, "objective": float(obj[0]), "npivots": k,
            "pivots": [tuple(int(v) for v in r) for r in log[:min(k, log_cap)]],
            "obj_by_var": ov[:nv], "obj_const": float(ov[nv]), "const_by_var": cv}


def verify_bounds(A, b, x, tolerance: float = 1e-3) -> int:
    """First row on which the reference's Simplex.verify_bounds (simplex.py:663-671) would assert, else -1."""
    A, b, x = _d(A), _d(b), _d(x)
    n, m = b.shape[0], x.shape[0]
    assert A.shape == (n, m)
    return int(lib().lp_oracle_verify_bounds(n, m, _p(A), _p(b), _p(x), tolerance))


def driveouts(reset: bool = True) -> int:
    """x_aux drive-out pivots the rule="bland" restatement performed since the last reset."""
    return int(lib().lp_oracle_driveouts(int(reset)))


def simplex_batch(A, b, c, tol: float = 1e-9, max_pivots: int = 1_000_000, threads: int = 0,
                  rule: str = "reference"):
    """Batch [B,n,m],[B,n],[B,m] on host threads.  Returns (status, x, objective, npivots, threads)."""
    A, b, c = _d(A), _d(b), _d(c)
    B, n, m = A.shape
    status = np.zeros(B, dtype=np.int32)
    x = np.zeros((B, m))
    obj = np.zeros(B)
    npiv = np.zeros(B, dtype=np.int32)
    used = lib().lp_oracle_simplex_batch_rule(B, n, m, _p(A), _p(b), _p(c), tol, max_pivots, RULES[rule],
                                              _p(status, _ip), _p(x), _p(obj), _p(npiv, _ip), threads)
    return status, x, obj, npiv, int(used)


PIVOTING = {"reference": 0, "partial": 1}


def gauss_jordan(a, b, full_sweep: bool = True, pivoting: str = "reference"):
    a, b = _d(a), _d(b)
    n = b.shape[0]
    x = np.zeros(n)
    if n:
        lib().lp_oracle_gauss_jordan_pivot(n, _p(a), _p(b), _p(x), int(full_sweep), PIVOTING[pivoting])
    return x


def gauss_jordan_batch(a, b, full_sweep: bool = False, threads: int = 0, pivoting: str = "reference"):
    a, b = _d(a), _d(b)
    B, n = b.shape
    x = np.zeros((B, n))
    used = lib().lp_oracle_gauss_jordan_batch_pivot(B, n, _p(a), _p(b), _p(x), int(full_sweep),
                                                    PIVOTING[pivoting], threads)
    return x, int(used)


def pyset_replay(lo: int, hi: int, ops):
    """Replay [(remove, add), ...] on set(range(lo, hi)); returns (iteration order, table size)."""
    ops = np.ascontiguousarray(ops, dtype=np.int32).reshape(-1, 2)
    cap = (hi - lo) + len(ops) + 1
    order = np.zeros(cap, dtype=np.int32)
    size = np.zeros(1, dtype=np.int32)
    cnt = lib().lp_oracle_pyset_replay(lo, hi, len(ops), _p(ops, _ip), _p(order, _ip), _p(size, _ip))
    if cnt < 0:
        raise KeyError("remove of a key that is not in the set")
    return [int(v) for v in order[:cnt]], int(size[0])


# ---- the unmodified reference Gauss-Jordan (oracle/_ref), when it was built here ----
def ref_gauss_jordan_available() -> bool:
    return os.path.exists(REF_GJ_LIB_PATH)


def ref_gauss_jordan(a, b):
    """SolveEquation of the reference itself (energy_values.cc:166), called in-process."""
    global _ref_gj
    if _ref_gj is None:
        R = ctypes.CDLL(REF_GJ_LIB_PATH)
        R.lp_ref_gauss_jordan.restype = None
        R.lp_ref_gauss_jordan.argtypes = [ctypes.c_int, _dp, _dp, _dp]
        _ref_gj = R
    a, b = _d(a), _d(b)
    n = b.shape[0]
    x = np.zeros(n)
    if n:
        _ref_gj.lp_ref_gauss_jordan(n, _p(a), _p(b), _p(x))
    return x
