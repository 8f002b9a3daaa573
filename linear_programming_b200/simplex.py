"""Host-side mirror of the reference's LP interface, running on the B200 engine.

Reference surface mirrored (file:line in /root/reference/week-2/diet/simplex.py):
  * ``Simplex(A, b, c, n, m)``            :496-505   same attribute names
  * ``Simplex.solve() -> (anst, ansx)``   :586-630   anst in {-1, 0, 1}; ``.optimum``,
                                                     ``.assignment``, ``.anst`` side attributes
  * ``Simplex.answer_type_str(anst)``     :507-512
  * ``Simplex.parse(stream, file)``       :707-718   "n m" / n rows of A / b / c
  * error behaviour: inputs on which the reference dies with AttributeError (:626),
    IndexError (:541) or TypeError (:409) raise the same exception type here.
and the C++ stubs' call shape ``allocate_ads(n, m, A, b, c)`` (ad_allocation.cpp:9-14) /
``solve_diet_problem(n, m, A, b, c)`` (diet.cc:10).

Everything numeric happens in liblp_b200.so (hand-written sm_100a CUDA) through ctypes;
PyTorch is used only to hand CUDA tensors across (data_ptr + current stream).
"""
from __future__ import annotations

import ctypes
import sys
from dataclasses import dataclass

import numpy as np

from . import _lib

STATUS_BOUNDED = 0
STATUS_NO_SOLUTION = -1
STATUS_INFINITY = 1
STATUS_ITERATION_LIMIT = 2
STATUS_REF_ATTRIBUTE_ERROR = 3
STATUS_REF_INDEX_ERROR = 4
STATUS_REF_TYPE_ERROR = 5

DEFAULT_TOLERANCE = 1e-09  # global_tolerance, simplex.py:16
SMEM_LIMIT_BYTES = 227 * 1024


class IterationLimitError(RuntimeError):
    """max_pivots reached (the reference has no cap: it would keep cycling)."""


@dataclass
class BatchResult:
    status: object      # int32 [B]
    x: object           # float64 [B, m]  (zeros where status != 0)
    objective: object   # float64 [B]
    npivots: object     # int32 [B]
    pivot_log: object   # int32 [B, log_cap, 2] or None; rows beyond npivots are -1

    def pivots(self, p: int):
        """(entering, leaving) pairs of problem p, like the instrumented reference records."""
        if self.pivot_log is None:
            raise ValueError("solve_batch was called with log_cap=0")
        k = int(self.npivots[p])
        log = self.pivot_log[p]
        if hasattr(log, "cpu"):
            log = log.cpu().numpy()
        return [(int(e), int(l)) for e, l in log[:min(k, log.shape[0])]]


def default_max_pivots(n: int, m: int) -> int:
    return max(10000, 50 * (n + m))


def fits_batch_kernel(n: int, m: int) -> bool:
    return int(_lib.lib().lpb200_simplex_batch_smem_bytes(n, m)) <= SMEM_LIMIT_BYTES


def _is_torch_cuda(t) -> bool:
    return type(t).__module__.startswith("torch") and getattr(t, "is_cuda", False)


def _vp(arr) -> ctypes.c_void_p:
    return ctypes.c_void_p(arr.ctypes.data)


# Element types the batched kernel takes for A (include/lp_b200.h LPB200_DTYPE_*).  The reference
# parses integers (simplex.py:712-717); an integer or float32 A is widened to fp64 on the GPU,
# exactly, so the answers are those of the float64 call at 1/8 .. 1/2 of the input bytes.
DTYPE_F64, DTYPE_F32, DTYPE_I32, DTYPE_I16, DTYPE_I8 = 0, 1, 2, 3, 4
_TYPED_A = {"float64": DTYPE_F64, "float32": DTYPE_F32, "int32": DTYPE_I32, "int16": DTYPE_I16, "int8": DTYPE_I8}


def _a_dtype_code(dtype) -> int | None:
    """LPB200_DTYPE_* of a numpy / torch dtype, None when A has to be converted to float64."""
    return _TYPED_A.get(str(dtype).replace("torch.", ""))


RULE_REFERENCE, RULE_BLAND, RULE_DANTZIG, RULE_DEVEX = 0, 1, 2, 3
_RULES = {"reference": RULE_REFERENCE, "bland": RULE_BLAND, "dantzig": RULE_DANTZIG, "devex": RULE_DEVEX}


def _rule_code(rule: str) -> int:
    if rule not in _RULES:
        raise ValueError("rule must be 'reference', 'bland', 'dantzig' or 'devex'")
    return _RULES[rule]


def solve_batch(A, b, c, tol: float = DEFAULT_TOLERANCE, max_pivots: int | None = None,
                log_cap: int = 0, mode: str = "auto", rule: str = "reference") -> BatchResult:
    """Solve B independent LPs  max c.x  s.t. A x <= b, x >= 0  with the reference's semantics.

    A[B,n,m], b[B,n], c[B,m]: numpy / nested lists (host path: copies in and out) or CUDA
    float64 torch tensors (zero-copy, enqueued on the current stream, results stay on the GPU).
    A may also be int8 / int16 / int32 / float32 (numpy array or CUDA tensor): the batch kernel
    reads it as is and widens on the GPU -- same answers, a fraction of the bytes over PCIe.
    rule: "reference" reproduces the reference bit for bit, bugs included; "bland" is the
    corrected mode (include/lp_b200.h LPB200_RULE_BLAND: Bland's rule, zero ratios kept, x_aux
    driven out, no crash classes) and "dantzig" the same with largest-coefficient pricing (far
    fewer pivots; Bland takes over on degenerate streaks) -- not reference behaviour; both kernels
    serve every rule with identical decisions.  "devex" prices by Devex reference-framework weights
    (approximate steepest edge, about a third fewer pivots than "dantzig").
    mode: "batch" (one CTA per LP, tableau in shared memory), "single" (grid-wide kernel, one
    LP after the other), "auto" (batch when the tableau fits shared memory).
    """
    L = _lib.lib()
    if _is_torch_cuda(A):
        return _solve_batch_torch(L, A, b, c, tol, max_pivots, log_cap, mode, rule)
    a_code = _a_dtype_code(getattr(A, "dtype", None))
    A = np.ascontiguousarray(A) if a_code is not None else np.ascontiguousarray(A, dtype=np.float64)
    a_code = a_code or DTYPE_F64
    b = np.ascontiguousarray(b, dtype=np.float64)
    c = np.ascontiguousarray(c, dtype=np.float64)
    if A.ndim != 3 or b.ndim != 2 or c.ndim != 2:
        raise ValueError("expected A[B,n,m], b[B,n], c[B,m]")
    B, n, m = A.shape
    if b.shape != (B, n) or c.shape != (B, m):
        raise ValueError("shape mismatch: A%s b%s c%s" % (A.shape, b.shape, c.shape))
    if max_pivots is None:
        max_pivots = default_max_pivots(n, m)
    status = np.zeros(B, dtype=np.int32)
    x = np.zeros((B, m), dtype=np.float64)
    obj = np.zeros(B, dtype=np.float64)
    npiv = np.zeros(B, dtype=np.int32)
    log = np.full((B, log_cap, 2), -1, dtype=np.int32) if log_cap > 0 else None
    if mode == "auto":
        mode = "batch" if fits_batch_kernel(n, m) else "single"
    r_code = _rule_code(rule)
    if mode == "batch":
        _lib.check(L.lpb200_simplex_batch_ex(B, n, m, _vp(A), a_code, _vp(b), _vp(c), tol, max_pivots, r_code,
                                             _vp(status), _vp(x), _vp(obj), _vp(npiv),
                                             _vp(log) if log is not None else None, log_cap))
    elif mode == "single":
        if a_code != DTYPE_F64:
            A = A.astype(np.float64)
        for p in range(B):
            lp = log[p] if log is not None else None
            _lib.check(L.lpb200_simplex_single_ex(n, m, _vp(A[p]), _vp(b[p]), _vp(c[p]), tol, max_pivots, r_code,
                                                  _vp(status[p:p + 1]), _vp(x[p]), _vp(obj[p:p + 1]),
                                                  _vp(npiv[p:p + 1]), _vp(lp) if lp is not None else None,
                                                  log_cap, None))
    else:
        raise ValueError("mode must be 'auto', 'batch' or 'single'")
    return BatchResult(status, x, obj, npiv, log)


def _solve_batch_torch(L, A, b, c, tol, max_pivots, log_cap, mode, rule="reference"):
    import torch
    a_code = _a_dtype_code(A.dtype)
    if a_code is None or not A.is_contiguous():
        raise ValueError("CUDA A must be a contiguous float64/float32/int32/int16/int8 tensor")
    for t in (b, c):
        if not (_is_torch_cuda(t) and t.dtype == torch.float64 and t.is_contiguous() and t.device == A.device):
            raise ValueError("CUDA b and c must be contiguous float64 tensors on the device of A")
    B, n, m = A.shape
    if tuple(b.shape) != (B, n) or tuple(c.shape) != (B, m):
        raise ValueError("shape mismatch")
    if max_pivots is None:
        max_pivots = default_max_pivots(n, m)
    dev = A.device
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        status = torch.empty(B, dtype=torch.int32, device=dev)
        x = torch.empty((B, m), dtype=torch.float64, device=dev)
        obj = torch.empty(B, dtype=torch.float64, device=dev)
        npiv = torch.empty(B, dtype=torch.int32, device=dev)
        log = torch.full((B, log_cap, 2), -1, dtype=torch.int32, device=dev) if log_cap > 0 else None
        logp = log.data_ptr() if log is not None else None
        if mode == "auto":
            mode = "batch" if fits_batch_kernel(n, m) else "single"
        r_code = _rule_code(rule)
        if mode == "batch":
            _lib.check(L.lpb200_simplex_batch_ex_device(B, n, m, A.data_ptr(), a_code, b.data_ptr(),
                                                        c.data_ptr(), tol, max_pivots, r_code,
                                                        status.data_ptr(), x.data_ptr(), obj.data_ptr(),
                                                        npiv.data_ptr(), logp, log_cap, stream))
        elif mode == "single":
            if a_code != DTYPE_F64:
                A = A.to(torch.float64)
            wsb = int(L.lpb200_simplex_single_workspace_bytes(n, m))
            ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
            for p in range(B):
                _lib.check(L.lpb200_simplex_single_ex_device(
                    n, m, A[p].data_ptr(), b[p].data_ptr(), c[p].data_ptr(), tol, max_pivots, r_code,
                    status[p:].data_ptr(), x[p].data_ptr(), obj[p:].data_ptr(), npiv[p:].data_ptr(),
                    (log[p].data_ptr() if log is not None else None), log_cap, ws.data_ptr(), wsb, stream))
        else:
            raise ValueError("mode must be 'auto', 'batch' or 'single'")
    return BatchResult(status, x, obj, npiv, log)


def verify_bounds_batch(A, b, x, status=None, tolerance: float = 1e-3):
    """Batched ``Simplex.verify_bounds`` (simplex.py:663-671): for every LP p with status[p] == 0 (all, when
    status is None) the first constraint r with  sum_i A[r][i]*x[i] > b[r] + tolerance  (summed left to right like
    the reference), -1 when the answer passes.  numpy in -> numpy int32[B]; CUDA float64 tensors in -> CUDA
    int32 tensor (enqueued on the current stream).  The reference hard-codes 1e-3 and ignores its argument."""
    L = _lib.lib()
    if _is_torch_cuda(A):
        import torch
        B, n, m = A.shape
        for t, shape in ((A, (B, n, m)), (b, (B, n)), (x, (B, m))):
            if not (_is_torch_cuda(t) and t.dtype == torch.float64 and t.is_contiguous() and tuple(t.shape) == shape):
                raise ValueError("CUDA A[B,n,m], b[B,n], x[B,m] must be contiguous float64 tensors")
        if status is not None and not (_is_torch_cuda(status) and status.dtype == torch.int32 and status.is_contiguous()):
            raise ValueError("CUDA status must be a contiguous int32 tensor")
        with torch.cuda.device(A.device):
            out = torch.empty(B, dtype=torch.int32, device=A.device)
            _lib.check(L.lpb200_verify_bounds_batch_device(
                B, n, m, A.data_ptr(), b.data_ptr(), x.data_ptr(), status.data_ptr() if status is not None else None,
                tolerance, out.data_ptr(), torch.cuda.current_stream(A.device).cuda_stream))
        return out
    A = np.ascontiguousarray(A, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    x = np.ascontiguousarray(x, dtype=np.float64)
    if A.ndim != 3 or b.shape != A.shape[:2] or x.shape != (A.shape[0], A.shape[2]):
        raise ValueError("expected A[B,n,m], b[B,n], x[B,m]")
    B, n, m = A.shape
    st = np.ascontiguousarray(status, dtype=np.int32) if status is not None else None
    out = np.full(B, -1, dtype=np.int32)
    _lib.check(L.lpb200_verify_bounds_batch(B, n, m, _vp(A), _vp(b), _vp(x), _vp(st) if st is not None else None,
                                            tolerance, _vp(out)))
    return out


def solve(A, b, c, tol: float = DEFAULT_TOLERANCE, max_pivots: int | None = None, mode: str = "auto",
          rule: str = "reference"):
    """Per-problem surface: solve(A, b, c) -> (status, x).  x is a list of m floats when
    status == 0, else None (like the reference's (anst, ansx))."""
    A = np.asarray(A, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64).reshape(-1)
    c = np.asarray(c, dtype=np.float64).reshape(-1)
    n, m = b.shape[0], c.shape[0]
    A = A.reshape(n, m)
    r = solve_batch(A[None], b[None], c[None], tol=tol, max_pivots=max_pivots, mode=mode, rule=rule)
    st = int(r.status[0])
    return st, ([float(v) for v in r.x[0]] if st == STATUS_BOUNDED else None)


class Simplex:
    """Drop-in for the reference's ``Simplex`` (simplex.py:494-718) on the B200 engine."""

    def __init__(self, A, b, c, n, m):
        self.A = A
        self.b = b
        self.c = c
        self.v = 0
        self.n = n
        self.m = m
        self.tolerance = DEFAULT_TOLERANCE
        self.max_pivots = None
        self.rule = "reference"  # "bland": the corrected mode (engine-only attribute, see solve_batch)
        self.npivots = None
        self.pivots = None

    @staticmethod
    def answer_type_str(anst):
        try:
            return ["No solution", "Bounded solution", "Infinity"][anst + 1]
        except Exception:
            return "Unrecognized Answer : %s" % anst

    def solve(self, record_pivots: bool = False):
        n, m = self.n, self.m
        A = np.asarray(self.A, dtype=np.float64).reshape(n, m)
        b = np.asarray(self.b, dtype=np.float64).reshape(n)
        c = np.asarray(self.c, dtype=np.float64).reshape(m)
        cap = self.max_pivots if self.max_pivots is not None else default_max_pivots(n, m)
        r = solve_batch(A[None], b[None], c[None], tol=self.tolerance, max_pivots=cap,
                        log_cap=(cap if record_pivots else 0), rule=self.rule)
        st = int(r.status[0])
        self.npivots = int(r.npivots[0])
        if record_pivots:
            self.pivots = r.pivots(0)
        if st == STATUS_BOUNDED:
            self.optimum = float(r.objective[0])
            self.assignment = [float(v) for v in r.x[0]]
            self.anst = 0
            return (0, self.assignment)
        if st == STATUS_INFINITY:
            self.anst = 1
            return (1, None)
        if st == STATUS_NO_SOLUTION:
            self.anst = -1
            return (-1, None)
        # the classes in which the reference itself raises
        if st == STATUS_REF_ATTRIBUTE_ERROR:
            raise AttributeError("'Simplex' object has no attribute 'assignment'")   # simplex.py:626
        if st == STATUS_REF_INDEX_ERROR:
            raise IndexError("list index out of range")                              # simplex.py:183
        if st == STATUS_REF_TYPE_ERROR:
            raise TypeError("unsupported operand type(s) for *: 'int' and 'NoneType'")  # simplex.py:409
        if st == STATUS_ITERATION_LIMIT:
            raise IterationLimitError("no optimum after %d pivots" % self.npivots)
        raise RuntimeError("unknown status %d" % st)

    def verify_bounds(self, tolerance=1e-3):
        """simplex.py:663-671: assert A x <= b + 1e-3 row by row on the last answer (the reference reads the
        module-global ``ansx`` and ignores ``tolerance``; here the answer is ``self.assignment``)."""
        n, m = self.n, self.m
        A = np.asarray(self.A, dtype=np.float64).reshape(1, n, m)
        b = np.asarray(self.b, dtype=np.float64).reshape(1, n)
        x = np.asarray(self.assignment, dtype=np.float64).reshape(1, m)
        assert int(verify_bounds_batch(A, b, x, tolerance=1e-3)[0]) < 0

    @staticmethod
    def parse(stream=sys.stdin, file=None):
        if file:
            stream = open(file, "r")
        with stream as stream:
            n, m = list(map(int, stream.readline().split()))
            A = []
            for _ in range(n):
                A += [list(map(int, stream.readline().split()))]
            b = list(map(int, stream.readline().split()))
            c = list(map(int, stream.readline().split()))
            return Simplex(A, b, c, n, m)

    def __repr__(self):
        return "A:\n%s\nb:\n%s\nc:\n%s\n" % (self.A, self.b, self.c)


def allocate_ads(n, m, A, b, c):
    """week-2/ad_allocation/ad_allocation.cpp:9-14 -- returns (status, x[m])."""
    st, x = Simplex(A, b, c, n, m).solve()
    return st, (x if x is not None else [0.0] * m)


solve_diet_problem = allocate_ads  # week-2/diet/diet.cc:10
