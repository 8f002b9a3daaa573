"""Host-side mirror of the reference's Gauss-Jordan interface on the B200 engine.

Reference surface (file:line in /root/reference/week-2/energy_values/energy_values.cc):
  * ``Column SolveEquation(Equation{a, b})``   :166-197   -> ``solve_equation(a, b)``
  * ``ReadEquation()``                         :66-80     -> ``read_equation(text)``
  * ``PrintColumn(column)``                    :199-205   -> ``format_column(x)``
plus the batch entry point the engine adds (B independent systems, one CTA each).
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib


def _is_torch_cuda(t) -> bool:
    return type(t).__module__.startswith("torch") and getattr(t, "is_cuda", False)


PIVOT_REFERENCE, PIVOT_PARTIAL = 0, 1
_PIVOTING = {"reference": PIVOT_REFERENCE, "partial": PIVOT_PARTIAL}


def solve_equation_batch(a, b, pivoting: str = "reference", return_status: bool = False):
    """a[B,n,n], b[B,n] -> x[B,n].  numpy in -> numpy out (host path); CUDA float64 torch
    tensors in -> CUDA tensor out (zero-copy, current stream).  Inputs are not modified.
    pivoting: "reference" = the reference's rule (diagonal pivot, swap on an exact zero only,
    energy_values.cc:85-114); "partial" = largest |a[i][k]| in the column (not reference
    behaviour; for ill-conditioned systems).
    return_status=True returns (x, status): status[B] int32 is -1 when every step found a non-zero pivot, else
    the first step at which none existed (the reference then divides by zero: that system's x is inf/nan)."""
    if pivoting not in _PIVOTING:
        raise ValueError("pivoting must be 'reference' or 'partial'")
    pv = _PIVOTING[pivoting]
    L = _lib.lib()
    if _is_torch_cuda(a):
        import torch
        if not (a.dtype == torch.float64 and b.dtype == torch.float64 and a.is_contiguous()
                and b.is_contiguous() and _is_torch_cuda(b) and a.device == b.device):
            raise ValueError("CUDA inputs must be contiguous float64 tensors on one device")
        B, n, n2 = a.shape
        if n != n2 or tuple(b.shape) != (B, n):
            raise ValueError("expected a[B,n,n], b[B,n]")
        with torch.cuda.device(a.device):
            x = torch.empty_like(b)
            wsb = int(L.lpb200_gauss_jordan_workspace_bytes(B, n))
            ws = torch.empty(max(wsb, 1), dtype=torch.uint8, device=a.device)
            stream = torch.cuda.current_stream(a.device).cuda_stream
            status = torch.empty(B, dtype=torch.int32, device=a.device) if return_status else None
            _lib.check(L.lpb200_gauss_jordan_batch_status_device(
                B, n, a.data_ptr(), b.data_ptr(), x.data_ptr(), pv, status.data_ptr() if return_status else None,
                ws.data_ptr(), wsb, stream))
        return (x, status) if return_status else x
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    if a.ndim != 3 or b.ndim != 2 or a.shape[1] != a.shape[2] or b.shape != a.shape[:2]:
        raise ValueError("expected a[B,n,n], b[B,n]")
    B, n = b.shape
    x = np.zeros((B, n), dtype=np.float64)
    status = np.full(B, -1, dtype=np.int32)
    _lib.check(L.lpb200_gauss_jordan_batch_status(B, n, ctypes.c_void_p(a.ctypes.data), ctypes.c_void_p(b.ctypes.data),
                                                  ctypes.c_void_p(x.ctypes.data), pv,
                                                  ctypes.c_void_p(status.ctypes.data) if return_status else None))
    return (x, status) if return_status else x


def solve_equation(a, b, pivoting: str = "reference"):
    """SolveEquation (energy_values.cc:166): one square system, returns the solution column."""
    b = np.asarray(b, dtype=np.float64).reshape(-1)
    n = b.shape[0]
    if n == 0:
        return np.zeros(0, dtype=np.float64)
    a = np.asarray(a, dtype=np.float64).reshape(n, n)
    return solve_equation_batch(a[None], b[None], pivoting=pivoting)[0]


def read_equation(text: str):
    """ReadEquation (energy_values.cc:66-80): "size" then size rows of size+1 numbers (a | b)."""
    toks = text.split()
    size = int(toks[0])
    vals = np.array([float(t) for t in toks[1:1 + size * (size + 1)]], dtype=np.float64).reshape(size, size + 1)
    return np.ascontiguousarray(vals[:, :size]), np.ascontiguousarray(vals[:, size])


def format_column(x) -> str:
    """PrintColumn (energy_values.cc:199-205): cout.precision(20), one value per line
    (ostream default float format == printf %.20g)."""
    return "".join("%.20g\n" % float(v) for v in x)
