"""ctypes binding of liblp_b200.so -- the only route from Python to the engine.

There is NO CPU fallback: if the library is missing this raises, and if no CUDA device is
visible every compute entry point returns LPB200_E_NO_DEVICE which is raised as EngineError.
The symbols bound here are exactly the ones include/lp_b200.h declares.
"""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LPB_LIB_PATH") or os.path.join(HERE, "liblp_b200.so")  # override: kernel experiments

c_dp = ctypes.POINTER(ctypes.c_double)
c_ip = ctypes.POINTER(ctypes.c_int32)
c_vp = ctypes.c_void_p

# name -> (restype, argtypes); keep in sync with include/lp_b200.h (tests/test_abi.py checks)
SIGNATURES = {
    "lpb200_version": (ctypes.c_char_p, []),
    "lpb200_last_error": (ctypes.c_char_p, []),
    "lpb200_device_count": (ctypes.c_int, []),
    "lpb200_simplex_batch_smem_bytes": (ctypes.c_int64, [ctypes.c_int, ctypes.c_int]),
    "lpb200_simplex_batch_device": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp,
                                                   ctypes.c_double, ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp,
                                                   ctypes.c_int, c_vp]),
    "lpb200_simplex_batch": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp,
                                            ctypes.c_double, ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp,
                                            ctypes.c_int]),
    "lpb200_simplex_batch_ex_device": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_vp,
                                                      ctypes.c_int, c_vp, c_vp, ctypes.c_double, ctypes.c_int,
                                                      ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp, ctypes.c_int,
                                                      c_vp]),
    "lpb200_simplex_batch_ex": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_vp, ctypes.c_int,
                                               c_vp, c_vp, ctypes.c_double, ctypes.c_int, ctypes.c_int, c_vp,
                                               c_vp, c_vp, c_vp, c_vp, ctypes.c_int]),
    "lpb200_verify_bounds_batch_device": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp,
                                                         c_vp, ctypes.c_double, c_vp, c_vp]),
    "lpb200_verify_bounds_batch": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, c_vp,
                                                  ctypes.c_double, c_vp]),
    "lpb200_narrow_f64_to_i8": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int64]),
    "lpb200_host_threads": (ctypes.c_int, []),
    "lpb200_simplex_batch_last_lanes": (ctypes.c_int, [c_vp]),
    "lpb200_simplex_single_workspace_bytes": (ctypes.c_int64, [ctypes.c_int, ctypes.c_int]),
    "lpb200_simplex_single_state_device": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, c_vp]),
    "lpb200_simplex_single_device": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, ctypes.c_double,
                                                    ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp, ctypes.c_int,
                                                    c_vp, ctypes.c_int64, c_vp]),
    "lpb200_simplex_single": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, ctypes.c_double,
                                             ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp, ctypes.c_int, c_vp]),
    "lpb200_simplex_single_ex_device": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, ctypes.c_double,
                                                       ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp,
                                                       ctypes.c_int, c_vp, ctypes.c_int64, c_vp]),
    "lpb200_simplex_single_ex": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, ctypes.c_double,
                                                ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp,
                                                ctypes.c_int, c_vp]),
    "lpb200_comm_handle_bytes": (ctypes.c_int64, []),
    "lpb200_comm_create": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(c_vp), c_vp]),
    "lpb200_comm_connect": (ctypes.c_int, [c_vp, c_vp]),
    "lpb200_comm_destroy": (ctypes.c_int, [c_vp]),
    "lpb200_simplex_sharded_columns": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_ip, c_ip]),
    "lpb200_simplex_sharded_workspace_bytes": (ctypes.c_int64, [ctypes.c_int, ctypes.c_int, ctypes.c_int]),
    "lpb200_simplex_sharded_device": (ctypes.c_int, [c_vp, ctypes.c_int, ctypes.c_int, c_vp, ctypes.c_int64, c_vp,
                                                     c_vp, ctypes.c_double, ctypes.c_int, c_vp, c_vp, c_vp, c_vp,
                                                     c_vp, ctypes.c_int, c_vp, ctypes.c_int64, ctypes.c_double,
                                                     c_vp]),
    "lpb200_simplex_sharded_emulated": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp,
                                                       ctypes.c_double, ctypes.c_int, c_vp, c_vp, c_vp, c_vp,
                                                       c_vp, ctypes.c_int]),
    "lpb200_simplex_sharded_ex_device": (ctypes.c_int, [c_vp, ctypes.c_int, ctypes.c_int, c_vp, ctypes.c_int64,
                                                        c_vp, c_vp, ctypes.c_double, ctypes.c_int, ctypes.c_int,
                                                        c_vp, c_vp, c_vp, c_vp, c_vp, ctypes.c_int, c_vp,
                                                        ctypes.c_int64, ctypes.c_double, c_vp]),
    "lpb200_simplex_sharded_emulated_ex": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_vp, c_vp,
                                                          c_vp, ctypes.c_double, ctypes.c_int, ctypes.c_int, c_vp,
                                                          c_vp, c_vp, c_vp, c_vp, ctypes.c_int]),
    "lpb200_simplex_sharded_emulated_state": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, c_vp, c_vp,
                                                             c_vp, ctypes.c_double, ctypes.c_int, ctypes.c_int, c_vp,
                                                             c_vp, c_vp, c_vp, c_vp, ctypes.c_int, c_vp, c_vp]),
    "lpb200_gauss_jordan_workspace_bytes": (ctypes.c_int64, [ctypes.c_int, ctypes.c_int]),
    "lpb200_gauss_jordan_batch_device": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, c_vp,
                                                        ctypes.c_int64, c_vp]),
    "lpb200_gauss_jordan_batch": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp]),
    "lpb200_gauss_jordan_batch_ex_device": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, ctypes.c_int,
                                                           c_vp, ctypes.c_int64, c_vp]),
    "lpb200_gauss_jordan_batch_ex": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, ctypes.c_int]),
    "lpb200_gauss_jordan_batch_status_device": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp,
                                                               ctypes.c_int, c_vp, c_vp, ctypes.c_int64, c_vp]),
    "lpb200_gauss_jordan_batch_static_device": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp,
                                                               ctypes.c_int, c_vp, c_vp]),
    "lpb200_gauss_jordan_batch_status": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, ctypes.c_int,
                                                        c_vp]),
}


class EngineError(RuntimeError):
    """A liblp_b200 call failed (bad argument, CUDA error, or no CUDA device)."""

    def __init__(self, code: int, message: str):
        super().__init__("liblp_b200 error %d: %s" % (code, message))
        self.code = code


_lib = None


def lib():
    """Load liblp_b200.so (built in-tree by linear_programming_b200.build).  Raises if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "liblp_b200.so is not built (%s). Run `python -m linear_programming_b200.build` "
                "(needs nvcc). The engine has no CPU fallback." % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise EngineError(rc, lib().lpb200_last_error().decode("utf-8", "replace"))


def device_count() -> int:
    return int(lib().lpb200_device_count())
