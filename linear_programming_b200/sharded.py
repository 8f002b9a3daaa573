"""Multi-GPU host logic: one process per GPU, torch.distributed for plumbing.

* Batched problems shard by problem index with NO data-path collective: ``problem_shard``.
* One large LP is column-sharded: rank g owns tableau column positions [g*W, (g+1)*W).  The
  per-pivot exchange (8-byte entering-key all-to-all + pivot-column broadcast) happens INSIDE
  the persistent CUDA kernel over peer-mapped exchange buffers (simplex_single.cu); the host
  only (1) creates the communicator, (2) all-gathers the 64-byte IPC handles with
  torch.distributed (any backend: nccl on the GPUs, gloo in the CPU tests) and (3) launches.

The reference has no distributed path; this replaces nothing in it -- it is the engine's
scale-out of ``Simplex.solve()`` (week-2/diet/simplex.py:586-630) for one LP too large or too
slow for a single GPU.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from .simplex import DEFAULT_TOLERANCE, BatchResult, _rule_code, default_max_pivots


def problem_shard(total: int, world: int, rank: int):
    """Contiguous block of problem indices [lo, hi) owned by ``rank`` (strong scaling: the
    first ``total % world`` ranks get one extra problem)."""
    if world < 1 or not (0 <= rank < world) or total < 0:
        raise ValueError("need world >= 1, 0 <= rank < world, total >= 0")
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def column_shard(m: int, world: int, rank: int):
    """(c_lo, c_count): the columns of A owned by ``rank`` when an n x m LP is column-sharded
    over ``world`` GPUs (positions are dealt in blocks of W = ceil16(ceil((m+1)/world)))."""
    lo = ctypes.c_int32()
    cnt = ctypes.c_int32()
    _lib.check(_lib.lib().lpb200_simplex_sharded_columns(m, world, rank, ctypes.byref(lo), ctypes.byref(cnt)))
    return int(lo.value), int(cnt.value)


def gather_handles(handle: bytes, group=None) -> bytes:
    """All-gather one fixed-size opaque handle per rank, concatenated in rank order."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    mine = torch.frombuffer(bytearray(handle), dtype=torch.uint8).to(dev)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return b"".join(bytes(t.cpu().numpy().tobytes()) for t in out)


class ShardedComm:
    """Peer-mapped exchange buffers of the column-sharded solver (one per process/GPU)."""

    def __init__(self, max_n: int, group=None):
        import torch.distributed as dist
        L = _lib.lib()
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.max_n = int(max_n)
        hb = int(L.lpb200_comm_handle_bytes())
        buf = ctypes.create_string_buffer(hb)
        self._comm = ctypes.c_void_p()
        _lib.check(L.lpb200_comm_create(self.rank, self.world, self.max_n, ctypes.byref(self._comm), buf))
        handles = gather_handles(buf.raw, group)
        _lib.check(L.lpb200_comm_connect(self._comm, handles))
        dist.barrier(group)  # every rank has mapped every buffer before anyone launches

    def close(self):
        if self._comm:
            _lib.lib().lpb200_comm_destroy(self._comm)
            self._comm = ctypes.c_void_p()

    def solve(self, A_cols, b, c, n: int, m: int, tol: float = DEFAULT_TOLERANCE, max_pivots=None,
              log_cap: int = 0, timeout_s: float = 20.0, workspace=None, rule: str = "reference") -> BatchResult:
        """Collective: every rank calls it with ITS columns of A (CUDA float64 tensor [n, c_count],
        any row stride) and the full b, c.  Enqueues on the current stream; results (replicated
        on every rank) stay on the GPU.  rule: as in solve_batch (every rank must pass the same)."""
        import torch
        L = _lib.lib()
        if max_pivots is None:
            max_pivots = default_max_pivots(n, m)
        dev = b.device
        lo, cnt = column_shard(m, self.world, self.rank)
        if cnt > 0:
            if tuple(A_cols.shape) != (n, cnt) or A_cols.dtype != torch.float64 or A_cols.stride(1) != 1:
                raise ValueError("A_cols must be float64 [n, %d] with unit column stride" % cnt)
            a_ptr, lda = A_cols.data_ptr(), A_cols.stride(0)
        else:
            a_ptr, lda = None, 0
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev).cuda_stream
            wsb = int(L.lpb200_simplex_sharded_workspace_bytes(n, m, self.world))
            ws = workspace if workspace is not None else torch.empty(wsb, dtype=torch.uint8, device=dev)
            status = torch.empty(1, dtype=torch.int32, device=dev)
            x = torch.empty((1, m), dtype=torch.float64, device=dev)
            obj = torch.empty(1, dtype=torch.float64, device=dev)
            npiv = torch.empty(1, dtype=torch.int32, device=dev)
            log = torch.full((1, log_cap, 2), -1, dtype=torch.int32, device=dev) if log_cap > 0 else None
            _lib.check(L.lpb200_simplex_sharded_ex_device(
                self._comm, n, m, a_ptr, lda, b.data_ptr(), c.data_ptr(), tol, max_pivots, _rule_code(rule),
                status.data_ptr(),
                x.data_ptr(), obj.data_ptr(), npiv.data_ptr(), log.data_ptr() if log is not None else None,
                log_cap, ws.data_ptr(), wsb, timeout_s, stream))
        res = BatchResult(status, x, obj, npiv, log)
        res._keepalive = ws
        return res


def solve_sharded_emulated(A, b, c, nranks: int, tol: float = DEFAULT_TOLERANCE, max_pivots=None,
                           log_cap: int = 0, rule: str = "reference", state: bool = False):
    """The nranks-way column-sharded protocol with every rank emulated as a CTA group of one
    cooperative launch on the current GPU.  Returns one BatchResult whose leading axis is the
    RANK (all ranks must agree); used by the parity tests and on boxes with fewer GPUs.
    state=True also exports the tableau the solve stopped on (lpb200_simplex_sharded_emulated_state):
    ``res.obj_by_var[rank, n+m+2]`` (each rank reports the columns it owns) and
    ``res.const_by_var[rank, n+m+1]`` (replicated)."""
    L = _lib.lib()
    A = np.ascontiguousarray(A, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64).reshape(-1)
    c = np.ascontiguousarray(c, dtype=np.float64).reshape(-1)
    n, m = b.shape[0], c.shape[0]
    if A.shape != (n, m):
        raise ValueError("expected A[n,m], b[n], c[m]")
    if max_pivots is None:
        max_pivots = default_max_pivots(n, m)
    G = int(nranks)
    status = np.zeros(G, dtype=np.int32)
    x = np.zeros((G, m), dtype=np.float64)
    obj = np.zeros(G, dtype=np.float64)
    npiv = np.zeros(G, dtype=np.int32)
    log = np.full((G, log_cap, 2), -1, dtype=np.int32) if log_cap > 0 else None

    def vp(a):
        return ctypes.c_void_p(a.ctypes.data)
    if state:
        nv = n + m + 1
        ov = np.zeros((G, nv + 1), dtype=np.float64)
        cv = np.zeros((G, nv), dtype=np.float64)
        _lib.check(L.lpb200_simplex_sharded_emulated_state(G, n, m, vp(A), vp(b), vp(c), tol, max_pivots,
                                                           _rule_code(rule), vp(status), vp(x), vp(obj), vp(npiv),
                                                           vp(log) if log is not None else None, log_cap, vp(ov),
                                                           vp(cv)))
        res = BatchResult(status, x, obj, npiv, log)
        res.obj_by_var, res.const_by_var = ov, cv
        return res
    _lib.check(L.lpb200_simplex_sharded_emulated_ex(G, n, m, vp(A), vp(b), vp(c), tol, max_pivots, _rule_code(rule),
                                                    vp(status), vp(x), vp(obj), vp(npiv),
                                                    vp(log) if log is not None else None, log_cap))
    return BatchResult(status, x, obj, npiv, log)
