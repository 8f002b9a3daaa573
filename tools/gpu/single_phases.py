#!/usr/bin/env python3
"""Per-phase time of one pivot of the grid-wide kernel (needs a -DLPB_TIMING build:
python -m linear_programming_b200.build --variant timing -DLPB_TIMING, then
LPB_LIB_PATH=.../variants/liblp_b200_timing.so python tools/gpu/single_phases.py).

A SMALL LP, so the sweep moves almost nothing and what is left is the fixed cost of a pivot -- grid barriers,
selection phases, exchange -- the part the column-sharded LP pays on top of (bytes / HBM rate).  Run with the full
grid on one GPU (LPB_DEBUG_SINGLE_CPR forces the CTAs per rank) and with 8 emulated ranks."""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

import linear_programming_b200 as lp  # noqa: E402
from linear_programming_b200 import _lib, workloads  # noqa: E402
from linear_programming_b200.sharded import solve_sharded_emulated  # noqa: E402

NAMES = ["P1 entering scan", "grid.sync", "reduce + key exchange", "P2 column + ratio", "grid.sync", "reduce",
         "P3 pivot row", "grid.sync", "bookkeeping + sweep", "grid.sync"]


def dump(tag):
    L = _lib.lib()
    out = (ctypes.c_ulonglong * (32 * 12))()
    L.lpb_debug_single_timing.restype = ctypes.c_int
    assert L.lpb_debug_single_timing(out) == 0
    t = np.array(out[:], dtype=np.float64).reshape(32, 12)
    rows = {}
    for r in range(32):
        if t[r, 10] > 0:
            rows[r] = t[r, :10] / t[r, 10] / 1e3
    if not rows:
        print(tag, "no pivots recorded")
        return
    mean = np.mean(list(rows.values()), axis=0)
    print("%s: %d rank(s), %d pivots, %.2f us per pivot" % (tag, len(rows), int(t[min(rows), 10]), mean.sum()))
    print("   " + "  ".join("%s %.2f" % (n, v) for n, v in zip(NAMES, mean)))
    return {"tag": tag, "us_per_pivot": float(mean.sum()), "phases_us": dict(zip(NAMES, [float(v) for v in mean]))}


def main_sharded():
    """Under torchrun: the config-5 LP column-sharded over the ranks, 60 pivots; every rank prints its own timers
    (pipelined mode: slots 0..5 are the look-ahead's sub-phases, 'bookkeeping + sweep' its total on the control CTA)."""
    import torch
    import torch.distributed as dist
    from linear_programming_b200.sharded import ShardedComm, column_shard
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    n, m = int(os.environ.get("PH_N", "16383")), int(os.environ.get("PH_M", "32767"))
    comm = ShardedComm(max_n=n)
    lo, cnt = column_shard(m, world, rank)
    A, b, c = workloads.single_lp(n, m, seed=5, col_lo=lo, col_cnt=cnt)
    dA, db, dc = (torch.from_numpy(v).to(dev) for v in (A, b, c))
    for rep in range(2):
        comm.solve(dA, db, dc, n, m, max_pivots=60)
        torch.cuda.synchronize()
        dist.barrier()
        for rk in range(world):
            if rk == rank and rep == 1 and rank in (0, 1, world - 1):
                global NAMES
                NAMES = ["LA bookkeeping+bcol", "LA objective row", "LA key exchange", "LA column from A", "LA broadcast/receive",
                         "LA ratio", "-", "-", "look-ahead total", "barrier wait"]
                dump("rank %d of %d, %dx%d" % (rank, world, n, m))
            elif rk == rank:
                dump_quiet()
            dist.barrier()
    comm.close()
    dist.destroy_process_group()


def dump_quiet():
    L = _lib.lib()
    out = (ctypes.c_ulonglong * (32 * 12))()
    L.lpb_debug_single_timing.restype = ctypes.c_int
    L.lpb_debug_single_timing(out)


def main():
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        return main_sharded()
    n, m, cap = 100, 100, 1500
    A, b, c = (v[0] for v in workloads.simplex_batch("D2", 1, n, m, seed=9))
    res = []
    for cpr in (296, 37):
        os.environ["LPB_DEBUG_SINGLE_CPR"] = str(cpr)
        lp.solve_batch(A[None], b[None], c[None], mode="single", max_pivots=cap)
        dump("warm-up")
        lp.solve_batch(A[None], b[None], c[None], mode="single", max_pivots=cap)
        res.append(dump("1 GPU, %d CTAs, %dx%d" % (cpr, n, m)))
    for ranks in (2, 8):
        os.environ["LPB_DEBUG_SINGLE_CPR"] = str(296 // ranks)
        solve_sharded_emulated(A, b, c, ranks, max_pivots=cap)
        dump("warm-up")
        solve_sharded_emulated(A, b, c, ranks, max_pivots=cap)
        res.append(dump("%d emulated ranks x %d CTAs, %dx%d" % (ranks, 296 // ranks, n, m)))
    print(json.dumps(res))


if __name__ == "__main__":
    main()
