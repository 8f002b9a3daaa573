#!/usr/bin/env python3
"""Config 5: one large dense LP on the grid-wide kernel -- ms/pivot and achieved HBM GB/s.

    python tools/bench_single.py [--n 16383] [--m 32767] [--pivots 200] [--reps 2]

Algorithmic bytes per pivot = 16*(n+1)*(m+1) (every condensed-tableau element read and
written once).  Inputs are generated on the host, copied once; the timed region
(CUDA events on the launching stream) is the whole cooperative kernel including tableau set-up,
so the set-up cost (one 2x4.3 GB pass) is also measured separately with pivots=0.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from linear_programming_b200 import _lib, workloads  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=16383)
    ap.add_argument("--m", type=int, default=32767)
    ap.add_argument("--pivots", type=int, default=200)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--dist", default="D1")
    args = ap.parse_args()
    n, m = args.n, args.m
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    A, b, c = workloads.single_lp(n, m, seed=5, dist=args.dist)
    dA = torch.from_numpy(A).to(dev)
    db = torch.from_numpy(b).to(dev)
    dc = torch.from_numpy(c).to(dev)
    del A
    wsb = int(L.lpb200_simplex_single_workspace_bytes(n, m))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    st = torch.zeros(1, dtype=torch.int32, device=dev)
    npv = torch.zeros(1, dtype=torch.int32, device=dev)
    x = torch.zeros(m, dtype=torch.float64, device=dev)
    obj = torch.zeros(1, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream

    def run(cap):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(L.lpb200_simplex_single_device(n, m, dA.data_ptr(), db.data_ptr(), dc.data_ptr(), 1e-9, cap,
                                                  st.data_ptr(), x.data_ptr(), obj.data_ptr(), npv.data_ptr(),
                                                  None, 0, ws.data_ptr(), wsb, stream))
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), int(npv.item()), int(st.item())

    run(2)  # warm-up
    setup_ms = min(run(0)[0] for _ in range(args.reps))
    best = None
    for _ in range(args.reps):
        ms, piv, status = run(args.pivots)
        if best is None or ms < best[0]:
            best = (ms, piv, status)
    ms, piv, status = best
    per_pivot_ms = (ms - setup_ms) / max(piv, 1)
    alg = 16.0 * (n + 1) * (m + 1)
    peak = 6549.4
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    gbs = alg / (per_pivot_ms * 1e-3) / 1e9
    print(json.dumps({
        "workload": "configs[4]: single dense LP, condensed tableau %d x %d (n=%d, m=%d), %s" % (n + 1, m + 1, n, m, args.dist),
        "pivots": piv, "status": status, "total_ms": ms, "setup_ms": setup_ms, "ms_per_pivot": per_pivot_ms,
        "algorithmic_bytes_per_pivot": alg, "achieved_GBps": gbs, "peak_GBps": peak, "frac": gbs / peak,
        "pivots_per_s": 1e3 / per_pivot_ms}))


if __name__ == "__main__":
    main()
