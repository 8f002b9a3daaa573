#!/usr/bin/env python3
"""Pivots/s of the batched kernel with a fixed pivot budget per LP (every LP stops at --cap pivots, so the rate
does not depend on the pivot path -- usable with experiment builds whose arithmetic is deliberately altered).

    [LPB_LIB_PATH=...] [LPB_BATCH_KERNEL=w] python tools/gpu/pivot_rate.py [--batch 4096] [--cap 400] [--dist D2]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from linear_programming_b200 import _lib, workloads  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--cap", type=int, default=400)
    ap.add_argument("--dist", default="D2")
    ap.add_argument("--n", type=int, default=100)
    ap.add_argument("--m", type=int, default=100)
    args = ap.parse_args()
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    B, n, m = args.batch, args.n, args.m
    A, b, c = (torch.from_numpy(v).to(dev) for v in workloads.simplex_batch(args.dist, B, n, m, seed=2))
    st = torch.empty(B, dtype=torch.int32, device=dev)
    npv = torch.empty(B, dtype=torch.int32, device=dev)
    x = torch.empty((B, m), dtype=torch.float64, device=dev)
    obj = torch.empty(B, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream

    def run():
        _lib.check(L.lpb200_simplex_batch_device(B, n, m, A.data_ptr(), b.data_ptr(), c.data_ptr(), 1e-9, args.cap,
                                                 st.data_ptr(), x.data_ptr(), obj.data_ptr(), npv.data_ptr(), None, 0,
                                                 stream))
    run()
    torch.cuda.synchronize()
    best = None
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    piv = int(npv.sum().item())
    print(json.dumps({"lib": os.environ.get("LPB_LIB_PATH", "default"), "kernel": os.environ.get("LPB_BATCH_KERNEL", "auto"),
                      "ms": best, "pivots": piv, "Mpivots_per_s": piv / best / 1e3,
                      "cycles_per_pivot_per_cta": 2 * 148 * 1.965e9 / (piv / (best * 1e-3))}))


if __name__ == "__main__":
    main()
