#!/usr/bin/env python3
"""Config 3: batched Gauss-Jordan, B systems of n x n fp64 -- systems/s and effective GB/s.

    python tools/bench_gj.py [--batch 4096] [--n 512] [--reps 3] [--check 8]

Contract bytes per system: right-looking 8*n^2*(n+3); the reference's full
sweep 16*n^2*(n+1) is printed beside it.  The kernel is panel-blocked (K = 16), so its real
HBM traffic is ~K times lower than either figure; 'effective' GB/s may exceed the HBM peak.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from linear_programming_b200 import _lib, workloads  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--n", type=int, default=512)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--check", type=int, default=8)
    ap.add_argument("--cpu", type=int, default=32, help="systems timed on the host cores (0 = skip)")
    args = ap.parse_args()
    B, n = args.batch, args.n
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    a, b = workloads.gauss_jordan_batch(B, n=n, seed=1)
    da, db = torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev)
    x = torch.empty_like(db)
    wsb = int(L.lpb200_gauss_jordan_workspace_bytes(B, n))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream

    def run():
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(L.lpb200_gauss_jordan_batch_device(B, n, da.data_ptr(), db.data_ptr(), x.data_ptr(),
                                                      ws.data_ptr(), wsb, stream))
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    run()
    ms = min(run() for _ in range(args.reps))
    out = {"workload": "configs[2]: batched Gauss-Jordan, %d systems %dx%d fp64" % (B, n, n),
           "ms": ms, "systems_per_s": B / (ms * 1e-3),
           "effective_GBps_right_looking": 8.0 * n * n * (n + 3) * B / (ms * 1e-3) / 1e9,
           "effective_GBps_full_sweep": 16.0 * n * n * (n + 1) * B / (ms * 1e-3) / 1e9,
           "fp64_ops_per_s": 2.0 * n * n * (n + 1) / 2 * B / (ms * 1e-3)}
    import oracle
    oracle.build()
    if args.check:
        want, _ = oracle.gauss_jordan_batch(a[:args.check], b[:args.check])
        got = x[:args.check].cpu().numpy()
        out["bit_exact_vs_oracle"] = bool(np.array_equal(got.view(np.int64), want.view(np.int64)))
        out["checked_systems"] = args.check
    if args.cpu:
        k = min(args.cpu, B)
        t0 = time.perf_counter()
        _, used = oracle.gauss_jordan_batch(a[:k], b[:k], full_sweep=True, threads=0)
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": k / dt, "unit": "systems/s", "cores": used, "kind": "port",
                               "sample": "%d systems, full sweep like energy_values.cc" % k}
        if oracle.ref_gauss_jordan_available():
            t0 = time.perf_counter()
            for i in range(min(4, k)):
                oracle.ref_gauss_jordan(a[i], b[i])
            dt = time.perf_counter() - t0
            out["cpu_reference_1core"] = {"value": min(4, k) / dt, "unit": "systems/s", "cores": 1,
                                          "kind": "reference", "sample": "%d systems through oracle/_ref" % min(4, k)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
