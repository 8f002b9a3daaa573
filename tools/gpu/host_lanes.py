#!/usr/bin/env python3
"""End-to-end rate of lpb200_simplex_batch from pinned fp64 host memory, with and without the second
(host-narrowing) lane, plus the raw narrowing rate of the host pool: Config 4, D1.

    python tools/gpu/host_lanes.py [--batch 65536] [--threads 1,4,8,16]
Each configuration runs in a child process (the pool size is fixed at first use)."""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def child(batch):
    import ctypes
    import numpy as np
    import torch
    from linear_programming_b200 import _lib, workloads
    L = _lib.lib()
    n = m = 100
    A, b, c = workloads.simplex_batch("D1", batch, n, m, seed=2)
    hA, hb, hc = (torch.from_numpy(v).pin_memory() for v in (A, b, c))
    st = torch.empty(batch, dtype=torch.int32).pin_memory()
    x = torch.empty((batch, m), dtype=torch.float64).pin_memory()
    obj = torch.empty(batch, dtype=torch.float64).pin_memory()
    npv = torch.empty(batch, dtype=torch.int32).pin_memory()
    dst = np.empty(A.size, dtype=np.int8)
    t0 = time.perf_counter()
    ok = L.lpb200_narrow_f64_to_i8(hA.data_ptr(), dst.ctypes.data, A.size)
    t0 = time.perf_counter()
    ok = L.lpb200_narrow_f64_to_i8(hA.data_ptr(), dst.ctypes.data, A.size)
    narrow_gbs = A.nbytes / (time.perf_counter() - t0) / 1e9

    def step():
        _lib.check(L.lpb200_simplex_batch(batch, n, m, hA.data_ptr(), hb.data_ptr(), hc.data_ptr(), 1e-9, 20000,
                                          st.data_ptr(), x.data_ptr(), obj.data_ptr(), npv.data_ptr(), None, 0))
    step()
    times = []
    for _ in range(4):
        torch.cuda.synchronize()
        t = time.perf_counter()
        step()
        times.append(time.perf_counter() - t)
    lanes = (ctypes.c_int64 * 4)()
    L.lpb200_simplex_batch_last_lanes(lanes)
    import zlib
    print(json.dumps({"host_threads": int(L.lpb200_host_threads()), "narrow_env": os.environ.get("LPB_HOST_NARROW", "1"),
                      "narrow_ok": int(ok), "narrow_only_GBps": round(narrow_gbs, 1),
                      "e2e_LPs_per_s": round(batch / min(times)), "ms_best": round(min(times) * 1e3, 2),
                      "ms_all": [round(t * 1e3, 1) for t in times], "lanes": [int(v) for v in lanes],
                      "host_GBps_effective": round(A.nbytes / min(times) / 1e9, 1),
                      "x_crc": zlib.crc32(x.numpy().tobytes()), "status_crc": zlib.crc32(st.numpy().tobytes())}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=65536)
    ap.add_argument("--threads", default="0,4,8,16,32")
    ap.add_argument("--child", action="store_true")
    args = ap.parse_args()
    if args.child:
        return child(args.batch)
    print(json.dumps({"cpu_count": os.cpu_count()}))
    for t in args.threads.split(","):
        env = dict(os.environ)
        if t == "0":
            env["LPB_HOST_NARROW"] = "0"
        else:
            env["LPB_HOST_THREADS"] = t
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", "--batch", str(args.batch)],
                             env=env, capture_output=True, text=True)
        print(out.stdout.strip() or out.stderr[-800:], flush=True)


if __name__ == "__main__":
    main()
