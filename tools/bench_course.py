#!/usr/bin/env python3
"""Configs 0 and 1 of BASELINE.json (SURVEY 8(d) Config 1/2): the course instances and the diet-size LPs.

  * every week-2 fixture held in tests/golden/simplex_golden.json: status / pivot count / pivot order / x
    parity with the unmodified reference (golden), single-LP latency through the Python API (host
    buffers, one problem per call) and of the kernel alone (device-resident, CUDA events), beside the
    reference's own wall time per file (tests/golden/reference_course_times.json, measured where the
    reference exists, see tools/time_reference_course.py);
  * Config 2: 10^5 LPs from the stress.py:23-26 distribution (7 % of b forced to 0), solved grouped by
    shape through the host API (copies included): LPs/s, parity rate against the CPU oracle, and the
    oracle's own LPs/s on all host threads.
Prints one JSON object.
"""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import linear_programming_b200 as lp
from linear_programming_b200 import workloads


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stress", type=int, default=100000)
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    import oracle  # checker + CPU baseline leg only
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "simplex_golden.json")))
    ref_times = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_course_times.json")))
    cases = [c for c in golden["cases"] if c["name"].startswith("week-2/") and "/fail/" not in c["name"]]
    rows = []
    for case in cases:
        n, m = case["n"], case["m"]
        A = np.array(case["A"], dtype=np.float64).reshape(1, n, m)
        b = np.array(case["b"], dtype=np.float64).reshape(1, n)
        c = np.array(case["c"], dtype=np.float64).reshape(1, m)
        cap = case["npivots"] + 8
        res = lp.solve_batch(A, b, c, log_cap=cap, max_pivots=cap)
        ok = (int(res.status[0]) == case["status"] and int(res.npivots[0]) == case["npivots"]
              and [list(q) for q in res.pivots(0)] == case["pivots"]
              and (case["status"] != 0 or [float(v).hex() for v in res.x[0]] == case["x"]))
        host = []
        for _ in range(args.reps):
            t0 = time.perf_counter()
            lp.solve_batch(A, b, c)
            host.append(time.perf_counter() - t0)
        dA, db, dc = (torch.from_numpy(v).cuda() for v in (A, b, c))
        lp.solve_batch(dA, db, dc)
        torch.cuda.synchronize()
        dev = []
        for _ in range(args.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            lp.solve_batch(dA, db, dc)
            e1.record()
            torch.cuda.synchronize()
            dev.append(e0.elapsed_time(e1) * 1e-3)
        t0 = time.perf_counter()
        oracle.simplex(A[0], b[0], c[0])
        t_or = time.perf_counter() - t0
        ref = ref_times["cases"].get(case["name"], {})
        rows.append({"name": case["name"], "n": n, "m": m, "status": case["status"], "npivots": case["npivots"],
                     "parity_with_reference": bool(ok),
                     "gpu_host_api_us": statistics.median(host) * 1e6,
                     "gpu_device_us": statistics.median(dev) * 1e6,
                     "cpu_oracle_us": t_or * 1e6,
                     "reference_process_s": ref.get("seconds")})
    # ---- Config 2: stress distribution ----
    probs = workloads.stress_batch(args.stress, seed=0)
    groups = {}
    for A, b, c in probs:
        groups.setdefault(A.shape, []).append((A, b, c))
    packed = {k: tuple(np.stack([q[i] for q in v]) for i in range(3)) for k, v in groups.items()}
    for A, b, c in packed.values():   # warm-up (arena, module load)
        lp.solve_batch(A[:8], b[:8], c[:8])
    t0 = time.perf_counter()
    timed = {k: lp.solve_batch(A, b, c, max_pivots=5000) for k, (A, b, c) in packed.items()}
    t_gpu = time.perf_counter() - t0
    # untimed second pass with the pivot log, for the pivot-ORDER check
    results = {k: lp.solve_batch(A, b, c, max_pivots=5000, log_cap=64) for k, (A, b, c) in packed.items()}
    for k in results:
        assert np.array_equal(results[k].status, timed[k].status)
        assert np.array_equal(results[k].x.view(np.int64), timed[k].x.view(np.int64))
    # the same 49 launches with the inputs already on the device (CUDA events, no sync in between)
    dev_in = {k: tuple(torch.from_numpy(v).cuda() for v in t) for k, t in packed.items()}
    for k, (A, b, c) in dev_in.items():
        lp.solve_batch(A, b, c, max_pivots=5000)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    keep = [lp.solve_batch(A, b, c, max_pivots=5000) for (A, b, c) in dev_in.values()]
    e1.record()
    torch.cuda.synchronize()
    t_dev = e0.elapsed_time(e1) * 1e-3
    t0 = time.perf_counter()
    want = {k: oracle.simplex_batch(A, b, c, max_pivots=5000) for k, (A, b, c) in packed.items()}
    t_cpu = time.perf_counter() - t0
    total = agree = 0
    hist = {}
    for k, res in results.items():
        st, x, obj, npiv, _ = want[k]
        same = (res.status == st) & (res.npivots == npiv) & (res.x.view(np.int64) == x.view(np.int64)).all(axis=1)
        total += len(st)
        agree += int(same.sum())
        for s in st:
            hist[int(s)] = hist.get(int(s), 0) + 1
    # pivot ORDER against the oracle on a sample (the oracle's batch driver keeps no log)
    order_checked = order_ok = 0
    for k, (A, b, c) in packed.items():
        res = results[k]
        for p in range(min(40, A.shape[0])):
            o = oracle.simplex(A[p], b[p], c[p], max_pivots=5000)
            order_checked += 1
            order_ok += int(res.pivots(p) == o["pivots"][:64])
    out = {"config_1_course_instances": rows,
           "config_2_stress": {"lps": total, "shape_groups": len(packed), "gpu_host_api_seconds": t_gpu,
                               "gpu_lps_per_s": total / t_gpu, "gpu_device_resident_seconds": t_dev,
                               "gpu_device_resident_lps_per_s": total / t_dev, "cpu_oracle_seconds": t_cpu,
                               "cpu_oracle_lps_per_s": total / t_cpu, "cpu_threads": os.cpu_count(),
                               "status_x_pivotcount_agreement": agree / total,
                               "pivot_order_checked": order_checked, "pivot_order_agreement": order_ok / order_checked,
                               "status_histogram": {str(k): v for k, v in sorted(hist.items())}},
           "reference_note": ref_times.get("note"), "reference_interpreter_startup_s": ref_times.get("interpreter_startup_seconds")}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
