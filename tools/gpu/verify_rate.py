#!/usr/bin/env python3
"""HBM rate of the batched verify_bounds kernel on Config 4 (65536 x 100x100, device-resident): the kernel reads
A once, 8*n*m bytes per LP."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import linear_programming_b200 as lp
from linear_programming_b200 import workloads
B, n, m = 65536, 100, 100
A, b, c = (torch.from_numpy(v).cuda() for v in workloads.simplex_batch("D1", B, n, m, seed=2))
res = lp.solve_batch(A, b, c, max_pivots=20000)
lp.simplex.verify_bounds_batch(A, b, res.x, res.status); torch.cuda.synchronize()
best = 1e9
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fb = lp.simplex.verify_bounds_batch(A, b, res.x, res.status); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "MEASURED_PEAKS.json")))["hbm_gbs"]
gbs = B * n * m * 8 / (best * 1e-3) / 1e9
print(json.dumps({"kernel": "verify_bounds_kernel", "lps": B, "ms": best, "LPs_per_s": B / (best * 1e-3), "GBps": gbs,
                  "frac_of_measured_hbm_peak": gbs / peak, "bounded": int((res.status == 0).sum()),
                  "violations": int((fb >= 0).sum())}))
