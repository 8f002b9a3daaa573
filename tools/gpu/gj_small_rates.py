"""Batched Gauss-Jordan throughput on small systems (device-resident, CUDA events)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import linear_programming_b200 as lp
from linear_programming_b200 import workloads
for n, B in ((4, 200000), (20, 100000), (64, 20000), (128, 8000), (256, 2000), (512, 1184)):
    a, b = workloads.gauss_jordan_batch(B, n=n, seed=1, swap_frac=0.01 if n > 2 else 0.0)
    da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    lp.solve_equation_batch(da, db); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); x = lp.solve_equation_batch(da, db); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print("n=%d B=%d  %.2f ms  %.3f M systems/s  %.1f GFLOP-equivalent/s (2n^3)" % (n, B, ms, B / ms / 1e3, 2.0 * n**3 * B / ms / 1e6), flush=True)
