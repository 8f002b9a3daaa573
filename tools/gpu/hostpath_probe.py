import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import linear_programming_b200 as lp
from linear_programming_b200 import workloads
sizes = [(3, 3, 2000), (5, 4, 2100), (7, 7, 2200), (2, 6, 1900), (3, 3, 2000), (7, 7, 2200), (5, 4, 2100)]
for rep in range(2):
    for n, m, B in sizes:
        A, b, c = workloads.simplex_batch("D3", B, n, m, seed=1)
        t0 = time.perf_counter()
        lp.solve_batch(A, b, c, max_pivots=5000, log_cap=64)
        print(rep, n, m, B, "%.3f ms" % ((time.perf_counter() - t0) * 1e3), flush=True)
