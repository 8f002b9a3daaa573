import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import linear_programming_b200 as lp
from linear_programming_b200 import workloads
import oracle
B = int(sys.argv[1]); dist = sys.argv[2]; start = int(sys.argv[3]) if len(sys.argv) > 3 else 0
A, b, c = workloads.simplex_batch(dist, B, 100, 100, seed=2, start=start)
res = lp.solve_batch(A, b, c, max_pivots=50000)
st, x, obj, npiv, _ = oracle.simplex_batch(A, b, c, max_pivots=50000)
bad = np.nonzero((res.status != st) | (res.npivots != npiv) | (res.x.view(np.int64) != x.view(np.int64)).any(axis=1))[0]
print(B, dist, "mismatches", len(bad), bad[:10], "pivots", int(npiv.sum()))
