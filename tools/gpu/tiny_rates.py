"""Device-resident LPs/s on diet-size problems (Config 2 distribution), two-warp variant vs the 8+1 warp kernel."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import linear_programming_b200 as lp
from linear_programming_b200 import workloads
import oracle
probs = workloads.stress_batch(200000, seed=0)
groups = {}
for A, b, c in probs:
    groups.setdefault(A.shape, []).append((A, b, c))
packed = {k: tuple(np.stack([q[i] for q in v]) for i in range(3)) for k, v in groups.items()}
dev = {k: tuple(torch.from_numpy(v).cuda() for v in t) for k, t in packed.items()}
for which in ("smem", "tiny", "smem", "tiny"):
    if which == "smem": os.environ["LPB_BATCH_KERNEL"] = "smem"
    else: os.environ.pop("LPB_BATCH_KERNEL", None)
    for (A, b, c) in dev.values(): lp.solve_batch(A, b, c, max_pivots=5000)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res = {k: lp.solve_batch(A, b, c, max_pivots=5000) for k, (A, b, c) in dev.items()}
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    ok = True
    for k in list(res)[:12]:
        st, x, obj, npiv, _ = oracle.simplex_batch(*packed[k], max_pivots=5000)
        ok &= bool(np.array_equal(res[k].status.cpu().numpy(), st) and np.array_equal(res[k].x.cpu().numpy().view(np.int64), x.view(np.int64)) and np.array_equal(res[k].npivots.cpu().numpy(), npiv))
    print(which, "%.2f ms  %.1f M LPs/s  parity(12 groups)=%s" % (ms, len(probs) / ms / 1e3, ok), flush=True)
# a mid-size shape that still qualifies (30x30) and one that does not (31x31)
for n, m in ((30, 30), (20, 28)):
    A, b, c = workloads.simplex_batch("D3", 20000, n, m, seed=3)
    dA, db, dc = (torch.from_numpy(v).cuda() for v in (A, b, c))
    for which in ("smem", "tiny"):
        if which == "smem": os.environ["LPB_BATCH_KERNEL"] = "smem"
        else: os.environ.pop("LPB_BATCH_KERNEL", None)
        lp.solve_batch(dA, db, dc); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); r = lp.solve_batch(dA, db, dc); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(n, m, which, "%.2f ms  %.2f M LPs/s  %.1f M pivots/s" % (ms, 20000 / ms / 1e3, int(r.npivots.sum()) / ms / 1e3), flush=True)
