"""Which batched-kernel variant for which shape: pivots/s of every eligible variant over a grid of shapes
(LPB_BATCH_KERNEL forces one), and what the library picks on its own."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import json
import numpy as np, torch
import linear_programming_b200 as lp
from linear_programming_b200 import workloads
shapes = [(7, 7), (16, 16), (30, 30), (40, 40), (50, 50), (62, 62), (50, 100), (80, 80), (100, 100)]
out = {}
for n, m in shapes:
    B = max(4000, min(40000, 80000000 // (n * m * 8)))
    A, b, c = workloads.simplex_batch("D3", B, n, m, seed=3)
    dA, db, dc = (torch.from_numpy(v).cuda() for v in (A, b, c))
    row = {}
    for name in ("smem", "tiny", "small", "mid", "auto"):
        if name in ("tiny", "small") and (m + 3) // 2 > 32:
            continue
        os.environ.pop("LPB_BATCH_KERNEL", None)
        if name != "auto":
            os.environ["LPB_BATCH_KERNEL"] = name
        lp.solve_batch(dA, db, dc, max_pivots=20000); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); r = lp.solve_batch(dA, db, dc, max_pivots=20000); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        row[name] = {"Mpivots_per_s": int(r.npivots.sum()) / ms / 1e3, "MLPs_per_s": B / ms / 1e3}
    out["%dx%d" % (n, m)] = {"batch": B, **row}
    print("%dx%d B=%d | " % (n, m, B) + " | ".join("%s %.0f" % (k, v["Mpivots_per_s"]) for k, v in row.items()) + "  (M pivots/s)", flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/variant_grid.json", "w"), indent=1)
