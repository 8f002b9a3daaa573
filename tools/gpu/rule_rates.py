"""LPs/s and pivots/s of the four pivoting rules on Config 4 (device-resident inputs, CUDA events)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import linear_programming_b200 as lp
from linear_programming_b200 import workloads

out = {}
for dist, B in (("D1", 65536), ("D2", 8192)):
    A, b, c = workloads.simplex_batch(dist, B, 100, 100, seed=2)
    dA = torch.from_numpy(A).to(torch.int8).cuda()
    db, dc = torch.from_numpy(b).cuda(), torch.from_numpy(c).cuda()
    for rule in ("reference", "bland", "dantzig", "devex"):
        lp.solve_batch(dA, db, dc, rule=rule)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        res = lp.solve_batch(dA, db, dc, rule=rule)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        piv = int(res.npivots.sum().item())
        st = torch.bincount(res.status + 1, minlength=4).tolist()
        out["%s_%s" % (dist, rule)] = {"batch": B, "ms": ms, "lps_per_s": B / ms * 1e3, "pivots": piv,
                                       "pivots_per_s": piv / ms * 1e3,
                                       "status_counts[-1,0,1,2]": st}
        print(dist, rule, json.dumps(out["%s_%s" % (dist, rule)]), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/rule_rates.json", "w"), indent=1)
