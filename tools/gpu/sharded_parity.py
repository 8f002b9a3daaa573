#!/usr/bin/env python3
"""Multi-process parity check of the column-sharded solver (run under torchrun, one rank per
GPU): every rank must return the oracle's status, pivot sequence and x bits.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tools/gpu/sharded_parity.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from linear_programming_b200 import workloads  # noqa: E402
from linear_programming_b200.sharded import ShardedComm, column_shard  # noqa: E402
import oracle  # noqa: E402  (checker)


def main():
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    comm = ShardedComm(max_n=400)
    ok = True
    report = []
    for dist_name, n, m, cap in (("D1", 300, 5000, 120), ("D2", 96, 700, 200), ("D3", 120, 1000, 200),
                                 ("D3", 7, 5, 50), ("D2", 200, 40000, 60)):
        A, b, c = workloads.simplex_batch(dist_name, 1, n, m, seed=11)
        A, b, c = A[0], b[0], c[0]
        lo, cnt = column_shard(m, world, rank)
        dA = torch.from_numpy(np.ascontiguousarray(A[:, lo:lo + cnt])).to(dev)
        for rule in ("reference", "bland", "dantzig", "devex"):  # the corrected rules: Dantzig's key is two packets per rank
            want = oracle.simplex(A, b, c, max_pivots=cap, rule=rule)
            res = None
            for _ in range(2):  # twice on the same communicator: epochs carry over between solves
                res = comm.solve(dA, torch.from_numpy(b).to(dev), torch.from_numpy(c).to(dev), n, m,
                                 max_pivots=cap, log_cap=cap, rule=rule)
                torch.cuda.synchronize()
            good = (int(res.status[0]) == want["status"] and res.pivots(0) == want["pivots"]
                    and np.array_equal(res.x[0].cpu().numpy().view(np.int64), want["x"].view(np.int64)))
            ok = ok and good
            report.append({"dist": dist_name, "n": n, "m": m, "rule": rule, "status": want["status"],
                           "pivots": want["npivots"], "match": bool(good)})
    flags = [None] * world
    dist.all_gather_object(flags, bool(ok))
    if rank == 0:
        print(json.dumps({"world": world, "all_ranks_match_oracle": all(flags), "cases": report}), flush=True)
    comm.close()
    dist.destroy_process_group()
    sys.exit(0 if all(flags) else 1)


if __name__ == "__main__":
    main()
