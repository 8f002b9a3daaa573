"""GPU parity tests of the COLUMN-SHARDED single-LP path (BASELINE configs[4]).

The multi-GPU protocol (8-byte entering-key all-to-all + pivot-column broadcast as flagged
packets, inside the persistent kernel) is exercised on ONE GPU by running every rank as a
CTA group of one cooperative launch (lpb200_simplex_sharded_emulated) -- the very kernel,
packet format and exchange schedule the multi-process path uses.  Bar: every rank returns
the reference's status, pivot sequence and x bits."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case_arrays(case):
    A = np.array(case["A"], dtype=np.float64).reshape(case["n"], case["m"])
    return A, np.array(case["b"], dtype=np.float64), np.array(case["c"], dtype=np.float64)


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_golden_cases_on_emulated_ranks(simplex_golden, nranks):
    """Reference fixtures + seeded random LPs of every outcome class (incl. phase 1, where the
    auxiliary column lives on the LAST rank and is dropped in place)."""
    from linear_programming_b200.sharded import solve_sharded_emulated
    cases = [c for c in simplex_golden["cases"] if c["npivots"] <= 120]
    # a spread over shapes/outcomes, bounded so the test stays short (each case is one launch)
    by_status = {}
    for c in cases:
        by_status.setdefault(c["status"], []).append(c)
    picked = []
    for st, lst in sorted(by_status.items()):
        lst = sorted(lst, key=lambda c: -c["npivots"])
        picked += lst[:10] + lst[-4:]
    assert len(picked) >= 40
    for case in picked:
        A, b, c = _case_arrays(case)
        cap = case["npivots"] + 4
        res = solve_sharded_emulated(A, b, c, nranks, max_pivots=cap, log_cap=cap)
        for g in range(nranks):
            name = "%s rank %d/%d" % (case["name"], g, nranks)
            assert int(res.status[g]) == case["status"], name
            assert int(res.npivots[g]) == case["npivots"], name
            assert [list(q) for q in res.pivots(g)] == case["pivots"], name
            if case["status"] == 0:
                assert [float(v).hex() for v in res.x[g]] == case["x"], name
                assert float(res.objective[g]).hex() == case["optimum"], name


@pytest.mark.parametrize("nranks,dist,n,m", [(2, "D1", 150, 700), (4, "D2", 96, 300), (8, "D3", 120, 260),
                                             (8, "D2", 60, 2100)])
def test_mid_size_vs_oracle(oracle_lib, nranks, dist, n, m):
    from linear_programming_b200 import workloads
    from linear_programming_b200.sharded import solve_sharded_emulated
    A, b, c = workloads.simplex_batch(dist, 1, n, m, seed=9)
    A, b, c = A[0], b[0], c[0]
    cap = 250
    want = oracle_lib.simplex(A, b, c, max_pivots=cap)
    res = solve_sharded_emulated(A, b, c, nranks, max_pivots=cap, log_cap=cap)
    for g in range(nranks):
        assert int(res.status[g]) == want["status"]
        assert res.pivots(g) == want["pivots"]
        assert np.array_equal(res.x[g].view(np.int64), want["x"].view(np.int64))
        assert float(res.objective[g]).hex() == float(want["objective"]).hex()


def test_one_rank_emulated_equals_single_mode(oracle_lib):
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    from linear_programming_b200.sharded import solve_sharded_emulated
    A, b, c = workloads.simplex_batch("D3", 1, 40, 90, seed=4)
    r1 = solve_sharded_emulated(A[0], b[0], c[0], 1, max_pivots=300, log_cap=300)
    r2 = lp.solve_batch(A, b, c, mode="single", max_pivots=300, log_cap=300)
    assert int(r1.status[0]) == int(r2.status[0]) and r1.pivots(0) == r2.pivots(0)
    assert np.array_equal(r1.x[0].view(np.int64), r2.x[0].view(np.int64))


def test_two_processes_two_gpus_vs_oracle():
    """The real multi-process path (CUDA IPC peer mapping, one rank per GPU) when the box has
    at least two GPUs; on a one-GPU box the emulated-rank tests above cover the protocol."""
    import os
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    import socket
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with socket.socket() as sk:  # a free rendezvous port instead of a fixed one
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(root, "tools", "gpu", "sharded_parity.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert '"all_ranks_match_oracle": true' in out.stdout
