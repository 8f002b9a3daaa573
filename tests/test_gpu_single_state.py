"""GPU parity of the grid-wide / column-sharded kernel beyond the answer: the TABLEAU it stops on.

lpb200_simplex_single_state_device exports, by variable id, the objective row (what pick_entering_idx scans,
simplex.py:390-392) and the constant column (the b_i of the ratio test, :414); with max_pivots = K that is the
state after the first K pivots, compared here with oracle/simplex_oracle.c lp_oracle_simplex_state.  This is how
BASELINE.json configs[4] -- a 16384 x 32768 tableau no solver finishes inside a test -- is checked at FULL SIZE:
index arithmetic beyond 2^31 bytes, the 131072-slot set table (m = 32768 never resizes) and the
512-column tiling, on one GPU and on 8 column-sharded ranks."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _gpu_state(A, b, c, cap, rule=0):
    import torch
    from linear_programming_b200 import _lib
    L = _lib.lib()
    n, m = A.shape
    dev = torch.device("cuda", 0)
    dA, db, dc = (torch.from_numpy(np.ascontiguousarray(v)).to(dev) for v in (A, b, c))
    wsb = int(L.lpb200_simplex_single_workspace_bytes(n, m))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    st = torch.zeros(1, dtype=torch.int32, device=dev)
    npv = torch.zeros(1, dtype=torch.int32, device=dev)
    x = torch.zeros(m, dtype=torch.float64, device=dev)
    obj = torch.zeros(1, dtype=torch.float64, device=dev)
    log = torch.full((max(cap, 1), 2), -1, dtype=torch.int32, device=dev)
    nv = n + m + 1
    ov = torch.zeros(nv + 1, dtype=torch.float64, device=dev)
    cv = torch.zeros(nv, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(L.lpb200_simplex_single_ex_device(n, m, dA.data_ptr(), db.data_ptr(), dc.data_ptr(), 1e-9, cap, rule,
                                                 st.data_ptr(), x.data_ptr(), obj.data_ptr(), npv.data_ptr(),
                                                 log.data_ptr(), max(cap, 1), ws.data_ptr(), wsb, stream))
    _lib.check(L.lpb200_simplex_single_state_device(n, m, ws.data_ptr(), ov.data_ptr(), cv.data_ptr(), stream))
    torch.cuda.synchronize()
    k = int(npv.item())
    return {"status": int(st.item()), "npivots": k, "pivots": [tuple(int(v) for v in r) for r in log[:k].cpu().numpy()],
            "obj_by_var": ov[:nv].cpu().numpy(), "obj_const": float(ov[nv].item()), "const_by_var": cv.cpu().numpy(),
            "x": x.cpu().numpy(), "objective": float(obj.item())}


def _check(got, want, name):
    assert got["status"] == want["status"], name
    assert got["pivots"] == want["pivots"], name
    # == on doubles: exact for every non-zero; a never-touched -0.0 (= -A[i][j] with A[i][j] == 0) equals +0.0
    assert np.array_equal(got["obj_by_var"], want["obj_by_var"]), name
    assert np.array_equal(got["const_by_var"], want["const_by_var"]), name
    assert np.float64(got["obj_const"]).view(np.int64) == np.float64(want["obj_const"]).view(np.int64), name


def _check_emulated(A, b, c, cap, nranks, want, name):
    from linear_programming_b200.sharded import solve_sharded_emulated
    res = solve_sharded_emulated(A, b, c, nranks, max_pivots=cap, log_cap=max(cap, 1), state=True)
    nv = A.shape[0] + A.shape[1] + 1
    for g in range(nranks):
        assert int(res.status[g]) == want["status"], name
        assert res.pivots(g) == want["pivots"], name
        assert np.array_equal(res.const_by_var[g], want["const_by_var"]), name
        assert res.obj_by_var[g, nv] == want["obj_const"], name
    # every column is owned by exactly one rank: the sum over ranks is the full objective row
    assert np.array_equal(res.obj_by_var[:, :nv].sum(axis=0), want["obj_by_var"]), name
    owners = (res.obj_by_var[:, :nv] != 0).sum(axis=0)
    assert owners.max() <= 1, name


@pytest.mark.parametrize("dist,n,m,cap", [("D1", 200, 900, 20), ("D2", 150, 1300, 60), ("D3", 180, 400, 75),
                                          ("D3", 64, 40, 100000)])
def test_state_after_k_pivots_vs_oracle(oracle_lib, dist, n, m, cap):
    from linear_programming_b200 import workloads
    A, b, c = (v[0] for v in workloads.simplex_batch(dist, 1, n, m, seed=13))
    want = oracle_lib.simplex_state(A, b, c, max_pivots=cap, threads=2)
    _check(_gpu_state(A, b, c, cap), want, "single %s %dx%d" % (dist, n, m))
    for nranks in (3, 8):
        _check_emulated(A, b, c, cap, nranks, want, "emulated x%d %s %dx%d" % (nranks, dist, n, m))


def test_state_of_a_crash_class_is_zero(oracle_lib):
    A = np.array([[1.0, 2.0], [3.0, 1.0]])
    b, c = np.zeros(2), np.array([1.0, 1.0])
    got = _gpu_state(A, b, c, 10)
    assert got["status"] == 3 and not got["obj_by_var"].any() and not got["const_by_var"].any()


def test_config5_full_size_first_pivots_vs_oracle(oracle_lib):
    """BASELINE.json configs[4] at the size it names: 16383 x 32767 (tableau 16384 x 32768, 4.3 GB), first pivots
    on one GPU and on 8 emulated column-sharded ranks against the oracle (rows swept on all host threads)."""
    import torch
    from linear_programming_b200 import workloads
    if torch.cuda.get_device_properties(0).total_memory < 40e9:
        pytest.skip("needs ~20 GB of device memory")
    n, m, cap = 16383, 32767, 10
    A, b, c = workloads.single_lp(n, m, seed=5)
    want = oracle_lib.simplex_state(A, b, c, max_pivots=cap, threads=0)
    assert want["npivots"] == cap and want["status"] == 2
    _check(_gpu_state(A, b, c, cap), want, "single, full size")
    _check_emulated(A, b, c, cap, 8, want, "8 emulated ranks, full size")
