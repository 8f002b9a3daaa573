"""Simplex.verify_bounds (week-2/diet/simplex.py:663-671), batched on the GPU (lpb200_verify_bounds_batch).

CPU part: the C restatement against the LIVE reference method when /root/reference is mounted (the reference
asserts on the module-global ``ansx``), and against a straight Python transcription everywhere.
GPU part: the kernel against the restatement -- staged and direct variants, skipped statuses, first failing row."""
import numpy as np
import pytest


def _transcription(A, b, x):
    """simplex.py:663-671 line by line; returns the first row whose assertion fires, -1 if none."""
    r = 0
    for row in A:
        dot = [row[i] * x[i] for i in range(len(row))]
        s = 0
        for v in dot:
            s += v
        if not (s <= (b[r]) + (1e-3)):
            return r
        r += 1
    return -1


def _cases(seed, count, n, m, spoil):
    rng = np.random.default_rng(seed)
    A = rng.integers(-100, 101, size=(count, n, m)).astype(np.float64)
    x = rng.uniform(0, 5, size=(count, m))
    x[rng.random((count, m)) < 0.5] = 0.0
    b = np.einsum("pij,pj->pi", A, x) + rng.uniform(0.0, 3.0, size=(count, n))
    for p in range(0, count, spoil):  # one violated constraint (by 0.002 .. 1) in every spoil-th LP
        r = int(rng.integers(0, n))
        b[p, r] -= rng.uniform(3.002, 4.0)
    return A, b, x


def test_oracle_matches_transcription_and_live_reference(oracle_lib):
    from oracle import ref_harness
    A, b, x = _cases(1, 60, 7, 9, 3)
    ref = ref_harness.load_reference() if ref_harness.available() else None
    bad = 0
    for p in range(len(A)):
        got = oracle_lib.verify_bounds(A[p], b[p], x[p])
        assert got == _transcription(A[p].tolist(), b[p].tolist(), x[p].tolist())
        bad += got >= 0
        if ref is not None:  # the unmodified method: AssertionError <=> some row fails
            ref.ansx = x[p].tolist()
            s = ref.Simplex([[int(v) for v in row] for row in A[p]], b[p].tolist(), [0.0] * 9, 7, 9)
            try:
                s.verify_bounds()
                fired = False
            except AssertionError:
                fired = True
            assert fired == (got >= 0)
    assert 10 <= bad <= 30


def test_oracle_boundary_is_inclusive(oracle_lib):
    A = np.array([[1.0, 0.0], [0.0, 1.0]])
    x = np.array([2.0, 3.0])
    assert oracle_lib.verify_bounds(A, np.array([2.0, 3.0]), x) == -1
    assert oracle_lib.verify_bounds(A, np.array([1.9, 3.0]), x) == 0
    assert oracle_lib.verify_bounds(A, np.array([2.0, 2.9]), x) == 1
    assert oracle_lib.verify_bounds(A, np.array([2.0, float("nan")]), x) == 1  # not (s <= nan)


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["default", "staged", "direct"])
@pytest.mark.parametrize("n,m,count", [(7, 9, 300), (100, 100, 200), (1, 1, 64), (3, 257, 40), (160, 170, 12),
                                       (300, 310, 6), (256, 33, 9), (100, 32, 50), (100, 31, 1500)])
def test_gpu_kernel_vs_oracle(oracle_lib, monkeypatch, kernel, n, m, count):
    """Three kernels: the pipelined one (default for n <= 256: column blocks of 32 through a cp.async double buffer),
    and the two of round 1 forced through LPB_DEBUG_VERIFY -- whole LP staged in shared memory / rows read straight
    from global memory ((300, 310) fits neither the pipelined nor the staged kernel)."""
    import linear_programming_b200 as lp
    if kernel != "default":
        monkeypatch.setenv("LPB_DEBUG_VERIFY", kernel)
    A, b, x = _cases(7 + n, count, n, m, 3)
    status = np.zeros(count, dtype=np.int32)
    status[1::4] = 1     # "Infinity": no x, skipped
    status[2::8] = -1
    want = np.array([oracle_lib.verify_bounds(A[p], b[p], x[p]) if status[p] == 0 else -1 for p in range(count)],
                    dtype=np.int32)
    got = lp.simplex.verify_bounds_batch(A, b, x, status)
    assert np.array_equal(got, want)
    assert (want >= 0).sum() >= 1
    allp = lp.simplex.verify_bounds_batch(A, b, x)  # status None: every LP
    assert np.array_equal(allp, np.array([oracle_lib.verify_bounds(A[p], b[p], x[p]) for p in range(count)]))


@pytest.mark.gpu
def test_gpu_device_tensors_and_solver_answers(oracle_lib):
    """End of the fuzz loop on the GPU: solve a stress batch, then verify every bounded answer in place."""
    import torch
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch("D3", 256, 24, 31, seed=5)
    dA, db, dc = (torch.from_numpy(v).cuda() for v in (A, b, c))
    res = lp.solve_batch(dA, db, dc, max_pivots=20000)
    fb = lp.simplex.verify_bounds_batch(dA, db, res.x, res.status)
    torch.cuda.synchronize()
    st, x = res.status.cpu().numpy(), res.x.cpu().numpy()
    want = np.array([oracle_lib.verify_bounds(A[p], b[p], x[p]) if st[p] == 0 else -1 for p in range(256)])
    assert np.array_equal(fb.cpu().numpy(), want)
    assert (st == 0).sum() > 20


@pytest.mark.gpu
def test_simplex_object_verify_bounds():
    import linear_programming_b200 as lp
    s = lp.Simplex([[1, 1], [1, -1]], [4, 2], [1, 1], 2, 2)
    assert s.solve()[0] == 0
    s.verify_bounds()
    s.assignment = [10.0, 0.0]
    with pytest.raises(AssertionError):
        s.verify_bounds()
