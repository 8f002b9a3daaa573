"""GPU parity tests (run on the B200 box with -m gpu): the CUDA engine, called through the
C ABI, against the golden vectors of the unmodified reference and against the oracle.
Bar: status, pivot count and the full pivot sequence identical; x and optimum BIT-identical."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case_arrays(case):
    A = np.array(case["A"], dtype=np.float64).reshape(case["n"], case["m"])
    return A, np.array(case["b"], dtype=np.float64), np.array(case["c"], dtype=np.float64)


def _check_against_case(res, p, case):
    name = case["name"]
    assert int(res.status[p]) == case["status"], name
    assert int(res.npivots[p]) == case["npivots"], name
    assert [list(q) for q in res.pivots(p)] == case["pivots"], name
    if case["status"] == 0:
        assert [float(v).hex() for v in res.x[p]] == case["x"], name
        assert float(res.objective[p]).hex() == case["optimum"], name
    else:
        assert not np.any(res.x[p]), name


@pytest.mark.parametrize("mode", ["batch", "single"])
def test_every_golden_case(simplex_golden, mode):
    """All fixtures of week-2/diet/tests, week-2/diet/fail, week-2/ad_allocation/tests, the
    unit-test LPs and the seeded random LPs (all outcome classes) -- grouped by shape so
    each group is one batched launch."""
    import linear_programming_b200 as lp
    groups = {}
    for case in simplex_golden["cases"]:
        if mode == "single" and case["npivots"] > 400:
            continue  # the grid-wide kernel pays 4 grid syncs per pivot; keep the test short
        groups.setdefault((case["n"], case["m"]), []).append(case)
    checked = 0
    for (n, m), cases in sorted(groups.items()):
        A = np.stack([_case_arrays(c)[0] for c in cases])
        b = np.stack([_case_arrays(c)[1] for c in cases])
        c_ = np.stack([_case_arrays(c)[2] for c in cases])
        cap = max(c["npivots"] for c in cases) + 8
        res = lp.solve_batch(A, b, c_, log_cap=cap, max_pivots=cap, mode=mode)
        for p, case in enumerate(cases):
            _check_against_case(res, p, case)
            checked += 1
    assert checked >= (430 if mode == "batch" else 400)


def test_known_answer_optima():
    # week-2/diet/test_simplex.py:16-50
    import linear_programming_b200 as lp
    s = lp.Simplex([[-1, -1], [1, 0], [0, 1]], [-1, 2, 2], [-1, 2], 3, 2)
    s.solve()
    assert s.optimum == 4.0
    s = lp.Simplex([[1, 1, 3], [2, 2, 5], [4, 1, 2]], [30, 24, 36], [3, 1, 2], 3, 3)
    s.solve()
    assert s.optimum == 28
    s = lp.Simplex([[2, -1], [1, -5]], [2, -4], [2, -1], 2, 2)
    s.solve()
    assert s.optimum == 2.0


def test_reference_crash_classes_raise_like_the_reference(simplex_golden):
    import linear_programming_b200 as lp
    want = {3: AttributeError, 4: IndexError, 5: TypeError}
    seen = set()
    for case in simplex_golden["cases"]:
        if case["status"] in want and case["status"] not in seen:
            seen.add(case["status"])
            with pytest.raises(want[case["status"]]):
                lp.Simplex(case["A"], case["b"], case["c"], case["n"], case["m"]).solve()
    assert seen == {3, 4, 5}


@pytest.mark.parametrize("dist,n,m,count", [("D1", 100, 100, 96), ("D2", 100, 100, 24), ("D3", 100, 100, 12),
                                            ("D2", 37, 61, 64), ("D3", 64, 20, 64), ("D3", 150, 130, 6)])
def test_random_batches_vs_oracle(oracle_lib, dist, n, m, count):
    """Config-4 distributions at sizes the oracle finishes in seconds: bit-exact vs the oracle
    (itself pinned to the reference)."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch(dist, count, n, m, seed=2)
    cap = 20000
    res = lp.solve_batch(A, b, c, log_cap=0, max_pivots=cap)
    st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=cap)
    assert np.array_equal(res.status, st)
    assert np.array_equal(res.npivots, npiv)
    assert np.array_equal(res.x.view(np.int64), x.view(np.int64))
    assert np.array_equal(res.objective.view(np.int64), obj.view(np.int64))


def test_stress_distribution_vs_oracle(oracle_lib):
    """Config 2: the stress.py:23-26 distribution with zeros in b (every crash class)."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    probs = workloads.stress_batch(4000, seed=3)
    groups = {}
    for A, b, c in probs:
        groups.setdefault(A.shape, []).append((A, b, c))
    seen = set()
    for (n, m), lst in groups.items():
        A = np.stack([q[0] for q in lst]); b = np.stack([q[1] for q in lst]); c = np.stack([q[2] for q in lst])
        res = lp.solve_batch(A, b, c, log_cap=64, max_pivots=5000)
        for p in range(len(lst)):
            o = oracle_lib.simplex(A[p], b[p], c[p], max_pivots=5000)
            assert int(res.status[p]) == o["status"]
            assert res.pivots(p) == o["pivots"][:64]
            assert np.array_equal(res.x[p].view(np.int64), o["x"].view(np.int64))
            seen.add(o["status"])
    assert {0, -1, 1, 3} <= seen


def test_iteration_limit():
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch("D2", 4, 40, 40, seed=2)
    res = lp.solve_batch(A, b, c, max_pivots=5, log_cap=8)
    assert set(int(s) for s in res.status) <= {2, 1, 0}
    assert 2 in set(int(s) for s in res.status)
    assert all(int(k) <= 5 for k in res.npivots)


def test_torch_device_tensors_zero_copy(oracle_lib):
    import torch
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch("D1", 300, 100, 100, seed=2)
    dA, db, dc = (torch.from_numpy(v).cuda() for v in (A, b, c))
    res = lp.solve_batch(dA, db, dc)
    torch.cuda.synchronize()
    assert res.x.is_cuda
    st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c)
    assert np.array_equal(res.status.cpu().numpy(), st)
    assert np.array_equal(res.npivots.cpu().numpy(), npiv)
    assert np.array_equal(res.x.cpu().numpy().view(np.int64), x.view(np.int64))


def test_single_mode_mid_size_vs_oracle(oracle_lib):
    """Grid-wide kernel on a tableau that spans several tiles/strips (D1 and a phase-1 LP)."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    for dist, n, m in (("D1", 300, 1200), ("D3", 200, 700)):
        A, b, c = workloads.single_lp(n, m, seed=5, dist="D1") if dist == "D1" else \
            tuple(v[0] for v in workloads.simplex_batch("D3", 1, n, m, seed=9))
        cap = 600
        res = lp.solve_batch(A[None], b[None], c[None], mode="single", max_pivots=cap, log_cap=cap)
        o = oracle_lib.simplex(A, b, c, max_pivots=cap)
        assert int(res.status[0]) == o["status"]
        assert res.pivots(0) == o["pivots"]
        assert np.array_equal(res.x[0].view(np.int64), o["x"].view(np.int64))
        assert float(res.objective[0]).hex() == float(o["objective"]).hex()


def test_tolerance_is_a_runtime_argument(oracle_lib):
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch("D2", 16, 30, 30, seed=6)
    for tol in (1e-6, 1e-12):
        res = lp.solve_batch(A, b, c, tol=tol, max_pivots=5000)
        st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, tol=tol, max_pivots=5000)
        assert np.array_equal(res.status, st) and np.array_equal(res.npivots, npiv)
        assert np.array_equal(res.x.view(np.int64), x.view(np.int64))


def test_full_gpu_two_ctas_per_sm_vs_oracle(oracle_lib):
    """Enough 100x100 LPs to keep two CTAs resident on every SM with several problems per CTA
    (persistent work queue): warp skew between co-resident CTAs once exposed a read of varpos[]
    racing the bookkeeping lane, which never shows with one CTA per SM."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    for dist, count, cap in (("D2", 900, 400), ("D1", 1500, 50000)):
        A, b, c = workloads.simplex_batch(dist, count, 100, 100, seed=2, start=4096)
        for _ in range(2):
            res = lp.solve_batch(A, b, c, max_pivots=cap)
            st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=cap)
            assert np.array_equal(res.status, st)
            assert np.array_equal(res.npivots, npiv)
            assert np.array_equal(res.x.view(np.int64), x.view(np.int64))
            assert np.array_equal(res.objective.view(np.int64), obj.view(np.int64))


@pytest.mark.parametrize("dtype", ["int8", "int16", "int32", "float32"])
@pytest.mark.parametrize("dist,n,m,count", [("D1", 100, 100, 64), ("D2", 100, 100, 16), ("D2", 37, 61, 48),
                                            ("D3", 30, 22, 48)])
def test_typed_constraint_matrix_matches_float64_and_oracle(oracle_lib, dtype, dist, n, m, count):
    """lpb200_simplex_batch_ex: an integer / float32 A (what a caller that parsed the reference's
    input with int(), simplex.py:712-717, holds) gives the answers of the float64 call bit for bit
    -- m % 4 == 0 takes the vector loads, the other shapes the element loads."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch(dist, count, n, m, seed=5)
    At = A.astype(dtype)
    assert np.array_equal(At.astype(np.float64), A)  # the distributions are integers in [-100, 100]
    cap = 20000
    res = lp.solve_batch(At, b, c, log_cap=0, max_pivots=cap)
    st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=cap)
    assert np.array_equal(res.status, st)
    assert np.array_equal(res.npivots, npiv)
    assert np.array_equal(res.x.view(np.int64), x.view(np.int64))
    assert np.array_equal(res.objective.view(np.int64), obj.view(np.int64))


def test_typed_device_tensor_and_golden_cases(simplex_golden):
    """Typed A as a CUDA int16 tensor (zero-copy), and every golden fixture of the reference whose
    coefficients fit int16 through the typed host call: same status, pivots, x as the fp64 goldens."""
    import torch
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch("D1", 200, 100, 100, seed=6)
    ref = lp.solve_batch(A, b, c)
    res = lp.solve_batch(torch.from_numpy(A.astype(np.int16)).cuda(), torch.from_numpy(b).cuda(),
                         torch.from_numpy(c).cuda())
    torch.cuda.synchronize()
    assert np.array_equal(res.status.cpu().numpy(), ref.status)
    assert np.array_equal(res.x.cpu().numpy().view(np.int64), ref.x.view(np.int64))
    groups = {}
    for case in simplex_golden["cases"]:
        Af = _case_arrays(case)[0]
        if np.array_equal(Af.astype(np.int16).astype(np.float64), Af):
            groups.setdefault((case["n"], case["m"]), []).append(case)
    checked = 0
    for (n, m), cases in sorted(groups.items()):
        A = np.stack([_case_arrays(k)[0] for k in cases]).astype(np.int16)
        b = np.stack([_case_arrays(k)[1] for k in cases])
        c_ = np.stack([_case_arrays(k)[2] for k in cases])
        cap = max(k["npivots"] for k in cases) + 8
        res = lp.solve_batch(A, b, c_, log_cap=cap, max_pivots=cap, mode="batch")
        for p, case in enumerate(cases):
            _check_against_case(res, p, case)
            checked += 1
    assert checked >= 400


@pytest.mark.parametrize("rule", ["bland", "dantzig", "devex"])
def test_bland_rule_stress_distribution_vs_oracle(oracle_lib, rule):
    """LPB200_RULE_BLAND (corrected mode): status, pivot sequence and x bit-identical to the CPU
    restatement of the same rule (itself checked against HiGHS in tests/test_oracle_bland.py),
    on the distribution whose zeros in b crash the reference; includes x_aux drive-outs."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    groups = {}
    for A, b, c in workloads.stress_batch(5000, seed=13):
        groups.setdefault(A.shape, []).append((A, b, c))
    seen = set()
    oracle_lib.driveouts(reset=True)
    for (n, m), lst in groups.items():
        A = np.stack([q[0] for q in lst]); b = np.stack([q[1] for q in lst]); c = np.stack([q[2] for q in lst])
        res = lp.solve_batch(A, b, c, log_cap=64, max_pivots=5000, rule=rule)
        for p in range(len(lst)):
            o = oracle_lib.simplex(A[p], b[p], c[p], max_pivots=5000, rule=rule)
            assert int(res.status[p]) == o["status"]
            assert int(res.npivots[p]) == o["npivots"]
            assert res.pivots(p) == o["pivots"][:64]
            assert np.array_equal(res.x[p].view(np.int64), o["x"].view(np.int64))
            assert float(res.objective[p]).hex() == float(o["objective"]).hex()
            seen.add(o["status"])
    assert seen == {0, -1, 1}
    assert oracle_lib.driveouts() > 0  # the x_aux drive-out path was exercised


@pytest.mark.parametrize("dist,n,m,count", [("D1", 100, 100, 64), ("D2", 100, 100, 8), ("D3", 100, 100, 12),
                                            ("D3", 37, 61, 64), ("D3", 64, 20, 64)])
@pytest.mark.parametrize("rule", ["bland", "dantzig", "devex"])
def test_bland_rule_random_batches_vs_oracle(oracle_lib, dist, n, m, count, rule):
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch(dist, count, n, m, seed=9)
    cap = 200000
    res = lp.solve_batch(A.astype(np.int16), b, c, max_pivots=cap, rule=rule)
    st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=cap, rule=rule)
    assert np.array_equal(res.status, st)
    assert np.array_equal(res.npivots, npiv)
    assert np.array_equal(res.x.view(np.int64), x.view(np.int64))
    assert np.array_equal(res.objective.view(np.int64), obj.view(np.int64))
    assert set(int(s) for s in st) <= {0, -1, 1}


@pytest.mark.parametrize("n,m,count", [(1, 1, 64), (1, 7, 64), (7, 1, 64), (2, 129, 32), (129, 2, 32), (3, 255, 16),
                                       (160, 160, 4), (100, 250, 4), (250, 100, 4), (20, 1000, 4)])
@pytest.mark.parametrize("rule", ["reference", "bland", "dantzig", "devex"])
def test_edge_shapes_vs_oracle(oracle_lib, n, m, count, rule):
    """Degenerate and extreme shapes of the one-CTA-per-LP kernel: a single row or column, odd widths,
    more than 64 column pairs (lanes loop), and the largest tableaus shared memory holds (one CTA/SM)."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    assert lp.simplex.fits_batch_kernel(n, m)
    A, b, c = workloads.simplex_batch("D3", count, n, m, seed=21)
    cap = 30000
    res = lp.solve_batch(A, b, c, max_pivots=cap, log_cap=32, mode="batch", rule=rule)
    st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=cap, rule=rule)
    assert np.array_equal(res.status, st)
    assert np.array_equal(res.npivots, npiv)
    assert np.array_equal(res.x.view(np.int64), x.view(np.int64))
    assert np.array_equal(res.objective.view(np.int64), obj.view(np.int64))


def test_auto_mode_falls_back_to_the_grid_wide_kernel_when_smem_is_too_small(oracle_lib):
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    n = m = 166
    assert not lp.simplex.fits_batch_kernel(n, m)
    A, b, c = workloads.simplex_batch("D1", 2, n, m, seed=22)
    res = lp.solve_batch(A, b, c, max_pivots=400)
    st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=400)
    assert np.array_equal(res.status, st) and np.array_equal(res.npivots, npiv)
    assert np.array_equal(res.x.view(np.int64), x.view(np.int64))
    with pytest.raises(lp.EngineError):
        lp.solve_batch(A, b, c, mode="batch")


def test_empty_batch_and_bad_arguments():
    import linear_programming_b200 as lp
    res = lp.solve_batch(np.zeros((0, 3, 4)), np.zeros((0, 3)), np.zeros((0, 4)))
    assert res.status.shape == (0,) and res.x.shape == (0, 4)
    with pytest.raises(ValueError):
        lp.solve_batch(np.zeros((2, 3, 4)), np.zeros((2, 4)), np.zeros((2, 4)))
    with pytest.raises(ValueError):
        lp.solve_batch(np.zeros((1, 3, 4)), np.zeros((1, 3)), np.zeros((1, 4)), rule="steepest")
    with pytest.raises(lp.EngineError):
        lp.solve_batch(np.zeros((1, 3, 4)), np.zeros((1, 3)), np.zeros((1, 4)), tol=-1.0)


@pytest.mark.parametrize("dist,n,m,count", [("D1", 100, 100, 320), ("D3", 100, 100, 16), ("D3", 37, 61, 64)])
def test_register_resident_variant_vs_oracle(oracle_lib, monkeypatch, dist, n, m, count):
    """The opt-in register-resident kernel variant (LPB_BATCH_KERNEL=reg; slower than the default
    shared-memory kernel on B200, kept as a measured experiment -- DESIGN.md 5.1) stays bit-exact."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    monkeypatch.setenv("LPB_BATCH_KERNEL", "reg")
    A, b, c = workloads.simplex_batch(dist, count, n, m, seed=23)
    cap = 20000
    res = lp.solve_batch(A, b, c, max_pivots=cap, log_cap=16)
    st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=cap)
    assert np.array_equal(res.status, st)
    assert np.array_equal(res.npivots, npiv)
    assert np.array_equal(res.x.view(np.int64), x.view(np.int64))
    assert np.array_equal(res.objective.view(np.int64), obj.view(np.int64))


@pytest.mark.parametrize("variant", ["tiny", "small", "mid", "smem", "w"])
def test_every_kernel_variant_is_bit_exact(oracle_lib, simplex_golden, monkeypatch, variant):
    """The library picks a CTA size per shape and batch size (1+1 warps for small problems, 4+1 for
    course-size ones, 8+1 for the 100x100 batch); every variant, forced through LPB_BATCH_KERNEL, gives
    the reference's answers on the golden fixtures and the oracle's on random batches."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    monkeypatch.setenv("LPB_BATCH_KERNEL", variant)
    groups = {}
    for case in simplex_golden["cases"]:
        groups.setdefault((case["n"], case["m"]), []).append(case)
    checked = 0
    for (n, m), cases in sorted(groups.items()):
        A = np.stack([_case_arrays(k)[0] for k in cases])
        b = np.stack([_case_arrays(k)[1] for k in cases])
        c_ = np.stack([_case_arrays(k)[2] for k in cases])
        cap = max(k["npivots"] for k in cases) + 8
        res = lp.solve_batch(A, b, c_, log_cap=cap, max_pivots=cap, mode="batch")
        for p, case in enumerate(cases):
            _check_against_case(res, p, case)
            checked += 1
    assert checked >= 430
    for dist, n, m, count in (("D3", 37, 61, 600), ("D3", 50, 50, 600), ("D2", 20, 28, 600), ("D3", 62, 30, 600),
                              ("D1", 100, 100, 600)):
        A, b, c = workloads.simplex_batch(dist, count, n, m, seed=31)
        res = lp.solve_batch(A.astype(np.int16), b, c, max_pivots=20000)
        st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=20000)
        assert np.array_equal(res.status, st)
        assert np.array_equal(res.npivots, npiv)
        assert np.array_equal(res.x.view(np.int64), x.view(np.int64))
        assert np.array_equal(res.objective.view(np.int64), obj.view(np.int64))


def test_automatic_variant_choice_large_batches_vs_oracle(oracle_lib):
    """Batches large enough (> 2 LPs per SM) for the throughput variants to be picked automatically."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    for dist, n, m, count in (("D3", 7, 5, 3000), ("D3", 40, 40, 700), ("D3", 50, 100, 500), ("D2", 80, 80, 400)):
        A, b, c = workloads.simplex_batch(dist, count, n, m, seed=32)
        res = lp.solve_batch(A, b, c, max_pivots=30000)
        st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=30000)
        assert np.array_equal(res.status, st)
        assert np.array_equal(res.npivots, npiv)
        assert np.array_equal(res.x.view(np.int64), x.view(np.int64))


def _lanes():
    import ctypes
    from linear_programming_b200 import _lib
    out = (ctypes.c_int64 * 4)()
    assert _lib.lib().lpb200_simplex_batch_last_lanes(out) == 0
    return [int(v) for v in out]


@pytest.mark.parametrize("kind", ["integers", "mixed", "fractional"])
def test_two_lane_host_path_is_bit_identical(oracle_lib, monkeypatch, kind):
    """The host-pointer call on a large fp64 batch runs two lanes (copy engine ships fp64 chunks, host cores
    narrow other chunks to int8 exactly -- include/lp_b200.h).  Chunks shrunk to 64 KB so that 700 small LPs
    make ~90 chunks.  'mixed': every 5th LP holds a fraction or a -0.0, so its chunk must go over as fp64;
    'fractional': nothing can be narrowed and the second lane turns into a second copy lane.  Answers must
    equal the single-lane call and the oracle bit for bit."""
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch("D3", 700, 24, 31, seed=41)
    if kind == "mixed":
        A[::5, 3, 7] = 0.5
        A[2::10, 0, 0] = -0.0
    elif kind == "fractional":
        A = A + 0.25
    monkeypatch.setenv("LPB_HOST_CHUNK_BYTES", str(64 * 1024))
    monkeypatch.setenv("LPB_HOST_NARROW", "0")
    one = lp.solve_batch(A, b, c, max_pivots=20000, log_cap=16)
    assert _lanes()[1] == 0 and _lanes()[0] > 50
    monkeypatch.setenv("LPB_HOST_NARROW", "1")
    two = lp.solve_batch(A, b, c, max_pivots=20000, log_cap=16)
    cl, nl, narrowed, abytes = _lanes()
    assert cl + nl == _ceil_div(700, (64 * 1024) // _lp_bytes(24, 31, 16))
    if kind == "integers":
        assert narrowed == nl and nl > 0 and abytes < A.nbytes
    elif kind == "mixed":
        assert 0 <= narrowed <= nl
    else:
        assert narrowed == 0 and abytes == A.nbytes
    # no override: the MEASURED gate decides (a probe call runs the copy engine alone for its first chunks, then both
    # lanes; the verdict holds for the next calls) -- whichever way it falls, every chunk is served exactly once
    monkeypatch.delenv("LPB_HOST_NARROW")
    gated = [lp.solve_batch(A, b, c, max_pivots=20000, log_cap=16) for _ in range(3)]
    cl, nl, _, _ = _lanes()
    assert cl + nl == _ceil_div(700, (64 * 1024) // _lp_bytes(24, 31, 16)) and cl > 0
    st, x, obj, npiv, _ = oracle_lib.simplex_batch(A, b, c, max_pivots=20000)
    for r in [one, two] + gated:
        assert np.array_equal(r.status, st) and np.array_equal(r.npivots, npiv)
        assert np.array_equal(r.x.view(np.int64), x.view(np.int64))
        assert np.array_equal(r.objective.view(np.int64), obj.view(np.int64))
    assert np.array_equal(one.pivot_log, two.pivot_log)


def _lp_bytes(n, m, log_cap):
    return n * m * 8 + (n + 2 * m + 1) * 8 + 8 + log_cap * 2 * 4


def _ceil_div(a, b):
    return (a + b - 1) // b


def test_devex_needs_fewer_pivots_than_dantzig(oracle_lib):
    import linear_programming_b200 as lp
    from linear_programming_b200 import workloads
    A, b, c = workloads.simplex_batch("D2", 64, 100, 100, seed=2)
    rd = lp.solve_batch(A.astype(np.int8), b, c, rule="dantzig", max_pivots=200000)
    rx = lp.solve_batch(A.astype(np.int8), b, c, rule="devex", max_pivots=200000)
    assert np.array_equal(rd.status, rx.status)
    ok = rd.status == 0
    assert np.allclose(rd.objective[ok], rx.objective[ok], rtol=1e-6)  # different pivot paths, same optimum
    assert rx.npivots.sum() * 5 < rd.npivots.sum() * 4  # at least a fifth fewer pivots
    with pytest.raises(lp.EngineError):
        lp.solve_batch(A[:1], b[:1], c[:1], rule="devex", mode="single", tol=-1.0)
