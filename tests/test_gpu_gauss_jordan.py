"""GPU parity tests for the batched Gauss-Jordan kernel (through the C ABI): bit-identical
to the compiled reference's outputs (golden) and to the oracle."""
import numpy as np
import pytest

from linear_programming_b200 import workloads
from test_oracle_gauss_jordan import _inputs

pytestmark = pytest.mark.gpu


def test_golden_outputs_of_the_reference(gj_golden):
    import linear_programming_b200 as lp
    for case in gj_golden["cases"]:
        a, b = _inputs(case)
        x = lp.solve_equation(a, b)
        got = [float(v).hex() for v in x]
        if case["name"].startswith("singular"):
            assert all(not np.isfinite(v) for v in x)  # the reference divides by zero too
            continue
        assert got == case["x"], case["name"]


def test_size_zero_prints_nothing():
    from linear_programming_b200 import gauss_jordan
    x = gauss_jordan.solve_equation(np.zeros((0, 0)), np.zeros(0))
    assert gauss_jordan.format_column(x) == ""


@pytest.mark.parametrize("n,count,frac", [(5, 40, 0.5), (16, 40, 0.5), (17, 40, 0.5), (100, 24, 0.3),
                                          (200, 8, 0.5), (257, 8, 0.5), (512, 12, 0.25)])
def test_batches_vs_oracle(oracle_lib, n, count, frac):
    import linear_programming_b200 as lp
    a, b = workloads.gauss_jordan_batch(count, n=n, seed=1, swap_frac=frac)
    x = lp.solve_equation_batch(a, b)
    want, _ = oracle_lib.gauss_jordan_batch(a, b, full_sweep=False)
    assert np.array_equal(x.view(np.int64), want.view(np.int64))
    # solves the system (size-independent property)
    resid = np.abs(np.einsum("bij,bj->bi", a, x) - b).max()
    assert resid < 1e-9


def test_consecutive_zero_pivots(oracle_lib):
    """week-2/energy_values/tests/05 style permutation systems: a swap at nearly every step,
    so panels are truncated again and again."""
    import linear_programming_b200 as lp
    rng = np.random.default_rng(4)
    n = 48
    a = np.zeros((6, n, n))
    b = rng.uniform(-1, 1, size=(6, n))
    for s in range(6):
        perm = rng.permutation(n)
        a[s, np.arange(n), perm] = rng.uniform(1, 2, size=n)
        a[s] += np.triu(rng.uniform(-0.1, 0.1, size=(n, n)), 1) * (rng.random((n, n)) < 0.2)
    x = lp.solve_equation_batch(a, b)
    want, _ = oracle_lib.gauss_jordan_batch(a, b, full_sweep=True)
    assert np.array_equal(x.view(np.int64), want.view(np.int64))


def test_torch_device_path(oracle_lib):
    import torch
    import linear_programming_b200 as lp
    a, b = workloads.gauss_jordan_batch(32, n=128, seed=1, swap_frac=0.2)
    x = lp.solve_equation_batch(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda())
    torch.cuda.synchronize()
    want, _ = oracle_lib.gauss_jordan_batch(a, b)
    assert np.array_equal(x.cpu().numpy().view(np.int64), want.view(np.int64))


@pytest.mark.parametrize("n,count", [(5, 40), (16, 40), (17, 40), (33, 30), (100, 16), (257, 6), (512, 6)])
def test_partial_pivoting_vs_oracle(oracle_lib, n, count):
    """LPB200_PIVOT_PARTIAL: the row exchanges happen inside the shared-memory panel and are
    replayed on the trailing columns -- bit-identical to the step-by-step CPU restatement of the
    same rule, on systems with tiny diagonal entries (an exchange on most steps)."""
    import linear_programming_b200 as lp
    from test_oracle_gauss_jordan import _hard_systems
    a, b, x_true = _hard_systems(count, n, seed=4)
    x = lp.solve_equation_batch(a, b, pivoting="partial")
    want, _ = oracle_lib.gauss_jordan_batch(a, b, full_sweep=False, pivoting="partial")
    assert np.array_equal(x.view(np.int64), want.view(np.int64))
    # and it is the accurate answer, unlike the reference's rule on the same input
    for s in range(min(count, 4)):
        ref = np.linalg.solve(a[s], b[s])
        assert np.abs(x[s] - ref).max() <= 1e-9 * max(1.0, np.abs(ref).max()) * max(1.0, np.linalg.cond(a[s]) * 1e-6)


def test_partial_pivoting_device_tensors_and_workload(oracle_lib):
    import torch
    import linear_programming_b200 as lp
    a, b = workloads.gauss_jordan_batch(8, n=256, seed=2, swap_frac=0.5)
    x = lp.solve_equation_batch(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), pivoting="partial")
    torch.cuda.synchronize()
    want, _ = oracle_lib.gauss_jordan_batch(a, b, full_sweep=False, pivoting="partial")
    assert np.array_equal(x.cpu().numpy().view(np.int64), want.view(np.int64))


@pytest.mark.parametrize("pairs", ["0", "1"])
@pytest.mark.parametrize("n,count,frac,pivoting", [(512, 6, 0.5, "reference"), (258, 6, 0.5, "reference"),
                                                   (300, 5, 1.0, "reference"), (34, 20, 0.5, "reference"),
                                                   (100, 10, 0.5, "reference"), (257, 4, 0.5, "reference"),
                                                   (386, 4, 0.0, "partial"), (512, 4, 0.0, "partial")])
def test_column_pair_variant_vs_oracle(oracle_lib, monkeypatch, pairs, n, count, frac, pivoting):
    """Both trailing-update variants forced on every shape class: the column-pair kernel (default for even
    n > 256) handles rows that are not a multiple of the block, pivot rows in the middle of a block, panels
    truncated by a zero pivot (odd first trailing column -> one-column fallback for that panel) and odd n
    (fallback throughout); the one-column kernel is the default elsewhere.  Same bits from both."""
    import linear_programming_b200 as lp
    monkeypatch.setenv("LPB_DEBUG_GJ_PAIRS", pairs)
    if pivoting == "partial":
        from test_oracle_gauss_jordan import _hard_systems
        a, b, _ = _hard_systems(count, n, seed=6)
    else:
        a, b = workloads.gauss_jordan_batch(count, n=n, seed=3, swap_frac=frac)
    x = lp.solve_equation_batch(a, b, pivoting=pivoting)
    want, _ = oracle_lib.gauss_jordan_batch(a, b, full_sweep=False, pivoting=pivoting)
    assert np.array_equal(x.view(np.int64), want.view(np.int64))


@pytest.mark.parametrize("n", [6, 20, 100, 300, 512])
@pytest.mark.parametrize("pivoting", ["reference", "partial"])
def test_per_system_status_finds_the_singular_systems(oracle_lib, n, pivoting):
    """lpb200_gauss_jordan_batch_status: -1 for a system whose every step found a pivot, else the first step whose
    column holds no non-zero at or below the diagonal -- where the reference's SelectPivotElement finds nothing
    to swap with (energy_values.cc:100-110) and SolveEquation goes on to divide by zero."""
    import linear_programming_b200 as lp
    count = 9
    a, b = workloads.gauss_jordan_batch(count, n=n, seed=5, swap_frac=0.4)
    want_status = np.full(count, -1, dtype=np.int32)
    for s, k in ((1, 0), (4, n // 2), (7, n - 1)):
        a[s, :, k] = 0.0  # column k is zero and stays zero under every earlier step: step k has no pivot
        want_status[s] = k
    x, status = lp.solve_equation_batch(a, b, pivoting=pivoting, return_status=True)
    assert np.array_equal(status, want_status)
    good = want_status < 0
    want, _ = oracle_lib.gauss_jordan_batch(a[good], b[good], pivoting=pivoting)
    assert np.array_equal(x[good].view(np.int64), want.view(np.int64))
    assert all((~np.isfinite(x[s])).any() for s in np.nonzero(~good)[0])
    # the device entry point reports the same
    import torch
    xd, sd = lp.solve_equation_batch(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), pivoting=pivoting,
                                     return_status=True)
    assert np.array_equal(sd.cpu().numpy(), want_status)
    assert np.array_equal(xd.cpu().numpy()[good].view(np.int64), want.view(np.int64))


def test_engine_vs_the_compiled_reference_at_config3_size(oracle_lib):
    """Straight against the UNMODIFIED reference (energy_values.cc compiled with its own flags, oracle/_ref -- it travels
    to the GPU box), not through the restatement: 8 systems of 512 x 512 (BASELINE.json configs[2]), most with a planted
    zero pivot, through both entry points of the engine."""
    import linear_programming_b200 as lp
    if not oracle_lib.ref_gauss_jordan_available():
        pytest.skip("oracle/_ref not built (no reference tree where the repository was built)")
    a, b = workloads.gauss_jordan_batch(8, n=512, seed=13, swap_frac=0.75)
    want = np.stack([oracle_lib.ref_gauss_jordan(a[i], b[i]) for i in range(8)])
    got = lp.solve_equation_batch(a, b)
    assert np.array_equal(got.view(np.int64), want.view(np.int64))
    import torch
    xd, st = lp.solve_equation_batch(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), return_status=True)
    assert np.array_equal(xd.cpu().numpy().view(np.int64), want.view(np.int64)) and bool((st == -1).all())
