"""lpb200_gauss_jordan_batch_static_device -- the entry point that allocates nothing (no counterpart in the reference;
added for callers that cannot call cudaMalloc).  The working copy of `a` lives
in a 32 MiB file-scope __device__ array of the library instead of a caller workspace.

Checked through the C ABI against the existing entry point lpb200_gauss_jordan_batch_status_device (itself pinned to
the compiled reference energy_values.cc and the oracle in tests/test_gpu_gauss_jordan.py), with the bar those tests
use: x BIT-identical, status equal -- on two calls in a row that share the static array (a full-size one, then a
smaller one of another shape over its leftovers), for both pivot rules, plus the size limit and the no-allocation claim.

This file is named to be collected LAST; it passes alone in a fresh process and after
the other tests (nothing here depends on what ran before: the array needs no clearing)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LIMIT = 32 * 1024 * 1024  # LPB200_GJ_STATIC_SCRATCH_BYTES, include/lp_b200.h


def _systems(count, n, seed, singular=()):
    from linear_programming_b200 import workloads
    a, b = workloads.gauss_jordan_batch(count, n=n, seed=seed, swap_frac=0.5)
    for s, k in singular:  # column k all zero: step k finds no pivot, the reference divides by zero (inf/nan)
        a[s, :, k] = 0.0
    return a, b


class _Case:
    """Inputs and outputs of one batch on the GPU, allocated up front (the caller's business, not the library's)."""

    def __init__(self, a, b):
        import torch
        self.B, self.n = b.shape
        self.a = torch.from_numpy(a).cuda()
        self.b = torch.from_numpy(b).cuda()
        self.a_copy = self.a.clone()
        self.x_new = torch.full_like(self.b, float("nan"))
        self.x_old = torch.full_like(self.b, float("nan"))
        self.st_new = torch.full((self.B,), 77, dtype=torch.int32, device="cuda")
        self.st_old = torch.full((self.B,), 77, dtype=torch.int32, device="cuda")
        from linear_programming_b200 import _lib
        self.wsb = int(_lib.lib().lpb200_gauss_jordan_workspace_bytes(self.B, self.n))
        self.ws = torch.empty(self.wsb, dtype=torch.uint8, device="cuda")

    def run_static(self, pivoting, stream):
        from linear_programming_b200 import _lib
        return _lib.lib().lpb200_gauss_jordan_batch_static_device(
            self.B, self.n, self.a.data_ptr(), self.b.data_ptr(), self.x_new.data_ptr(), pivoting,
            self.st_new.data_ptr(), stream)

    def run_existing(self, pivoting, stream):
        from linear_programming_b200 import _lib
        _lib.check(_lib.lib().lpb200_gauss_jordan_batch_status_device(
            self.B, self.n, self.a.data_ptr(), self.b.data_ptr(), self.x_old.data_ptr(), pivoting,
            self.st_old.data_ptr(), self.ws.data_ptr(), self.wsb, stream))

    def assert_same(self, what):
        import torch
        got, want = self.x_new.cpu().numpy(), self.x_old.cpu().numpy()
        assert np.array_equal(got.view(np.int64), want.view(np.int64)), what + ": x differs from the existing entry point"
        assert torch.equal(self.st_new, self.st_old), what + ": status differs"
        assert torch.equal(self.a, self.a_copy), what + ": the input matrix was modified"


@pytest.mark.parametrize("pivoting", [0, 1])
def test_two_calls_in_a_row_match_the_existing_entry_point(pivoting):
    import torch
    from linear_programming_b200 import _lib
    stream = torch.cuda.current_stream().cuda_stream
    # first call: 16 systems of 512 x 512 = exactly the 32 MiB the static array holds (column-pair kernel)
    big = _Case(*_systems(16, 512, seed=21, singular=[(5, 300)]))
    assert big.B * big.n * big.n * 8 == LIMIT
    # second call, right behind it on the same stream: another shape over the first call's leftovers
    # (course-size systems, the 64-thread kernel), with zero pivots to swap and two singular systems
    small = _Case(*_systems(300, 20, seed=22, singular=[(0, 0), (123, 19)]))
    for case, name in ((big, "first call (16 x 512 x 512)"), (small, "second call (300 x 20 x 20)")):
        assert case.run_static(pivoting, stream) == 0, _lib.lib().lpb200_last_error()
    for case in (big, small):
        case.run_existing(pivoting, stream)
    torch.cuda.synchronize()
    big.assert_same("first call, pivoting %d" % pivoting)
    small.assert_same("second call, pivoting %d" % pivoting)
    assert big.st_new.cpu().tolist()[5] == 300 and (big.st_new.cpu().numpy() == -1).sum() == 15
    st = small.st_new.cpu().numpy()
    assert st[0] == 0 and st[123] == 19 and (st == -1).sum() == 298
    # and against the CPU oracle, for the systems that have a solution
    import oracle
    oracle.build()
    good = st == -1
    names = {0: "reference", 1: "partial"}
    want, _ = oracle.gauss_jordan_batch(small.a.cpu().numpy()[good], small.b.cpu().numpy()[good], pivoting=names[pivoting])
    assert np.array_equal(small.x_new.cpu().numpy()[good].view(np.int64), want.view(np.int64))


def test_in_place_and_without_status():
    """x may alias b (the reference's SolveEquation returns b, energy_values.cc:196) and status may be NULL."""
    import torch
    from linear_programming_b200 import _lib
    L = _lib.lib()
    stream = torch.cuda.current_stream().cuda_stream
    case = _Case(*_systems(40, 96, seed=23))
    case.run_existing(0, stream)
    xb = case.b.clone()
    assert L.lpb200_gauss_jordan_batch_static_device(case.B, case.n, case.a.data_ptr(), xb.data_ptr(), xb.data_ptr(), 0,
                                                     None, stream) == 0, L.lpb200_last_error()
    torch.cuda.synchronize()
    assert torch.equal(xb.view(torch.int64), case.x_old.view(torch.int64))
    assert L.lpb200_gauss_jordan_batch_static_device(0, 96, None, None, None, 0, None, stream) == 0  # empty batch


def test_sizes_beyond_the_static_array_are_rejected():
    import torch
    from linear_programming_b200 import _lib
    L = _lib.lib()
    stream = torch.cuda.current_stream().cuda_stream
    x = torch.full((17, 512), 3.0, dtype=torch.float64, device="cuda")
    a = torch.zeros((17, 512, 512), dtype=torch.float64, device="cuda")
    st = torch.full((17,), 77, dtype=torch.int32, device="cuda")
    rc = L.lpb200_gauss_jordan_batch_static_device(17, 512, a.data_ptr(), x.data_ptr(), x.data_ptr(), 0, st.data_ptr(),
                                                   stream)
    msg = L.lpb200_last_error().decode()
    assert rc == -4, msg  # LPB200_E_TOO_LARGE
    assert str(17 * 512 * 512 * 8) in msg and str(LIMIT) in msg and "LPB200_GJ_STATIC_SCRATCH_BYTES" in msg
    torch.cuda.synchronize()
    assert bool((x == 3.0).all()) and bool((st == 77).all())  # nothing was enqueued
    # the size check does not overflow for absurd shapes
    assert L.lpb200_gauss_jordan_batch_static_device(2 ** 30, 2 ** 15, a.data_ptr(), x.data_ptr(), x.data_ptr(), 0, None,
                                                     stream) == -4
    assert L.lpb200_gauss_jordan_batch_static_device(1, 4, a.data_ptr(), x.data_ptr(), x.data_ptr(), 7, None, stream) == -1


def test_a_call_allocates_no_device_memory():
    """Free device memory (cudaMemGetInfo) is the same before and after the calls.  The sequence runs once first: the
    CUDA runtime loads a kernel's code lazily at its first launch, which is not the entry point's doing."""
    import torch
    from linear_programming_b200 import _lib
    stream = torch.cuda.current_stream().cuda_stream
    cases = [_Case(*_systems(16, 512, seed=24)), _Case(*_systems(64, 33, seed=25))]

    def sequence():
        for c in cases:
            for piv in (0, 1):
                assert c.run_static(piv, stream) == 0, _lib.lib().lpb200_last_error()
        torch.cuda.synchronize()
    sequence()
    reserved_before = torch.cuda.memory_reserved()
    # cudaMemGetInfo is device-wide: something else on the box (a monitor, another process) can move it, so the
    # claim is checked up to three times and has to hold once -- an allocation by the entry point would show every time
    deltas = []
    for _ in range(3):
        free_before, _ = torch.cuda.mem_get_info()
        sequence()
        sequence()
        free_after, _ = torch.cuda.mem_get_info()
        deltas.append(free_before - free_after)
        if deltas[-1] == 0:
            break
    assert 0 in deltas, "device memory in use changed on every attempt (bytes): %s" % deltas
    assert torch.cuda.memory_reserved() == reserved_before  # nor did anything go through torch's allocator
