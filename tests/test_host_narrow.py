"""Host-core half of the end-to-end path (linear_programming_b200/csrc/host_narrow.cpp): the exact
fp64 -> int8 narrowing the second lane of lpb200_simplex_batch applies to chunks of A.  No GPU needed.
Contract: the call says 1 only when widening dst reproduces src BIT FOR BIT (so the kernel's int8 path
gives the answers of the fp64 copy); anything else -- fractions, |v| > 127, -0.0, NaN, inf -- says 0."""
import ctypes

import numpy as np
import pytest


@pytest.fixture(scope="module")
def L():
    from linear_programming_b200 import _lib, build
    build.build()
    return _lib.lib()


def _narrow(L, src):
    src = np.ascontiguousarray(src, dtype=np.float64)
    dst = np.full(src.size + 64, 77, dtype=np.int8)
    rc = L.lpb200_narrow_f64_to_i8(src.ctypes.data, dst.ctypes.data, src.size)
    assert np.all(dst[src.size:] == 77), "wrote past the end"
    return rc, dst[:src.size]


@pytest.mark.parametrize("count", [0, 1, 15, 16, 17, 31, 1000, (1 << 18) - 1, (1 << 18) + 5, 3 * (1 << 18) + 77])
def test_exact_integers_round_trip(L, count):
    rng = np.random.default_rng(count)
    src = rng.integers(-127, 128, size=count).astype(np.float64)
    rc, dst = _narrow(L, src)
    assert rc == 1
    assert np.array_equal(dst.astype(np.float64).view(np.int64), src.view(np.int64))


@pytest.mark.parametrize("bad", [0.5, -0.0, 128.0, -128.0, 127.0000000001, 1e300, -1e300, float("nan"),
                                 float("inf"), float("-inf"), 5e-324, 2147483648.0, -2147483649.0, 256.0])
@pytest.mark.parametrize("count,pos", [(1, 0), (16, 15), (40, 33), (40, 3), ((1 << 18) + 40, (1 << 18) + 35),
                                       (2 * (1 << 18), 1 << 18)])
def test_anything_else_is_refused(L, bad, count, pos):
    src = np.arange(count, dtype=np.float64) % 100
    src[pos] = bad
    rc, _ = _narrow(L, src)
    assert rc == 0


def test_bad_arguments_and_pool(L):
    assert L.lpb200_narrow_f64_to_i8(None, None, 4) == -1
    assert L.lpb200_narrow_f64_to_i8(None, None, -1) == -1
    assert L.lpb200_narrow_f64_to_i8(None, None, 0) == 1
    assert 1 <= L.lpb200_host_threads() <= 32
    out = (ctypes.c_int64 * 4)(9, 9, 9, 9)
    assert L.lpb200_simplex_batch_last_lanes(out) == 0 and list(out) == [0, 0, 0, 0]
    assert L.lpb200_simplex_batch_last_lanes(None) == -1


def test_many_calls_from_two_threads(L):
    """The pool serialises parallel_for calls; results must not mix."""
    import threading
    rng = np.random.default_rng(5)
    good = rng.integers(-100, 101, size=600_000).astype(np.float64)
    bad = good.copy()
    bad[599_999] = 0.25
    res = {}

    def work(name, arr, want):
        ok = True
        for _ in range(20):
            rc, dst = _narrow(L, arr)
            ok = ok and rc == want and (want == 0 or np.array_equal(dst.astype(np.float64), arr))
        res[name] = ok
    t1 = threading.Thread(target=work, args=("good", good, 1))
    t2 = threading.Thread(target=work, args=("bad", bad, 0))
    t1.start(); t2.start(); t1.join(); t2.join()
    assert res == {"good": True, "bad": True}


def test_forked_child_runs_without_the_pool_threads(L):
    """fork() copies the pool object but not its worker threads: the child must neither hang nor crash."""
    import os
    src = np.arange(700_000, dtype=np.float64) % 100
    assert _narrow(L, src)[0] == 1  # the pool exists in the parent
    pid = os.fork()
    if pid == 0:
        try:
            rc, dst = _narrow(L, src)
            ok = rc == 1 and np.array_equal(dst.astype(np.float64), src)
            os._exit(0 if ok else 3)
        except BaseException:
            os._exit(4)
    for _ in range(300):
        done, st = os.waitpid(pid, os.WNOHANG)
        if done:
            break
        import time
        time.sleep(0.1)
    else:
        os.kill(pid, 9)
        pytest.fail("forked child hung in the host pool")
    assert os.WIFEXITED(st) and os.WEXITSTATUS(st) == 0
