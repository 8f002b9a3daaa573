"""The oracle's CPython-set model vs the real builtin `set` of this interpreter.

Anchors the entering order of week-2/diet/simplex.py:390-392 (iteration of
`self.independent`) and the re-substitution order of :538 (`aux_sf.dependent`).
"""
import random
import sys

import pytest


def _replay_real(lo, hi, ops):
    s = set(range(lo, hi))
    for rem, add in ops:
        if rem >= 0:
            s.remove(rem)
        if add >= 0:
            s.add(add)
    extra = sys.getsizeof(s) - sys.getsizeof(set())   # 0 while the embedded 8-slot table is in use
    return list(s), (extra // 16 if extra else 8)


def _pivot_like_ops(rng, lo, hi, universe, steps):
    """remove a member, add a non-member -- what SlackForm.pivot does to each set (:369-381)."""
    inside = list(range(lo, hi))
    outside = [k for k in range(universe) if not (lo <= k < hi)]
    ops = []
    for _ in range(steps):
        if not inside or not outside:
            break
        i = rng.randrange(len(inside))
        j = rng.randrange(len(outside))
        rem, add = inside[i], outside[j]
        inside[i], outside[j] = add, rem
        ops.append((rem, add))
    return ops


@pytest.mark.parametrize("seed", range(6))
def test_set_order_matches_cpython(oracle_lib, seed):
    rng = random.Random(seed)
    for _ in range(250):
        n = rng.randint(1, 130)
        m = rng.randint(1, 130)
        # independent = set(range(n, n+m)) mutated inside a universe of n+m (+1 aux) ids
        ops = _pivot_like_ops(rng, n, n + m, n + m + 1, rng.randint(0, 400))
        want, want_size = _replay_real(n, n + m, ops)
        got, got_size = oracle_lib.pyset_replay(n, n + m, ops)
        assert got == want
        assert got_size == want_size
        # dependent = set(range(0, n))
        ops = _pivot_like_ops(rng, 0, n, n + m + 1, rng.randint(0, 400))
        want, want_size = _replay_real(0, n, ops)
        got, got_size = oracle_lib.pyset_replay(0, n, ops)
        assert got == want
        assert got_size == want_size


def test_set_order_long_run_with_rebuilds(oracle_lib):
    """100x100 shape: 512-slot table that is rebuilt every ~200 fresh-slot adds."""
    rng = random.Random(99)
    ops = _pivot_like_ops(rng, 100, 200, 201, 6000)
    want, want_size = _replay_real(100, 200, ops)
    got, got_size = oracle_lib.pyset_replay(100, 200, ops)
    assert got == want and got_size == want_size == 512


def test_remove_only_then_order(oracle_lib):
    # aux_sf.independent.remove(nvars), simplex.py:574
    want, _ = _replay_real(5, 12, [(11, -1)])
    got, _ = oracle_lib.pyset_replay(5, 12, [(11, -1)])
    assert got == want
