"""Parity AT THE HEADLINE SHAPE against vectors the UNMODIFIED reference produced (tests/golden/make_bench_golden.py):
the first 96 / 6 / 6 problems of the D1 / D2 / D3 batches of 100 x 100 LPs that bench.py solves (BASELINE.json
configs[3]).  Status, the full pivot sequence (up to 1 700 pivots per LP), every bit of x and of the optimum.
CPU: the oracle (so that the checker the bench uses on all 65 536 LPs is itself pinned at this shape);
GPU: the batched kernel (product kernel and the opt-in variants) and the grid-wide kernel through its pipelined loop."""
import json
import os
import zlib

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from linear_programming_b200 import workloads


@pytest.fixture(scope="module")
def bench_golden():
    with open(os.path.join(GOLDEN_DIR, "bench_shape_golden.json")) as f:
        return json.load(f)


def _inputs(dist, count):
    return workloads.simplex_batch(dist, count, 100, 100, seed=2)


def _crc(A, b, c):
    return zlib.crc32(np.ascontiguousarray(c).tobytes(),
                      zlib.crc32(np.ascontiguousarray(b).tobytes(), zlib.crc32(np.ascontiguousarray(A).tobytes())))


def _by_dist(golden):
    out = {}
    for c in golden["cases"]:
        out.setdefault(c["dist"], []).append(c)
    for lst in out.values():
        lst.sort(key=lambda c: c["index"])
        assert [c["index"] for c in lst] == list(range(len(lst)))
    return out


def test_fixture_is_the_bench_batch(bench_golden):
    groups = _by_dist(bench_golden)
    assert {d: len(v) for d, v in groups.items()} == {"D1": 96, "D2": 6, "D3": 6}
    for dist, cases in groups.items():
        A, b, c = _inputs(dist, len(cases))
        for k in cases:
            assert _crc(A[k["index"]], b[k["index"]], c[k["index"]]) == k["input_crc"], (dist, k["index"])
    assert sum(k["npivots"] for k in bench_golden["cases"]) > 15000
    assert min(k["npivots"] for k in groups["D2"]) > 500 and max(k["npivots"] for k in groups["D3"]) > 1500


def test_oracle_at_the_headline_shape(oracle_lib, bench_golden):
    for dist, cases in _by_dist(bench_golden).items():
        A, b, c = _inputs(dist, len(cases))
        for k in cases:
            i = k["index"]
            r = oracle_lib.simplex(A[i], b[i], c[i], max_pivots=100000)
            assert r["status"] == k["status"], (dist, i)
            assert [list(p) for p in r["pivots"]] == k["pivots"], (dist, i)
            if k["status"] == 0:
                assert [float(v).hex() for v in r["x"]] == k["x"], (dist, i)
                assert float(r["objective"]).hex() == k["optimum"], (dist, i)


def _check_result(res, p, k, name):
    assert int(res.status[p]) == k["status"], name
    assert int(res.npivots[p]) == k["npivots"], name
    assert [list(q) for q in res.pivots(p)] == k["pivots"], name
    if k["status"] == 0:
        assert [float(v).hex() for v in np.asarray(res.x[p])] == k["x"], name
        assert float(res.objective[p]).hex() == k["optimum"], name


@pytest.mark.gpu
@pytest.mark.parametrize("variant", ["default", "w", "reg"])
def test_batched_kernel_at_the_headline_shape(bench_golden, monkeypatch, variant):
    import linear_programming_b200 as lp
    if variant != "default":
        monkeypatch.setenv("LPB_BATCH_KERNEL", variant)
    for dist, cases in _by_dist(bench_golden).items():
        A, b, c = _inputs(dist, len(cases))
        cap = max(k["npivots"] for k in cases) + 4
        res = lp.solve_batch(A, b, c, max_pivots=cap, log_cap=cap, mode="batch")
        for k in cases:
            _check_result(res, k["index"], k, "%s %s/%d" % (variant, dist, k["index"]))


@pytest.mark.gpu
def test_grid_wide_kernel_at_the_headline_shape(bench_golden, monkeypatch):
    """One LP per launch on the grid-wide kernel with 24 CTAs forced (control group + sweepers: the pipelined loop),
    on one rank and on 4 emulated column-sharded ranks."""
    import linear_programming_b200 as lp
    from linear_programming_b200.sharded import solve_sharded_emulated
    monkeypatch.setenv("LPB_DEBUG_SINGLE_CPR", "24")
    groups = _by_dist(bench_golden)
    picks = [("D1", i) for i in range(0, 96, 12)] + [("D2", 0), ("D2", 3), ("D3", 0), ("D3", 4)]
    for dist, i in picks:
        k = groups[dist][i]
        A, b, c = _inputs(dist, i + 1)
        cap = k["npivots"] + 4
        res = lp.solve_batch(A[i:i + 1], b[i:i + 1], c[i:i + 1], mode="single", max_pivots=cap, log_cap=cap)
        _check_result(res, 0, k, "single %s/%d" % (dist, i))
        if dist != "D1" and i == 0:
            sh = solve_sharded_emulated(A[i], b[i], c[i], 4, max_pivots=cap, log_cap=cap)
            for g in range(4):
                _check_result(sh, g, k, "4 ranks, rank %d %s/%d" % (g, dist, i))


def test_oracle_on_the_config5_generator(oracle_lib, bench_golden):
    """workloads.single_lp (the generator of BASELINE.json configs[4]) at the sizes the Python reference can hold."""
    for k in bench_golden["single_lp_cases"]:
        A, b, c = workloads.single_lp(k["n"], k["m"], seed=5)
        assert _crc(A, b, c) == k["input_crc"]
        r = oracle_lib.simplex(A, b, c, max_pivots=100000)
        assert r["status"] == k["status"] and [list(p) for p in r["pivots"]] == k["pivots"]
        assert [float(v).hex() for v in r["x"]] == k["x"] and float(r["objective"]).hex() == k["optimum"]


@pytest.mark.gpu
def test_grid_wide_and_sharded_kernel_on_the_config5_generator(bench_golden):
    """The grid-wide kernel (pipelined loop, default grid) and 2 / 8 emulated column-sharded ranks against the reference's
    own answers on config 5's generator: status, pivot sequence, x and optimum bits."""
    import linear_programming_b200 as lp
    from linear_programming_b200.sharded import solve_sharded_emulated
    for k in bench_golden["single_lp_cases"]:
        A, b, c = workloads.single_lp(k["n"], k["m"], seed=5)
        cap = k["npivots"] + 4
        res = lp.solve_batch(A[None], b[None], c[None], mode="single", max_pivots=cap, log_cap=cap)
        _check_result(res, 0, k, "single %dx%d" % (k["n"], k["m"]))
        for nranks in (2, 8):
            sh = solve_sharded_emulated(A, b, c, nranks, max_pivots=cap, log_cap=cap)
            for g in range(nranks):
                _check_result(sh, g, k, "%d ranks, rank %d, %dx%d" % (nranks, g, k["n"], k["m"]))
