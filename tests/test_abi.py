"""C-ABI checks that need no GPU: the library builds/loads and exports exactly what
include/lp_b200.h declares; without a device every compute call fails loudly (no CPU fallback)."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def engine_lib():
    from linear_programming_b200 import _lib, build
    build.build()
    return _lib


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "lp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lpb200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound(engine_lib):
    syms = _declared_symbols()
    assert len(syms) >= 12
    L = engine_lib.lib()
    for s in syms:
        assert hasattr(L, s), "liblp_b200.so does not export " + s
    assert sorted(engine_lib.SIGNATURES) == syms


def test_header_cites_reference_interfaces():
    text = open(os.path.join(ROOT, "include", "lp_b200.h")).read()
    for cite in ("simplex.py:496-505", "ad_allocation.cpp:9-19", "diet.cc:10-15", "energy_values.cc:166-197"):
        assert cite in text


def test_version_and_shape_queries(engine_lib):
    L = engine_lib.lib()
    assert b"sm_100a" in L.lpb200_version()
    assert 80000 < L.lpb200_simplex_batch_smem_bytes(100, 100) <= 113 * 1024   # 2 CTAs per SM
    assert L.lpb200_simplex_batch_smem_bytes(400, 400) > 227 * 1024
    assert L.lpb200_simplex_single_workspace_bytes(16383, 32767) > 16384 * 32768 * 8
    assert L.lpb200_gauss_jordan_workspace_bytes(4096, 512) == 4096 * 512 * 512 * 8


def test_no_cpu_fallback_without_device(engine_lib):
    import linear_programming_b200 as lp
    if engine_lib.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(engine_lib.EngineError) as ei:
        lp.solve([[1.0, 1.0]], [2.0], [1.0, 1.0])
    assert ei.value.code == -3
    with pytest.raises(engine_lib.EngineError):
        lp.solve_equation_batch(np.eye(2)[None], np.ones((1, 2)))


def test_invalid_arguments_are_reported(engine_lib):
    L = engine_lib.lib()
    rc = L.lpb200_simplex_batch(1, 0, 3, None, None, None, 1e-9, 10, None, None, None, None, None, 0)
    assert rc == -1 and b"n >= 1" in L.lpb200_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "linear_programming_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "lp_oracle" not in src, f
