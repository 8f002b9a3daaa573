"""The reference's differential stress loop (week-2/diet/stress.py) as a GPU-batched parity gate."""
import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu


def test_stress_distribution_30k_lps_bit_exact():
    from fuzz_stress import run
    out = run(count=30000, seed=3, nmax=7, scipy_sample=100)
    assert out["engine_vs_checker_mismatches"] == 0, out
    assert out["shapes"] == 49
    # every outcome class of SURVEY A.7 shows up in a run of this size
    assert {"-1", "0", "1", "3"} <= set(out["status_counts"]), out["status_counts"]


def test_wider_shapes_bit_exact():
    from fuzz_stress import run
    out = run(count=6000, seed=5, nmax=24, scipy_sample=0)
    assert out["engine_vs_checker_mismatches"] == 0, out
