"""The C/C++ side of the drop-in boundary, COMPILED: the bodies INTEGRATION.md sections 2-3 show a maintainer
(allocate_ads / solve_diet_problem -> lpb200_simplex_batch, SolveEquation -> lpb200_gauss_jordan_batch) built with
g++ against include/lp_b200.h and linked with -llp_b200, then fed the reference's own fixtures through the text
protocol of the reference's main() functions (ad_allocation.cpp:21-56, diet.cc:17-58, energy_values.cc:66-80,
199-212).  The build runs everywhere (no GPU needed to link); running the binaries is a GPU test."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CXX_DIR = os.path.join(ROOT, "tests", "cxx")
LIB_DIR = os.path.join(ROOT, "linear_programming_b200")
OUT_DIR = os.path.join(CXX_DIR, "_build")


def _build(name, std):
    from linear_programming_b200 import build as lpbuild
    lpbuild.build()
    os.makedirs(OUT_DIR, exist_ok=True)
    exe = os.path.join(OUT_DIR, name)
    src = os.path.join(CXX_DIR, name + ".cpp")
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(src),
                                                               os.path.getmtime(os.path.join(LIB_DIR, "liblp_b200.so"))):
        cmd = ["g++", "-pipe", "-O2", "-std=" + std, "-Wall", "-Wextra", src, "-I", os.path.join(ROOT, "include"),
               "-L", LIB_DIR, "-llp_b200", "-Wl,-rpath," + LIB_DIR, "-o", exe]
        res = subprocess.run(cmd, capture_output=True, text=True)
        assert res.returncode == 0, res.stdout + res.stderr
    return exe


def test_consumers_compile_and_link_against_the_c_abi():
    _build("ad_allocation_b200", "c++11")   # the reference's own flags: -pipe -O2 -std=c++11 / c++0x
    _build("energy_values_b200", "c++0x")
    res = subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-c", os.path.join(CXX_DIR, "header_is_c.c"),
                          "-I", os.path.join(ROOT, "include"), "-o", os.devnull], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr


@pytest.mark.gpu
def test_ad_allocation_and_diet_fixtures_through_the_cxx_consumer(simplex_golden):
    from test_cli_stream import _expected_stdout
    exe = _build("ad_allocation_b200", "c++11")
    cases = [c for c in simplex_golden["cases"] if c.get("stdin")]
    assert len(cases) >= 30
    seen = set()
    for case in cases:
        res = subprocess.run([exe], input=case["stdin"], capture_output=True, text=True, timeout=120)
        st = case["status"]
        seen.add(st)
        if st in (-1, 0, 1):
            assert res.returncode == 0, (case["name"], res.stderr)
            assert res.stdout == _expected_stdout(case), case["name"]
        else:  # the classes in which the reference itself raises: no answer, exit code 10 + status
            assert res.returncode == 10 + st and res.stdout == "", (case["name"], res.returncode)
    assert {-1, 0, 1} <= seen


@pytest.mark.gpu
def test_energy_values_fixtures_through_the_cxx_consumer(gj_golden):
    from linear_programming_b200 import gauss_jordan
    exe = _build("energy_values_b200", "c++0x")
    n = 0
    for case in gj_golden["cases"]:
        if case.get("stdin") is None or case["name"].startswith("singular"):
            continue
        res = subprocess.run([exe], input=case["stdin"], capture_output=True, text=True, timeout=120)
        assert res.returncode == 0, (case["name"], res.stderr)
        want = gauss_jordan.format_column([float.fromhex(h) for h in case["x"]])
        assert res.stdout == want, case["name"]
        n += 1
    assert n >= 5
