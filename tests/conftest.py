import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def simplex_golden():
    with open(os.path.join(GOLDEN_DIR, "simplex_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def gj_golden():
    with open(os.path.join(GOLDEN_DIR, "gauss_jordan_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def oracle_lib():
    import oracle
    oracle.build()
    return oracle
