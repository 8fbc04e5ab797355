import gzip
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    """Reference-generated vectors (tests/golden/make_golden.py)."""
    with gzip.open(os.path.join(ROOT, "tests", "golden", "sw_golden.json.gz"), "rb") as f:
        return json.loads(f.read().decode())


@pytest.fixture(scope="session")
def oracle():
    from oracle import sw_oracle
    sw_oracle.build()
    return sw_oracle
