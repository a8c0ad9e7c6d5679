import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def audit_golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "audit_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def liability_golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "liability_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def slashing_golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "slashing_golden.json")) as f:
        return json.load(f)
