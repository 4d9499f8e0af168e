import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    """Outputs of the reference's own kernels (oracle/_ref replay), committed by
    tests/golden/make_golden.py."""
    path = os.path.join(ROOT, "tests", "golden", "ref_vectors.npz")
    return dict(np.load(path))


@pytest.fixture(scope="session")
def golden_cases():
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_cases", os.path.join(ROOT, "tests", "golden", "make_golden.py"))
    # only the CASES dict is needed; parse it without importing oracle.ref
    src = open(spec.origin).read()
    start = src.index("CASES = {")
    end = src.index("\n}\n", start) + 3
    ns = {}
    exec(src[start:end], {"dict": dict}, ns)
    return ns["CASES"]
