import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_rows():
    with open(os.path.join(ROOT, "tests", "golden", "strings.json"), encoding="utf-8") as f:
        return json.load(f)["rows"]


def to_csr(strings):
    """list[str] -> (uint8 data, int64 off) as the C ABI takes them."""
    import numpy as np

    enc = [s.encode("utf-8") for s in strings]
    off = np.zeros(len(enc) + 1, np.int64)
    if enc:
        off[1:] = np.cumsum([len(b) for b in enc])
    data = np.frombuffer(b"".join(enc), np.uint8).copy() if enc else np.zeros(0, np.uint8)
    return data, off
