import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="session")
def pn15_stream():
    """The examples' bit stream: PN15 packed to 4096 bytes then unpacked
    (32768 bits, one trailing pad 0) -- examples/soqpsk_detection.py:28-30,45."""
    g = golden("g1_pn.npz")
    bits = np.unpackbits(g["pn15"])
    assert bits.size == 32768
    return bits
