"""pytest configuration: the ``gpu`` marker and shared helpers.

``-m "not gpu"`` covers the oracles against the golden vectors, the host logic and the C-ABI
surface; ``-m gpu`` are the parity tests proper (CUDA path through the C ABI vs the oracle).
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")
