"""pytest config: registers the ``gpu`` marker and puts the repo root on sys.path.

``-m "not gpu"`` runs on the CPU-only build container (oracle vs golden vectors,
host logic, C-ABI symbol check, gloo world_size-2 sharding); ``-m gpu`` runs on a
B200 and calls the CUDA path through the C-ABI library."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
