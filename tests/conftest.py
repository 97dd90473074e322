import glob
import json
import os
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


def golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN, "*.json")))


def load_golden(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        return os.path.join(GOLDEN, name + ".txt"), json.load(f)


@pytest.fixture(scope="session")
def b2():
    """The product's Python mirror of the C-ABI (ctypes over libb2bfs.so)."""
    import importlib

    return importlib.import_module("distributed-cuda-bfs_b200")
