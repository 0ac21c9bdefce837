import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA (sm_100a) device; run with -m gpu on the B200 box")


def pytest_collection_modifyitems(config, items):
    # keep the allocation-free entry point's test (tests/test_zz_*.py) last whatever the collection order of plugins
    items.sort(key=lambda it: (os.path.basename(str(it.fspath)).startswith("test_zz_"),))


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
