import os
import sys
from pathlib import Path

# before anything loads libgomp: the oracle's OpenMP workers spin briefly between the short regions of its pivot loop, then sleep
os.environ.setdefault("GOMP_SPINCOUNT", "3000000")

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def lib():
    from uno_b200 import build, capi
    build.build()
    return capi.load_library()
