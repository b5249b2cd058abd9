import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a); run with -m gpu")


@pytest.fixture(scope="session")
def pkg():
    from lgattn_pkg import pkg as _pkg
    return _pkg


@pytest.fixture(scope="session")
def golden_dir():
    return ROOT / "tests" / "golden"
