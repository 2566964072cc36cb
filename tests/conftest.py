import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: long CPU test, enabled with RTB_SLOW=1")


@pytest.fixture(scope="session")
def rtb():
    import raytracing_rust_b200 as m
    return m


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle_py as m
    return m
