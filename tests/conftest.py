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


def pytest_sessionfinish(session, exitstatus):
    """Under the self-checking build (RTB200_LIB = a -DRTB_DEBUG_CHECKS=1 library, profiles/r2_sanitizer.txt) every kernel
    bounds-checks the indices it forms; after a -m gpu run report the device counters and fail the session if any is set."""
    import os
    if "checks" not in os.environ.get("RTB200_LIB", ""):
        return
    try:
        import raytracing_rust_b200 as m
        if m.device_count() == 0:
            return
        c = m.check_failures()
    except Exception as e:  # a release build answers with an error: nothing to report
        print(f"\n[rtb200 self-checks] not available: {e}")
        return
    print(f"\n[rtb200 self-checks] device counters after the session: {c}")
    if any(c.values()):
        session.exitstatus = 1
