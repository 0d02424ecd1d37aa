import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (REPO, os.path.join(REPO, "mitsuba3-sunsky_b200"), os.path.join(REPO, "oracle"), os.path.dirname(__file__)):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: long-running statistical test")


@pytest.fixture(scope="session")
def ref():
    """The reference's own sunsky.cpp compiled in place (oracle/_ref). Skips when it was never built."""
    import ref_oracle
    if not ref_oracle.available():
        pytest.skip("oracle/_ref/libsunsky_ref.so not built (run oracle/build_ref.sh where /root/reference exists)")
    return ref_oracle


@pytest.fixture(scope="session")
def ssk():
    import sunsky_b200
    return sunsky_b200
