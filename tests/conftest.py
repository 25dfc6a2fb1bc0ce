import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def afk():
    """The product: ctypes view of libaflak_b200.so (built on demand, never a fallback)."""
    import aflak_b200  # builds libaflak_b200.so with nvcc when missing or stale; raises if that fails

    return aflak_b200


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure)."""
    import oracle

    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def gpu(afk):
    """GPU tests must not pass on a silent fallback: no device is a failure, not a skip."""
    n = afk.device_count()
    assert n > 0, "no CUDA device visible: -m gpu tests need the B200 box"
    afk.set_device(0)
    return afk
