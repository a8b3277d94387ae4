import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (REPO, os.path.join(REPO, "scenes"), os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "ref: needs /root/reference (only this container)")


@pytest.fixture(scope="session")
def ctx():
    """One CUDA context for the whole GPU session (stats build: counters on)."""
    import fs2_b200
    c = fs2_b200.Context(device=0, stats=True)
    yield c
    c.close()
