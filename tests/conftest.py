import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


@pytest.fixture(scope="session")
def renderer():
    """One library instance on cuda:0 for the whole GPU session (fails loudly without a GPU)."""
    import rtlib
    r = rtlib.pkg.Renderer(1)
    yield r
    r.close()
