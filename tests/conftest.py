import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


_shared = {"r": None}


@pytest.fixture
def renderer():
    """The library instance on cuda:0 shared by the GPU tests (fails loudly without a GPU).  The library state is
    process-global: a test module that makes and closes its own Renderer shuts it down, so it is re-initialised here
    when that has happened, and the modes are put back to their defaults."""
    import rtlib
    r = _shared["r"]
    if r is None or r.lib.rt_device_count() == 0:
        r = rtlib.pkg.Renderer(1)
        _shared["r"] = r
    r.set_mode(0)
    r.set_accel(0)
    r.set_max_depth(7)
    yield r


@pytest.fixture(scope="session", autouse=True)
def _close_shared_renderer():
    yield
    if _shared["r"] is not None and _shared["r"].lib.rt_device_count() != 0:
        _shared["r"].close()
