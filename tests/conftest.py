import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ptb():
    """One C-ABI instance per test session (the library keeps per-process state, like one cl::Context)."""
    from opencl_path_tracer_b200 import capi
    p = capi.Ptb(0)
    yield p
    p.shutdown()
