import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    """The product package; the library must already be built (make / __graft_entry__.build())."""
    import __graft_entry__ as ge
    if not os.path.exists(os.path.join(ROOT, "ray-tracer-challenge_b200", "librtc_b200.so")):
        ge.build()
    return importlib.import_module("ray-tracer-challenge_b200")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle_py
    oracle_py.lib()
    return oracle_py
