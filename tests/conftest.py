import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mts_b200  # registers the package
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))   # oracle_lib: the CPU oracle behind the package wrappers (test infrastructure)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the native libraries once per session if they are missing."""
    import mts_b200  # noqa
    from mitsuba0_6_b200 import abi  # noqa
    import oracle_lib
    if not (os.path.exists(abi.CUDA_LIB_PATH) and os.path.exists(oracle_lib.ORACLE_LIB_PATH)):
        import __graft_entry__ as g
        g.build()
    yield
