import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mts_b200  # registers the package
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the native libraries once per session if they are missing."""
    from mitsuba0_6_b200 import abi  # noqa
    import mts_b200  # noqa
    if not (os.path.exists(abi.LIB_PATHS["cuda"]) and os.path.exists(abi.LIB_PATHS["oracle"])):
        import __graft_entry__ as g
        g.build()
    yield
