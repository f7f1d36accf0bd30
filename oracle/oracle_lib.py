"""TEST INFRASTRUCTURE ONLY: puts the CPU oracle (oracle/libmtsoracle.so, the restatement of the reference's path
behind the C ABI of include/mtscuda.h with the prefix `mtsoracle_`) behind the package's ctypes wrappers, so that the
parity tests feed identical inputs to the oracle and to libmtscuda.  Only tests/, __graft_entry__.smoke() and
bench.py's CPU legs import this module; nothing in mitsuba0.6_b200/ knows that it exists."""
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))

import mts_b200 as M  # noqa: E402
from mitsuba0_6_b200 import abi  # noqa: E402

ORACLE_LIB_PATH = os.path.join(_HERE, "libmtsoracle.so")
ORACLE_PREFIX = "mtsoracle_"


def lib():
    return abi.bind(ORACLE_LIB_PATH, ORACLE_PREFIX)


def OraclePath(scene):
    """The CudaPath wrapper object driving the oracle library (same preprocess / render / cancel interface)."""
    return M.CudaPath(scene, lib=lib())
