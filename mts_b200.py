"""Import shim: the package directory is `mitsuba0.6_b200/` (a dot is not importable), so this
module loads it under the name `mts_b200` and re-exports its public names."""
import importlib.util as _u
import os as _os
import sys as _sys

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "mitsuba0.6_b200")
_name = "mitsuba0_6_b200"
if _name not in _sys.modules:
    _spec = _u.spec_from_file_location(_name, _os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
    _mod = _u.module_from_spec(_spec)
    _sys.modules[_name] = _mod
    _spec.loader.exec_module(_mod)
_pkg = _sys.modules[_name]
from mitsuba0_6_b200 import *  # noqa: F401,F403,E402
from mitsuba0_6_b200 import abi, hostmath, scene  # noqa: F401,E402
