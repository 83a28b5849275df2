"""Import alias: the package directory `svt-av1_pro-anchor-v2.1.0-_b200/` is not a valid Python
identifier, so this module loads it and registers it as `svtfilt_b200`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "svt-av1_pro-anchor-v2.1.0-_b200")
_spec = importlib.util.spec_from_file_location("svtfilt_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["svtfilt_b200"] = _mod
_spec.loader.exec_module(_mod)
