"""Import alias: loads the package in `jaxopt-3d-reconstruction_b200/` (not a valid Python
identifier) under the module name `b200ba`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "jaxopt-3d-reconstruction_b200")
_spec = importlib.util.spec_from_file_location("b200ba", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["b200ba"] = _mod
_spec.loader.exec_module(_mod)
