"""Import shim: loads the package directory `probability-of-collision-for-safe-planning_b200/`
(whose name is not a Python identifier) under the module name `pcoll_b200`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "probability-of-collision-for-safe-planning_b200")
_spec = importlib.util.spec_from_file_location(
    "pcoll_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["pcoll_b200"] = _mod
_spec.loader.exec_module(_mod)
