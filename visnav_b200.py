"""Import shim: `import visnav_b200` loads the package in ./visual-slam_b200/ (hyphenated dir)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "visual-slam_b200")
_spec = importlib.util.spec_from_file_location(
    "visnav_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["visnav_b200"] = _mod
_spec.loader.exec_module(_mod)
