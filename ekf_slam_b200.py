"""Import shim: exposes the package directory `ekf-slam_b200/` (not a valid Python identifier)
under the module name `ekf_slam_b200`."""
import importlib.util as _ilu
import os as _os
import sys as _sys

_root = _os.path.dirname(_os.path.abspath(__file__))
_pkg = _os.path.join(_root, "ekf-slam_b200")
_spec = _ilu.spec_from_file_location("ekf_slam_b200", _os.path.join(_pkg, "__init__.py"),
                                     submodule_search_locations=[_pkg])
_mod = _ilu.module_from_spec(_spec)
_sys.modules["ekf_slam_b200"] = _mod
_spec.loader.exec_module(_mod)
