"""Import shim: exposes the package directory
``learned-compression-of-point-cloud-geometry-and-attributes_b200/`` (not a valid identifier) as
the module ``pcc_b200``."""
import importlib.util
import os
import sys

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)),
                    "learned-compression-of-point-cloud-geometry-and-attributes_b200")
_spec = importlib.util.spec_from_file_location(
    "pcc_b200", os.path.join(_DIR, "__init__.py"), submodule_search_locations=[_DIR])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["pcc_b200"] = _mod
_spec.loader.exec_module(_mod)
