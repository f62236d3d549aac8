"""Import shim: the package directory is named `henjou-renderer_b200/` (not a Python identifier).
`import henjou_renderer_b200` loads that directory as the package of this name."""
import importlib.util
import os
import sys

_d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "henjou-renderer_b200")
_spec = importlib.util.spec_from_file_location(
    "henjou_renderer_b200", os.path.join(_d, "__init__.py"), submodule_search_locations=[_d])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["henjou_renderer_b200"] = _mod
_spec.loader.exec_module(_mod)
