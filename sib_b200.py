"""Import shim: the package directory is named `distributed-inference-system-based-storm_b200` (hyphens), which
Python cannot import by name.  `import sib_b200` loads it from that directory under this module name."""
import importlib.util
import os
import sys

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "distributed-inference-system-based-storm_b200")
_spec = importlib.util.spec_from_file_location("sib_b200", os.path.join(_DIR, "__init__.py"),
                                               submodule_search_locations=[_DIR])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["sib_b200"] = _mod
_spec.loader.exec_module(_mod)
