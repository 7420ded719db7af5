"""Import shim: the package directory is named `or-tools_b200` (not an importable
identifier), so `import pdlp_b200` loads it from there under this name."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "or-tools_b200")
_spec = importlib.util.spec_from_file_location("pdlp_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_module = importlib.util.module_from_spec(_spec)
sys.modules["pdlp_b200"] = _module
_spec.loader.exec_module(_module)
