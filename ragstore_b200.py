"""Import shim: the package directory is named `llm-kubernetes-minikube-sharp4dev_b200` (not a
valid Python identifier), so `import ragstore_b200` loads it from there under this name."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "llm-kubernetes-minikube-sharp4dev_b200")
_spec = importlib.util.spec_from_file_location("ragstore_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["ragstore_b200"] = _mod
_spec.loader.exec_module(_mod)
