"""Import shim: the package directory `4paradigm-k8s-vgpu-scheduler_b200/` is not a legal
Python module name, so it is loaded here under the name ``vgpu_b200``."""
import importlib.util
import os
import sys

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "4paradigm-k8s-vgpu-scheduler_b200")
_NAME = "vgpu_b200"

_spec = importlib.util.spec_from_file_location(_NAME, os.path.join(_DIR, "__init__.py"),
                                               submodule_search_locations=[_DIR])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[_NAME] = _mod
_spec.loader.exec_module(_mod)
