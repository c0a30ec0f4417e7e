"""Import shim: the package directory is named ``quantized.pytorch_b200`` (not a
valid dotted module name), so ``import qtorch_b200`` loads it from there."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "quantized.pytorch_b200")
_spec = importlib.util.spec_from_file_location(
    "qtorch_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["qtorch_b200"] = _mod
_spec.loader.exec_module(_mod)
