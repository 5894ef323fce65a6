"""Import shim: the package directory is named ``learnable-triangulation-pytorch_b200`` (not a valid python
identifier), so this module loads it and registers it as ``lt_b200``. ``import lt_b200`` then behaves like a normal
package import (``lt_b200.op``, ``lt_b200.multiview``, ...)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "learnable-triangulation-pytorch_b200")
_spec = importlib.util.spec_from_file_location("lt_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["lt_b200"] = _mod
_spec.loader.exec_module(_mod)
