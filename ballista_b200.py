"""Importable alias for the package directory `arrow-ballista_b200/` (a hyphen is not a valid module name)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("arrow-ballista_b200")
sys.modules[__name__] = _pkg
# submodules too, so that `ballista_b200.expr` and the package's own `.expr` are one module (one set of classes)
for _name, _mod in list(sys.modules.items()):
    if _name.startswith("arrow-ballista_b200."):
        sys.modules["ballista_b200." + _name.split(".", 1)[1]] = _mod
