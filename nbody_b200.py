"""Import alias: the package directory is `n-bodysimulation_b200/` (not a valid
identifier), so `import nbody_b200` loads it by path-string instead."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
sys.modules[__name__] = importlib.import_module("n-bodysimulation_b200")
