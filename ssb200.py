"""Import alias: the package directory is named `semantic-segmentation-server_b200` (not a Python
identifier), so `import ssb200` loads it under that name and returns the package itself."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("semantic-segmentation-server_b200")
sys.modules[__name__] = _pkg
