"""Import alias: the package directory is named `gcn-drug-repurposing_b200` (hyphenated, so not a
Python identifier).  `import gdr_b200` gives the same module object."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("gcn-drug-repurposing_b200")
sys.modules[__name__] = _pkg
