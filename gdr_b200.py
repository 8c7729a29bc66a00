"""Import alias: the package directory is named `gcn-drug-repurposing_b200` (hyphenated, so not a
Python identifier).  `import gdr_b200` gives the same module object."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_PKG = "gcn-drug-repurposing_b200"
_pkg = importlib.import_module(_PKG)
for _sub in ("trainer", "dist"):          # not imported by the package __init__
    importlib.import_module(f"{_PKG}.{_sub}")
# one module object per submodule under both names (a second import under the alias would otherwise
# create duplicate classes with separate state)
for _name, _mod in list(sys.modules.items()):
    if _name.startswith(_PKG + "."):
        sys.modules[__name__ + _name[len(_PKG):]] = _mod
sys.modules[__name__] = _pkg
