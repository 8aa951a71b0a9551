"""Importable alias of the package directory `iq-tree_b200/` (a hyphen is not a valid module name)."""
import importlib.util
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location(
    "iqtree_b200", os.path.join(_here, "iq-tree_b200", "__init__.py"),
    submodule_search_locations=[os.path.join(_here, "iq-tree_b200")])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["iqtree_b200"] = _mod
_spec.loader.exec_module(_mod)
