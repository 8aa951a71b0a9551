"""TEST INFRASTRUCTURE: loader of the CPU oracle library (host mirror + oracle/lk_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
The product path (iq-tree_b200/) never does.
"""
import ctypes
import os
import sys

_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _root not in sys.path:
    sys.path.insert(0, _root)
import iqtree_b200  # noqa: E402
from iqtree_b200 import _build  # noqa: E402
from iqtree_b200.phylotree import _bind, PhyloTree  # noqa: E402

_lib = None


def load_oracle_library():
    global _lib
    if _lib is None:
        _lib = _bind(ctypes.CDLL(_build.build_oracle()))
    return _lib


def OraclePhyloTree():
    """The same host mirror, computing on the CPU with the scalar restatement of the reference kernels."""
    return PhyloTree(lib=load_oracle_library())
