"""B200-native Felsenstein-pruning likelihood engine for IQ-TREE's PhyloTree kernel API.

Layout:
  csrc/            hand-written sm_100a kernels + the C ABI (include/lkcuda.h) + the host PhyloTree mirror (include/lktree.h)
  phylotree.py     ctypes binding with the reference's method names
  modelinputs.py   producers of the input contract (eigen-systems, Gamma rates, synthetic patterns) for benches/tests
  _build.py        in-tree nvcc / gcc build recipes

The directory name contains a hyphen; import it through the repo-root alias module `iqtree_b200`.
"""
from . import _build, modelinputs
from .phylotree import (PhyloTree, LkError, load_library, LKCUDA_SYMBOLS, LKTREE_SYMBOLS, KIND_NAMES,
                        FAMILY_PLAIN, FAMILY_MIXRATE, FAMILY_MIXTURE, SEQ_DNA, SEQ_PROTEIN, SEQ_BINARY, SEQ_MORPH,
                        SEQ_MULTISTATE, SEQ_CODON, WSL_RATECAT, WSL_MIXTURE, WSL_MIXTURE_RATECAT)

__all__ = ["PhyloTree", "LkError", "load_library", "modelinputs", "_build"]
