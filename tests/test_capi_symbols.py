"""CPU suite: the product C-ABI library loads without a GPU, exports every symbol include/*.h declares - and nothing of the test /
bench driver (harness/) -, and refuses to compute without a device (no CPU fallback)."""
import ctypes
import os
import re

import iqtree_b200
from iqtree_b200 import _build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared(header):
    path = os.path.join(ROOT, "include", header)
    if not os.path.exists(path):
        path = os.path.join(ROOT, "harness", header)
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lk(?:cuda|tree|pars)_[a-z0-9_]+)\s*\(", text)))


def test_cuda_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_build.build_cuda())
    names = declared("lkcuda.h") + declared("lkpars.h")
    assert len(names) > 60
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    # the product library is the engine alone: the PhyloTree mirror that drives it in tests / bench.py lives in harness/
    assert not hasattr(lib, "lktree_create")
    assert sorted(os.listdir(os.path.join(ROOT, "include"))) == ["lkcuda.h", "lkpars.h"]
    hlib = ctypes.CDLL(_build.build_harness())
    names += declared("lktree.h")
    missing = [n for n in names if not hasattr(hlib, n)]
    assert not missing, missing
    from iqtree_b200 import phylotree as pt
    bound = set(iqtree_b200.LKCUDA_SYMBOLS) | set(iqtree_b200.LKTREE_SYMBOLS) | set(pt.LKPARS_SYMBOLS)
    assert set(names) - {"lktree_allreduce_fn"} <= bound, sorted(set(names) - bound)


def test_oracle_library_exports_the_same_abi(oracle_lib):
    for n in declared("lkcuda.h") + declared("lktree.h") + declared("lkpars.h"):
        assert hasattr(oracle_lib, n), n


def test_no_cpu_fallback_without_device():
    lib = iqtree_b200.load_library()
    n = ctypes.c_int32(-1)
    rc = lib.lkcuda_device_count(ctypes.byref(n))
    if rc == 0 and n.value > 0:
        return      # running on a GPU box: nothing to check here
    import numpy as np
    from iqtree_b200 import modelinputs as mi
    from iqtree_b200 import phylotree as pt
    e = ctypes.c_void_p()
    mask = np.array([1, 2, 4, 8, 15], dtype=np.uint64)
    sites = np.zeros((3, 5), dtype=np.uint8)
    assert lib.lkpars_create(0, 4, 3, 5, sites.ctypes.data_as(pt.c_uint8_p), mask.ctypes.data_as(pt.c_uint64_p), 5, 2, ctypes.byref(e)) != 0
    assert b"no CUDA device" in lib.lkpars_last_error(None)
    t = iqtree_b200.PhyloTree(lib=lib)
    names = ["t%d" % i for i in range(4)]
    t.setAlignment(np.zeros((4, 8), dtype=np.uint8), np.ones(8), names=names)
    t.readTreeString("(t0:0.1,t1:0.1,(t2:0.1,t3:0.1):0.1);")
    m = mi.gtr_g4_inputs()
    t.setModel(0, 4, 1, m["eval"], m["evec"], m["inv_evec"], m["rates"], m["props"])
    try:
        t.computeLikelihood()
    except iqtree_b200.LkError as ex:
        assert "no CUDA device" in str(ex)
    else:
        raise AssertionError("the product path computed something without a GPU")
