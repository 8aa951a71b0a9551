#!/usr/bin/env python
"""Coverage / identity probe of the whole-traversal kernels: a fresh, NaN-poisoned engine (LKCUDA_POISON=1) runs ONE full traversal of a
synthetic DNA shape; every arena vector is then checked on the device for NaN entries and folded into an integer digest (partial_lh as
int64 words summed with wrap-around, scale_num summed).  Two builds of the library that compute the same bits print the same digest;
a pattern no kernel wrote, or one computed from unwritten data, shows up as NaN rows.
usage: LKCUDA_POISON=1 python tools/trv_probe.py NTAXA NPTN [runs=3] [NCAT=4]"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("LKCUDA_POISON", "1")
import iqtree_b200
from iqtree_b200 import modelinputs as mi


class _DevArray:
    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def probe(lib, names, nwk, st, mdl, ncat):
    """one fresh engine, one full traversal -> {"lnl", "vectors", "nan_rows", "digest", "bad"}"""
    import torch
    nptn = st.shape[1]
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(st, np.ones(nptn), names=names, seq_type=0, nstates=4, state_unknown=18)
    t.readTreeString(nwk)
    t.setModel(0, ncat, 1, mdl["eval"], mdl["evec"], mdl["inv_evec"], mdl["rates"], mdl["props"], None, mdl["state_freq"], 0.0)
    t.setParams(0.1 / nptn if nptn >= 100000 else 1e-6, 100.0)
    t.initializeAllPartialLh()
    t.clearAllPartialLH()
    lnl = t.computeLikelihood()
    e = t.engine(0)
    B = 4 * ncat
    digest, nan_rows, nvec, bad = 0, 0, 0, []
    for a, b in list(t.getBranches()) + [(y, x) for x, y in t.getBranches()]:
        vec, comp = t.branchVector(int(a), int(b))
        if vec < 0 or not comp:
            continue
        lhp, scp = C.c_void_p(), C.c_void_p()
        assert lib.lkcuda_vector_device_ptr(e, vec, C.byref(lhp), C.byref(scp)) == 0
        lh = torch.as_tensor(_DevArray(lhp.value, nptn * B, "<f8"), device="cuda:0")
        sc = torch.as_tensor(_DevArray(scp.value, nptn, "<i2"), device="cuda:0")
        isn = torch.isnan(lh).view(nptn, B).any(dim=1)
        rows = int(isn.sum().item())
        if rows:
            bad.append([int(a), int(b), vec, rows, int(isn.nonzero()[0].item())])
        nan_rows += rows
        digest = (digest * 1000003 + int(lh.view(torch.int64).sum().item()) + 7 * int(sc.to(torch.int64).sum().item())) & ((1 << 63) - 1)
        nvec += 1
    t.close()
    return {"lnl": lnl, "vectors": nvec, "nan_rows": nan_rows, "digest": "%016x" % digest, "bad": bad[:6]}


def shape(ntaxa, nptn, ncat=4):
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, 1, names)
    return names, nwk, mi.simulate_patterns(nwk, names, nptn, 4, 1), mi.gtr_g4_inputs(1, ncat=ncat)


def main():
    ntaxa, nptn = int(sys.argv[1]), int(sys.argv[2])
    runs = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    ncat = int(sys.argv[4]) if len(sys.argv) > 4 else 4
    names, nwk, st, mdl = shape(ntaxa, nptn, ncat)
    lib = iqtree_b200.load_library()
    for run in range(runs):
        r = probe(lib, names, nwk, st, mdl, ncat)
        r.update({"run": run, "shape": [ntaxa, nptn, ncat]})
        print(json.dumps(r))


if __name__ == "__main__":
    main()
