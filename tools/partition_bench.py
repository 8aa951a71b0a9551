#!/usr/bin/env python
"""BASELINE configs[4]: a 50-partition PartitionModel (25 DNA GTR+G4 + 25 protein LG+G4, NPTN patterns each, NTAXA taxa): the sum of
the partition trees' lnL, each partition an independent engine on its own stream (PhyloSuperTree::computeLikelihood,
phylosupertree.cpp:955-980).  Times one full evaluation (clearAllPartialLH + computeLikelihood of every partition) driven
sequentially and from a thread pool (the reference's `#pragma omp parallel for` over partitions).
LK_PART_GPUS=n spreads WHOLE PARTITIONS over n GPUs of the box (the second sharding axis: independent PhyloTree handles, the
host sums the partition likelihoods, phylosupertree.cpp:972-976) - DNA / protein pairs go to the same device so that the load is even.
usage: python tools/partition_bench.py [NTAXA] [NPTN] [NPART] [reps]"""
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import iqtree_b200
from iqtree_b200 import modelinputs as mi


def make_part(lib, ntaxa, nptn, S, seed, device=0):
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, 1, names)        # edge-linked partitions share the topology
    st = mi.simulate_patterns(nwk, names, nptn, S, seed)
    mdl = mi.gtr_g4_inputs(seed, ncat=4) if S == 4 else mi.random_reversible_inputs(S, seed, ncat=4)
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(st, np.ones(nptn), names=names, seq_type=0 if S == 4 else 1, nstates=S, state_unknown=18 if S == 4 else 23)
    t.readTreeString(nwk)
    t.setModel(0, 4, 1, mdl["eval"], mdl["evec"], mdl["inv_evec"], mdl["rates"], mdl["props"], None, mdl["state_freq"], 0.0)
    t.setParams(1e-6, 100.0)
    t.setDevices([device])
    t.initializeAllPartialLh()
    return t


def main():
    a = [int(x) for x in sys.argv[1:5]]
    ntaxa, nptn, npart, reps = (a + [100, 20000, 50, 5][len(a):])[:4]
    lib = iqtree_b200.load_library()
    ngpu = int(os.environ.get("LK_PART_GPUS", 1))
    parts = [make_part(lib, ntaxa, nptn, 4 if i % 2 == 0 else 20, 10 + i, (i // 2) % ngpu) for i in range(npart)]

    def one(t):
        t.clearAllPartialLH()
        return t.computeLikelihood()

    ref = sum(one(t) for t in parts)
    t0 = time.time()
    for _ in range(reps):
        s = sum(one(t) for t in parts)
    seq = (time.time() - t0) / reps
    assert s == ref
    out = {"gpus": ngpu, "partitions": npart, "ntaxa": ntaxa, "patterns_per_partition": nptn, "lnl": ref, "sequential_ms": seq * 1e3}
    for nthr in (4, 16, 32):
        with ThreadPoolExecutor(nthr) as ex:
            list(ex.map(one, parts))
            t0 = time.time()
            for _ in range(reps):
                s = sum(ex.map(one, parts))
            out["threads%d_ms" % nthr] = (time.time() - t0) / reps * 1e3
        assert s == ref
    updates = (ntaxa - 2) * nptn * 4 * npart
    best = min(v for k, v in out.items() if k.endswith("_ms"))
    out["updates_per_s"] = updates / (best * 1e-3)
    out["mem_gb"] = sum(t.memoryBytes() for t in parts) / 1e9
    print(json.dumps(out))
    for t in parts:
        t.close()


if __name__ == "__main__":
    main()
