#!/usr/bin/env python
"""SURVEY 8(f).4 measurement tool (not the headline bench): the stepwise-addition parsimony start tree
(PhyloTree::computeParsimonyTree, phylotreepars.cpp:309-426) on the device at BASELINE shape (DNA, 100 taxa), beside the UNMODIFIED
reference on the host cores (oracle/_ref/ref_harness dumppars, all threads) on a bounded sample of the same alignment.
Work unit: one candidate insertion scored on one site ("site-insertions"): sum over added taxa k of (2k-3) branches x sites.
usage: python tools/pars_bench.py [NSITES=1000000] [NTAXA=100] [reps=3] [REF_SITES=200000]"""
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import iqtree_b200
from iqtree_b200 import modelinputs as mi


def main():
    a = sys.argv[1:]
    nsites = int(a[0]) if len(a) > 0 else 1000000
    ntaxa = int(a[1]) if len(a) > 1 else 100
    reps = int(a[2]) if len(a) > 2 else 3
    ref_sites = int(a[3]) if len(a) > 3 else 200000
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, 1, names)
    st = mi.simulate_patterns(nwk, names, nsites, 4, 1)
    work = sum(2 * k - 3 for k in range(3, ntaxa))            # candidate insertions of one start tree
    lib = iqtree_b200.load_library()
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(np.zeros((ntaxa, 1), dtype=np.uint8), np.ones(1), names=names, seq_type=0, nstates=4, state_unknown=18)
    t.setParams(0.1 / nsites if nsites >= 100000 else 1e-6, 100.0)
    t0 = time.time()
    t.setParsimonySites(st, nsites)
    rng = np.random.default_rng(5)
    order = rng.permutation(ntaxa).astype(np.int32)
    score = t.computeParsimonyTree(order)                      # includes engine creation (upload + tip vectors)
    first = time.time() - t0
    walls, scores = [], [score]
    for r in range(reps):
        order = rng.permutation(ntaxa).astype(np.int32)
        t0 = time.time()
        scores.append(t.computeParsimonyTree(order))
        walls.append(time.time() - t0)
    n0, _ = t.parsLaunches()
    lib.lkpars_set_profiling(t.parsEngine(), 1)
    t.computeParsimonyTree(order)
    n1, dev_ms = t.parsLaunches()
    lib.lkpars_set_profiling(t.parsEngine(), 0)
    wall = float(np.median(walls))
    # bytes the kernels must move: a partial reads 2 and writes 1 vector, an insertion reads 3; about 2k partials are dirty per round
    vec_bytes = 4 * ((nsites + 31) // 32) * 4
    nbytes = sum((2 * k) * 3 * vec_bytes + (2 * k - 3) * 3 * vec_bytes for k in range(3, ntaxa))
    out = {"what": "parsimony_start_tree", "ntaxa": ntaxa, "nsites": nsites, "scores": scores,
           "first_call_s_with_upload": first, "wall_ms_per_tree": wall * 1e3, "launches_per_tree": n1 - n0, "device_ms_per_tree": dev_ms,
           "site_insertions_per_s": work * nsites / wall, "approx_gbps_device": nbytes / 1e9 / (dev_ms * 1e-3) if dev_ms else None}
    t.close()
    harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    if os.path.exists(harness) and ref_sites > 0:
        tmp = tempfile.mkdtemp(prefix="lk_pars_")
        ns = min(ref_sites, nsites)
        lut = np.array(list("ACGT"))
        aln = os.path.join(tmp, "sample.phy")
        with open(aln, "w") as f:
            f.write("%d %d\n" % (ntaxa, ns))
            for i, n in enumerate(names):
                f.write("%s  %s\n" % (n, "".join(lut[st[i, :ns]])))
        tre = os.path.join(tmp, "sample.nwk")
        open(tre, "w").write(nwk + "\n")
        threads = os.cpu_count() or 1
        cmd = [harness, "dumppars", os.path.join(tmp, "out"), "1", "-s", aln, "-m", "JC", "-te", tre, "-nt", str(threads), "-seed", "1", "-pre", os.path.join(tmp, "ref")]
        r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        try:
            m = json.load(open(os.path.join(tmp, "out", "metapars.json")))
            out["reference"] = {"sample_sites": ns, "informative_sites": m["nsites_informative"], "threads": m["threads"],
                                "stepwise_seconds": m["stepwise_seconds"],
                                "site_insertions_per_s": work * m["nsites_informative"] / m["stepwise_seconds"]}
            out["speedup_per_site"] = out["site_insertions_per_s"] / out["reference"]["site_insertions_per_s"]
        except Exception as ex:
            out["reference"] = {"error": str(ex), "tail": r.stdout[-500:]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
