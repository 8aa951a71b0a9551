#!/usr/bin/env python
"""Time the streaming reductions of the engine in isolation (branch lnL, theta_all + df/ddf, df/ddf, lnL from theta_all) on a DNA
GTR+G4 (or protein LG-like +G4) alignment at several pattern counts: CUDA-event time per launch (lkcuda_kernel_stats), algorithmic
bytes per launch and the fraction of the measured HBM peak.  Run it under `ncu --metrics gpu__time_duration.sum` for the
per-launch durations without the event overhead.
usage: python tools/reduce_bench.py [S=4] [reps=50] [NP,NP,...]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import iqtree_b200
from iqtree_b200 import modelinputs as mi


def peak_gbs():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        return 6542.1


def run(S, nptn, reps, lib):
    ntaxa, C = 16, 4
    SU = {4: 18, 20: 23}.get(S, S)
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, 1, names)
    st = mi.simulate_patterns(nwk, names, nptn, S, 1)
    mdl = mi.gtr_g4_inputs(1, ncat=C) if S == 4 else mi.random_reversible_inputs(S, 1, ncat=C)
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(st, np.ones(nptn), names=names, seq_type=0 if S == 4 else (1 if S == 20 else 4), nstates=S, state_unknown=SU)
    t.readTreeString(nwk)
    t.setModel(0, C, 1, mdl["eval"], mdl["evec"], mdl["inv_evec"], mdl["rates"], mdl["props"], None, mdl["state_freq"], 0.0)
    t.setParams(1e-6, 100.0)
    t.initializeAllPartialLh()
    t.clearAllPartialLH()
    lnl = t.computeLikelihood()
    ra, rb = [int(x) for x in t.rootBranch()]
    # an internal-internal branch: both ends own a vector
    ia = ib = None
    for a, b in t.getBranches():
        if t.branchVector(int(a), int(b))[0] >= 0 and t.branchVector(int(b), int(a))[0] >= 0:
            ia, ib = int(a), int(b)
            break
    out = {"S": S, "patterns": nptn, "lnl": lnl, "kernels": {}}
    peak = peak_gbs()

    def timed(name, kinds, fn, n=reps):
        for _ in range(3):
            fn()
        t.resetCounters(); t.setProfiling(True)
        for _ in range(n):
            fn()
        stats = t.kernelStats()
        t.setProfiling(False)
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        wall_us = (time.perf_counter() - t0) / n * 1e6
        launches = sum(stats[k][0] for k in kinds if k in stats)
        gb = sum(stats[k][1] for k in kinds if k in stats) / 1e9
        ms = sum(stats[k][2] for k in kinds if k in stats)
        out["kernels"][name] = {"launches_per_call": launches / n, "us_per_call": 1e3 * ms / n, "host_wall_us_per_call": wall_us, "algorithmic_mb_per_call": 1e3 * gb / n,
                                "gbps": gb / (ms * 1e-3) if ms > 0 else None, "frac_of_hbm_peak": gb / (ms * 1e-3) / peak if ms > 0 else None}

    timed("branch_lnl_tip_end", ["branch"], lambda: t.computeLikelihoodBranch(ra, rb))
    timed("theta_plus_derv_tip_end", ["theta", "derv"], lambda: t.computeLikelihoodDerv(ra, rb, reset_theta=True))
    timed("derv", ["derv"], lambda: t.computeLikelihoodDerv(ra, rb, reset_theta=False))
    timed("lnl_from_buffer", ["frombuf"], lambda: t.computeLikelihoodFromBuffer())
    if ia is not None:
        t.computeLikelihoodBranch(ia, ib)       # makes sure both directions are computed
        timed("branch_lnl_internal", ["branch"], lambda: t.computeLikelihoodBranch(ia, ib))
        timed("theta_plus_derv_internal", ["theta", "derv"], lambda: t.computeLikelihoodDerv(ia, ib, reset_theta=True))
    t.close()
    return out


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("-")]
    S = int(args[0]) if args else 4
    reps = int(args[1]) if len(args) > 1 else 50
    sizes = [int(x) for x in args[2].split(",")] if len(args) > 2 else [1000000, 125000, 8192]
    lib = iqtree_b200.load_library()
    for n in sizes:
        print(json.dumps(run(S, n, reps, lib)))


if __name__ == "__main__":
    main()
