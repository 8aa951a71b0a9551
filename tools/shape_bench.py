#!/usr/bin/env python
"""Time full-tree lnL / df-ddf / optimizeAllBranches(1) of the CUDA engine on a synthetic shape (tuning tool, not the headline bench).
usage: python tools/shape_bench.py NTAXA NPTN NSTATES NCAT [reps] [--oab] [--mix=M]   (--mix: M-component profile mixture, family 2)"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import iqtree_b200
from iqtree_b200 import modelinputs as mi


def main():
    ntaxa, nptn, S, C = [int(x) for x in sys.argv[1:5]]
    reps = int(sys.argv[5]) if len(sys.argv) > 5 and not sys.argv[5].startswith("-") else 5
    oab = "--oab" in sys.argv
    nmix = 1
    for a in sys.argv:
        if a.startswith("--mix="):
            nmix = int(a[6:])
    SU = {4: 18, 20: 23}.get(S, S)
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, 1, names)
    st = mi.simulate_patterns(nwk, names, nptn, S, 1)
    mdl = mi.gtr_g4_inputs(1, ncat=C) if S == 4 else mi.random_reversible_inputs(S, 1, ncat=C)
    lib = iqtree_b200.load_library()
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(st, np.ones(nptn), names=names, seq_type=0 if S == 4 else (1 if S == 20 else 4), nstates=S, state_unknown=SU)
    t.readTreeString(nwk)
    if nmix > 1:
        comps = [mi.random_reversible_inputs(S, 100 + m, ncat=C) for m in range(nmix)]
        t.setModel(2, C, nmix, np.concatenate([c["eval"].ravel() for c in comps]), np.concatenate([c["evec"].ravel() for c in comps]),
                   np.concatenate([c["inv_evec"].ravel() for c in comps]), mdl["rates"], mdl["props"], np.full(nmix, 1.0 / nmix), mdl["state_freq"], 0.0)
    else:
        t.setModel(0, C, 1, mdl["eval"], mdl["evec"], mdl["inv_evec"], mdl["rates"], mdl["props"], None, mdl["state_freq"], 0.0)
    t.setParams(0.1 / nptn if nptn >= 100000 else 1e-6, 100.0, device_newton="--devnewton" in sys.argv)
    t.initializeAllPartialLh()
    for _ in range(2):
        t.clearAllPartialLH(); lnl = t.computeLikelihood()
    t.resetCounters(); t.setProfiling(True)
    t0 = time.time()
    for _ in range(reps):
        t.clearAllPartialLH(); lnl = t.computeLikelihood()
    dt = (time.time() - t0) / reps
    stats = t.kernelStats()
    trav = t.getTraversal()
    a, b = int(trav[len(trav) // 2][0]), int(trav[len(trav) // 2][1])
    t.computeLikelihoodDerv(a, b, reset_theta=True)
    t0 = time.time()
    for _ in range(reps * 4):
        d = t.computeLikelihoodDerv(a, b, reset_theta=False)
    dd = (time.time() - t0) / (reps * 4)
    out = {"shape": [ntaxa, nptn, S, C, nmix], "lnl": lnl, "lnl_ms": dt * 1e3, "updates_per_s": (ntaxa - 2) * nptn * C * nmix / dt,
           "derv_ms": dd * 1e3, "mem_gb": t.memoryBytes() / 1e9,
           "kernels": {k: {"n": v[0], "ms": v[2], "alg_gbps": (v[1] / 1e9) / (v[2] * 1e-3) if v[2] > 0 else None} for k, v in stats.items() if v[0]}}
    flops = 0
    for k_ in ("partial",):
        pass
    out["eff_tflops_internal_internal_equiv"] = (ntaxa - 2) * nptn * C * nmix * (6 * S * S + S) / dt / 1e12
    if oab:
        t.clearAllPartialLH(); t.computeLikelihood()
        t0 = time.time(); l1 = t.optimizeAllBranches(1); out["oab1_s"] = time.time() - t0; out["oab1_lnl"] = l1
        t.resetCounters()
        t0 = time.time(); l2 = t.optimizeAllBranches(1); out["oab1_second_s"] = time.time() - t0; out["oab1_second_lnl"] = l2
        out["oab_kernels"] = {k: {"n": v[0], "ms": v[2]} for k, v in t.kernelStats().items() if v[0]}
    for a_ in sys.argv:
        if a_.startswith("--nni="):     # PhyloTree::getBestNNIForBran (both NNIs, 5 branches each) on the first n internal branches
            n = int(a_[6:])
            t.clearAllPartialLH(); t.computeLikelihood()
            deg = {}
            for x, y in t.getTraversal():
                deg[int(x)] = deg.get(int(x), 0) + 1; deg[int(y)] = deg.get(int(y), 0) + 1
            br = [(int(x), int(y)) for x, y in t.getTraversal() if deg[int(x)] > 1 and deg[int(y)] > 1][:n]
            t.getBestNNIForBran(*br[0])
            t.setProfiling(False)       # wall time without the per-launch events ...
            t0 = time.time()
            best = [max(m[0] for m in t.getBestNNIForBran(x, y)) for x, y in br]
            out["nni_branches"] = len(br); out["nni_ms_per_branch"] = (time.time() - t0) / len(br) * 1e3
            out["nni_queued_calls_fallbacks"] = list(t.nniQueuedStats())
            t.resetCounters(); t.setProfiling(True)     # ... then once more for the per-kind device times
            [t.getBestNNIForBran(x, y) for x, y in br]
            out["nni_best_delta"] = max(best) - lnl
            out["nni_kernels"] = {k: {"n": v[0], "ms": v[2]} for k, v in t.kernelStats().items() if v[0]}
    print(json.dumps(out))
    t.close()


if __name__ == "__main__":
    main()
