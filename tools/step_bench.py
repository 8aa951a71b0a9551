#!/usr/bin/env python
"""The timed step of bench.py alone (headline workload, one GPU, device-resident inputs): `clearAllPartialLH + computeLikelihood +
computeLikelihoodDerv(root branch, theta rebuilt)` and the lnL-only loop, with CUDA events, plus the per-launch device times of the
traversal launches of both - for A/B runs of library variants on one box (tools/variants_run.sh style).
usage: python tools/step_bench.py [steps=20] [reps=3]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import iqtree_b200
from iqtree_b200 import modelinputs as mi


def main():
    import torch
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    lib = iqtree_b200.load_library()
    names, nwk, states = bench.make_inputs(bench.NPTN)
    model = mi.gtr_g4_inputs(bench.SEED, ncat=bench.NCAT)
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(states, np.ones(bench.NPTN), names=names, seq_type=0, nstates=4, state_unknown=18)
    t.readTreeString(nwk)
    t.setModel(0, bench.NCAT, 1, model["eval"], model["evec"], model["inv_evec"], model["rates"], model["props"], None, model["state_freq"], 0.0)
    t.setParams(0.1 / bench.NPTN, 100.0)
    t.setDevices([0])
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    t.setStream(stream.cuda_stream)
    t.initializeAllPartialLh()
    root_leaf, root_nei = 0, None
    for a, b in t.getTraversal():
        if a == 0 or b == 0:
            root_nei = int(b if a == 0 else a)
            break
    t.benchSteps(root_leaf, root_nei, 5)
    out = {"steps": steps, "step_ms": [], "lnl_ms": []}
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        res = t.benchSteps(root_leaf, root_nei, steps)
        e1.record(stream)
        torch.cuda.synchronize()
        out["step_ms"].append(e0.elapsed_time(e1) / steps)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            t.clearAllPartialLH()
            t.computeLikelihood()
        e1.record(stream)
        torch.cuda.synchronize()
        out["lnl_ms"].append(e0.elapsed_time(e1) / steps)
    out["lnl"], out["df"], out["ddf"] = res
    # per-launch device times: the full traversal (lnL) and the traversal of the derivative's branch
    t.resetCounters(); t.setProfiling(True)
    for _ in range(steps):
        t.clearAllPartialLH(); t.computeLikelihood()
    full = t.kernelStats()["partial"]
    t.resetCounters()
    for _ in range(steps):
        t.clearAllPartialLH(); t.computeLikelihood(); t.computeLikelihoodDerv(root_leaf, root_nei, reset_theta=True)
    both = t.kernelStats()
    t.setProfiling(False)
    out["full_traversal_ms"] = full[2] / full[0]
    out["derv_traversal_ms"] = (both["partial"][2] - full[2] / full[0] * steps) / max(both["partial"][0] - steps, 1)
    out["launches_per_step"] = {k: v[0] / steps for k, v in both.items() if v[0]}
    out["kernel_ms_per_step"] = {k: v[2] / steps for k, v in both.items() if v[0]}
    print(json.dumps(out))
    t.close()


if __name__ == "__main__":
    main()
