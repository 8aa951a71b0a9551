#!/usr/bin/env python
"""SURVEY 8(f).3 measurement (tuning tool, not the headline bench): full-tree lnL + df/ddf under site-specific frequency models
(PMSF-like: one exchangeability matrix, NMODEL distinct frequency profiles -> NMODEL eigensystems, every pattern mapped to one).
usage: python tools/sitemodel_bench.py [NTAXA=100] [NPTN=50000] [NSTATES=20] [NCAT=4] [NMODEL=NPTN] [reps=5] [--cpu=NPTN_SAMPLE]
--cpu: also times the scalar C restatement (oracle, one core) on the first NPTN_SAMPLE patterns."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import iqtree_b200
from iqtree_b200 import modelinputs as mi


def build(lib, ntaxa, nptn, S, C, eig, pm, st, nwk, names):
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(st[:, :nptn], np.ones(nptn), names=names, seq_type=0 if S == 4 else (1 if S == 20 else 4), nstates=S,
                   state_unknown={4: 18, 20: 23}.get(S, S))
    t.readTreeString(nwk)
    r, p = mi.discrete_gamma(0.7, C)
    t.setModel(3, C, eig[0].shape[0], eig[0], eig[1], eig[2], r, p, None, None, 0.0)
    t.setSiteModels(pm[:nptn])
    t.initializeAllPartialLh()
    return t


def main():
    a = [x for x in sys.argv[1:] if not x.startswith("--")]
    ntaxa, nptn, S, C = (int(a[i]) if len(a) > i else d for i, d in enumerate([100, 50000, 20, 4]))
    nmodel = int(a[4]) if len(a) > 4 else nptn
    reps = int(a[5]) if len(a) > 5 else 5
    cpu = 0
    for x in sys.argv[1:]:
        if x.startswith("--cpu="):
            cpu = int(x[6:])
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, 1, names)
    st = mi.simulate_patterns(nwk, names, nptn, S, 1)
    rng = np.random.default_rng(7)
    nr = S * (S - 1) // 2
    rates = 0.05 + rng.gamma(0.8, 1.0, size=nr)
    ev = np.empty((nmodel, S)); evc = np.empty((nmodel, S, S)); inv = np.empty((nmodel, S, S))
    for m in range(nmodel):
        pi = rng.dirichlet(np.full(S, 0.8)); pi = np.clip(pi, 2e-3, None); pi /= pi.sum()
        e0, e1, e2 = mi.reversible_eigensystem(rates, pi)
        ev[m], evc[m], inv[m] = e0, np.asarray(e1).reshape(S, S), np.asarray(e2).reshape(S, S)
    pm = (np.arange(nptn) % nmodel).astype(np.int32)
    lib = iqtree_b200.load_library()
    t = build(lib, ntaxa, nptn, S, C, (ev, evc, inv), pm, st, nwk, names)
    for _ in range(2):
        t.clearAllPartialLH(); lnl = t.computeLikelihood()
    t.resetCounters(); t.setProfiling(True)
    t0 = time.time()
    for _ in range(reps):
        t.clearAllPartialLH(); lnl = t.computeLikelihood()
    dt = (time.time() - t0) / reps
    trav = t.getTraversal()
    n1, n2 = int(trav[len(trav) // 2][0]), int(trav[len(trav) // 2][1])
    t.computeLikelihoodDerv(n1, n2, reset_theta=True)
    t0 = time.time()
    for _ in range(reps * 4):
        d = t.computeLikelihoodDerv(n1, n2, reset_theta=False)
    dd = (time.time() - t0) / (reps * 4)
    stats = t.kernelStats()
    flops = (ntaxa - 2) * nptn * C * (6 * S * S + S)
    out = {"what": "sitemodel", "shape": [ntaxa, nptn, S, C, nmodel], "lnl": lnl, "lnl_ms": dt * 1e3, "derv_ms": dd * 1e3,
           "updates_per_s": (ntaxa - 2) * nptn * C / dt, "matvec_tflops": flops / dt / 1e12,
           "exps_per_s": (ntaxa - 2) * nptn * C * 2 * S / dt, "mem_gb": t.memoryBytes() / 1e9,
           "kernels": {k: {"n": v[0], "ms": v[2], "alg_gbps": (v[1] / 1e9) / (v[2] * 1e-3) if v[2] > 0 else None} for k, v in stats.items() if v[0]}}
    t.close()
    if cpu:
        from oracle.oracle_lib import load_oracle_library
        o = build(load_oracle_library(), ntaxa, cpu, S, C, (ev, evc, inv), pm, st, nwk, names)
        t0 = time.time(); lo = o.computeLikelihood(); dc = time.time() - t0
        out["cpu_port"] = {"nptn": cpu, "lnl_s": dc, "updates_per_s": (ntaxa - 2) * cpu * C / dc, "cores": 1, "kind": "port (oracle/lk_oracle.c, scalar)"}
        o.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
