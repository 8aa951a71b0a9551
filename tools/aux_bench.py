#!/usr/bin/env python
"""SURVEY 8(f).1 / 8(f).2 on the device at BASELINE shapes (tuning / measurement tool, not the headline bench):
   rell : UFBoot RELL - NSAMPLES bootstrap replicates x NPTN patterns (DNA 100 taxa GTR+G4): ms per candidate tree, GB/s of the matrix
          actually streamed (uint8 counts) and of the reference's float matrix (algorithmic), vs a numpy fp32 GEMV on the host cores
   em   : ModelMixture::optimizeWeights on LG+C20+G4-like mixtures: ms per EM step vs the host pass it replaces
   rem  : RateFree::optimizeWithEM (+R4 on DNA 100 taxa): wall time of the whole EM, likelihood evaluations, device share
usage: python tools/aux_bench.py rell [NPTN=1000000] [NSAMPLES=1000] [reps=10]
       python tools/aux_bench.py rem [NPTN=1000000]
       python tools/aux_bench.py em [NTAXA=100] [NPTN=10000] [reps=5]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import iqtree_b200
from iqtree_b200 import modelinputs as mi


def rell(nptn, nsamples, reps):
    ntaxa = 100
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, 1, names)
    st = mi.simulate_patterns(nwk, names, nptn, 4, 1)
    mdl = mi.gtr_g4_inputs(1, ncat=4)
    lib = iqtree_b200.load_library()
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(st, np.ones(nptn), names=names, seq_type=0, nstates=4, state_unknown=18)
    t.readTreeString(nwk)
    t.setModel(0, 4, 1, mdl["eval"], mdl["evec"], mdl["inv_evec"], mdl["rates"], mdl["props"], None, mdl["state_freq"], 0.0)
    t.setParams(0.1 / nptn if nptn >= 100000 else 1e-6, 100.0)
    t.initializeAllPartialLh()
    lnl, plh = t.computeLikelihood(pattern_lh=True)
    # multinomial resampling of nptn sites over nptn equally frequent patterns (iqtree.cpp:326-338 with unit pattern frequencies)
    rng = np.random.default_rng(3)
    boot = np.empty((nsamples, nptn), dtype=np.float32)
    for s in range(nsamples):
        boot[s] = np.bincount(rng.integers(0, nptn, size=nptn), minlength=nptn)
    t0 = time.time()
    t.setBootSamples(boot)
    t_up = time.time() - t0
    import ctypes
    k, nb = ctypes.c_int32(), ctypes.c_uint64()
    lib.lkcuda_boot_storage(t.engine(0), ctypes.byref(k), ctypes.byref(nb))
    r = t.computeRell()
    t.resetCounters(); t.setProfiling(True)
    t0 = time.time()
    for _ in range(reps):
        r = t.computeRell()
    wall = (time.time() - t0) / reps
    stats = t.kernelStats()
    n, alg, ms = stats["rell"]
    ker_ms = ms / max(n, 1)
    # host baseline: the same float GEMV with numpy (multi-threaded BLAS sgemv) on the float inputs the reference uses
    x = plh.astype(np.float32)
    t0 = time.time()
    hreps = max(1, min(3, reps))
    for _ in range(hreps):
        ref = boot @ x
    host = (time.time() - t0) / hreps
    exact = boot.astype(np.float64) @ x.astype(np.float64) if nsamples * nptn <= 2e8 else None
    out = {"what": "rell", "nptn": nptn, "nsamples": nsamples, "lnl": lnl, "bytes_per_entry": k.value, "matrix_gb": nb.value / 1e9,
           "upload_s": t_up, "kernel_ms": ker_ms, "wall_ms_per_tree": wall * 1e3,
           "streamed_gbps": (nb.value / 1e9) / (ker_ms * 1e-3) if ker_ms else None,
           "algorithmic_gbps_float_matrix": (alg / max(n, 1) / 1e9) / (ker_ms * 1e-3) if ker_ms else None,
           "host_numpy_sgemv_ms": host * 1e3, "host_cores": os.cpu_count(),
           "max_rel_err_vs_host_float": float(np.max(np.abs(r - ref) / np.abs(ref))),
           "max_rel_err_vs_fp64": float(np.max(np.abs(r - exact) / np.abs(exact))) if exact is not None else None}
    print(json.dumps(out))
    t.close()


def em(ntaxa, nptn, reps):
    S, C, M = 20, 4, 20
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, 1, names)
    st = mi.simulate_patterns(nwk, names, nptn, S, 1)
    mdl = mi.random_reversible_inputs(S, 1, ncat=C)
    comps = [mi.random_reversible_inputs(S, 100 + m, ncat=C) for m in range(M)]
    lib = iqtree_b200.load_library()
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(st, np.ones(nptn), names=names, seq_type=1, nstates=S, state_unknown=23)
    t.readTreeString(nwk)
    t.setModel(2, C, M, np.concatenate([c["eval"].ravel() for c in comps]), np.concatenate([c["evec"].ravel() for c in comps]),
               np.concatenate([c["inv_evec"].ravel() for c in comps]), mdl["rates"], mdl["props"], np.full(M, 1.0 / M), mdl["state_freq"], 0.0)
    t.initializeAllPartialLh()
    lnl0 = t.computeLikelihood()
    lnl1, w = t.optimizeMixtureWeights(M)
    t.resetCounters(); t.setProfiling(True)
    t0 = time.time()
    for _ in range(reps):
        lnl1, w = t.optimizeMixtureWeights(M)
    wall = (time.time() - t0) / reps
    stats = t.kernelStats()
    n, alg, ms = stats["em"]
    # what the reference does instead: fetch _pattern_lh_cat (P x C*M doubles) and run the two host loops per EM step
    t0 = time.time()
    raw = t.computePatternLhCat(4)
    fetch = time.time() - t0
    lc = raw.reshape(nptn, M, C).sum(axis=2)
    t0 = time.time()
    lk = lc.sum(axis=1)
    new_prop = (lc * (1.0 / lk)[:, None]).sum(axis=0) / nptn
    host_step = time.time() - t0
    out = {"what": "em", "ntaxa": ntaxa, "nptn": nptn, "ncat": C, "nmix": M, "lnl_before": lnl0, "lnl_after": lnl1, "weights_sum": float(w.sum()),
           "em_steps": n / reps, "em_kernel_ms_per_step": ms / max(n, 1), "optimizeWeights_wall_ms": wall * 1e3,
           "host_fetch_lh_cat_ms": fetch * 1e3, "host_numpy_em_step_ms": host_step * 1e3}
    print(json.dumps(out))
    t.close()


def rem(nptn):
    ntaxa = 100
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, 1, names)
    st = mi.simulate_patterns(nwk, names, nptn, 4, 1)
    mdl = mi.gtr_g4_inputs(1, ncat=4)
    lib = iqtree_b200.load_library()
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(st, np.ones(nptn), names=names, seq_type=0, nstates=4, state_unknown=18)
    t.readTreeString(nwk)
    t.setModel(0, 4, 1, mdl["eval"], mdl["evec"], mdl["inv_evec"], mdl["rates"], mdl["props"], None, mdl["state_freq"], 0.0)
    t.setParams(0.1 / nptn if nptn >= 100000 else 1e-6, 100.0)
    t.initializeAllPartialLh()
    lnl0 = t.computeLikelihood()
    n0 = t.numPartialLhComputations()
    t0 = time.time()
    lnl1, rates, props = t.optimizeRateFreeEM(4)
    wall = time.time() - t0
    out = {"what": "ratefree_em", "ntaxa": ntaxa, "nptn": nptn, "ncat": 4, "lnl_before": lnl0, "lnl_after": lnl1,
           "rates": [float(x) for x in rates], "props": [float(x) for x in props], "wall_s": wall,
           "main_tree_partial_computations": t.numPartialLhComputations() - n0,
           "note": "every EM step: 1 per-category likelihood pass (4 categories) + E-step on the device + 4 Brent searches of ~10-20 "
                   "one-category full-tree likelihoods each; the P x 4 posterior matrix never leaves the device"}
    # the unmodified reference's own EM on a bounded sample of the same alignment, all host threads (oracle/_ref/ref_harness dumpem)
    harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    ref_sites = int(os.environ.get("LK_REF_SITES", 100000))
    if os.path.exists(harness) and ref_sites > 0:
        import subprocess
        import tempfile
        tmp = tempfile.mkdtemp(prefix="lk_rem_")
        ns = min(ref_sites, nptn)
        lut = np.array(list("ACGT"))
        aln = os.path.join(tmp, "sample.phy")
        with open(aln, "w") as f:
            f.write("%d %d\n" % (ntaxa, ns))
            for i, n in enumerate(names):
                f.write("%s  %s\n" % (n, "".join(lut[st[i, :ns]])))
        tre = os.path.join(tmp, "sample.nwk")
        open(tre, "w").write(nwk + "\n")
        cmd = [harness, "dumpem", os.path.join(tmp, "out"), "-s", aln, "-m", "GTR+R4", "-te", tre, "-nt", str(os.cpu_count() or 1), "-seed", "1",
               "-pre", os.path.join(tmp, "ref")]
        r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        try:
            m = json.load(open(os.path.join(tmp, "out", "metaem.json")))
            out["reference"] = {"sample_sites": ns, "patterns": m["nptn"], "threads": m["threads"], "em_seconds": m["em_seconds"],
                                "lnl_before": m["lnl_before"], "lnl_after": m["lnl_after"],
                                "note": "its own initial +R4 parameters (not the ones above): a timing, not a parity check"}
            out["seconds_per_million_patterns"] = {"b200": wall * 1e6 / nptn, "reference": m["em_seconds"] * 1e6 / m["nptn"]}
        except Exception as ex:
            out["reference"] = {"error": str(ex), "tail": r.stdout[-400:]}
    print(json.dumps(out))
    t.close()


if __name__ == "__main__":
    a = sys.argv[1:]
    if a and a[0] == "rem":
        rem(int(a[1]) if len(a) > 1 else 1000000)
    elif a and a[0] == "em":
        em(int(a[1]) if len(a) > 1 else 100, int(a[2]) if len(a) > 2 else 10000, int(a[3]) if len(a) > 3 else 5)
    else:
        rell(int(a[1]) if len(a) > 1 else 1000000, int(a[2]) if len(a) > 2 else 1000, int(a[3]) if len(a) > 3 else 10)
