#!/usr/bin/env python
"""Wall time of the reference's complete command line (ModelFinder + tree search) on its own AVX + OpenMP kernels and on the engine
(-cuda), same files: `iqtree -s aln -m TEST -n N` through oracle/_ref/ref_harness[_cuda].  Reports chosen model, lnL, topology equality.
usage: python tools/dropin_cli_bench.py [NTAXA] [NPTN] [n_iter] [extra iqtree args...]"""
import json
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def run(harness, args, cwd, timeout):
    t0 = time.time()
    r = subprocess.run([harness, "iqtree"] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
    return time.time() - t0, r


def main():
    a = sys.argv[1:]
    ntaxa, nptn, niter = int(a[0]) if a else 50, int(a[1]) if len(a) > 1 else 50000, a[2] if len(a) > 2 else "2"
    extra = a[3:] if len(a) > 3 else ["-m", "TEST"]
    tmp = tempfile.mkdtemp(prefix="lkcli_")
    names, nwk, states = bench.make_inputs(nptn, ntaxa, 4, seed=11)
    aln = os.path.join(tmp, "a.phy")
    bench.write_phylip(aln, names, states, "ACGT")
    out = {"ntaxa": ntaxa, "patterns": nptn, "args": extra + ["-n", niter], "host_threads": os.cpu_count()}
    res = {}
    for tag, harness, more in (("reference", bench.HARNESS_REF, ["-nt", str(os.cpu_count() or 1)]), ("engine", bench.HARNESS_CUDA, ["-nt", "1", "-cuda"])):
        pre = os.path.join(tmp, tag)
        try:
            dt, r = run(harness, ["-s", aln] + extra + ["-n", niter, "-seed", "1", "-pre", pre] + more, tmp, int(os.environ.get("LK_CLI_TIMEOUT", 1500)))
        except subprocess.TimeoutExpired:
            out[tag] = {"timeout": True}
            continue
        rep = open(pre + ".iqtree").read() if os.path.exists(pre + ".iqtree") else ""
        m = re.search(r"Model of substitution: (\S+)", rep)
        l = re.search(r"Log-likelihood of the tree: (-?[0-9.]+)", rep)
        tree = re.sub(r":[0-9.eE+-]+", "", open(pre + ".treefile").read().strip()) if os.path.exists(pre + ".treefile") else None
        res[tag] = tree
        out[tag] = {"wall_s": dt, "rc": r.returncode, "model": m.group(1) if m else None, "lnl": float(l.group(1)) if l else None}
    if "reference" in res and "engine" in res:
        out["same_topology"] = res["reference"] == res["engine"]
        if out["reference"].get("wall_s") and out["engine"].get("wall_s"):
            out["speedup"] = out["reference"]["wall_s"] / out["engine"]["wall_s"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
