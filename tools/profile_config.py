#!/usr/bin/env python
"""Run ONE BASELINE shape through the reference's own code on the engine (ref_harness_cuda -cuda), for ncu:

    ncu --set full --clock-control none --import-source on -k regex:traverse_mma -c 2 -o gpurun_out/x \
        python tools/profile_config.py protein_lg_g4 [reps]

Generates the same synthetic files bench.py's `configs` leg uses (same seed), then execs the harness, so the kernels ncu sees are
the ones the bench line timed.  Keys: headline, protein_lg_g4, codon_gy, codon_mg_g4, lg_c20_g4.
"""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    key = sys.argv[1]
    reps = sys.argv[2] if len(sys.argv) > 2 else "2"
    tmp = tempfile.mkdtemp(prefix="lkprof_")
    if key == "headline":
        cfg = dict(nptn=bench.NPTN, ntaxa=bench.NTAXA, S=4, alphabet="ACGT", margs=["-m", "GTR+G4"])
        seed = bench.SEED
    else:
        cfg = [c for c in bench.config_list() if c["key"] == key][0]
        seed = bench.SEED + 3
    names, nwk, states = bench.make_inputs(cfg["nptn"], cfg["ntaxa"], cfg["S"], seed=seed)
    aln, tre = os.path.join(tmp, "a.phy"), os.path.join(tmp, "a.nwk")
    bench.write_phylip(aln, names, states, cfg["alphabet"])
    open(tre, "w").write(nwk + "\n")
    cmd = [bench.HARNESS_CUDA, "time", reps, "0", "0", "-s", aln] + cfg["margs"] + ["-te", tre, "-nt", "1", "-seed", "1", "-pre", os.path.join(tmp, "run"), "-cuda"]
    sys.stdout.flush()
    os.execv(cmd[0], cmd)


if __name__ == "__main__":
    main()
