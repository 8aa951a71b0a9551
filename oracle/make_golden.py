#!/usr/bin/env python
"""TEST INFRASTRUCTURE: generate tests/golden/*.npz from the UNMODIFIED reference.

Runs oracle/_ref/ref_harness (the reference IQ-TREE 1.4.3-beta compiled by oracle/Makefile.ref from
/root/reference) on small alignments and packs, per case, the exact kernel INPUTS the reference saw and the
reference's own OUTPUTS into one compressed .npz.  Needs /root/reference (alignment fixtures + sources), so it
runs only in the build container; the resulting fixtures are committed and travel to the GPU box.

    python oracle/make_golden.py            # all cases
    python oracle/make_golden.py dna_example_gtr_g4 ...
    python oracle/make_golden.py --nni      # tests/golden/nni/*.npz: PhyloTree::getBestNNIForBran on every internal branch (NNI_CASES)
    python oracle/make_golden.py --aux      # tests/golden/aux/*.npz: RELL, pattern rates, mixture-weight EM of AUX_CASES
    python oracle/make_golden.py --em       # tests/golden/em/*.npz: RateFree::optimizeWithEM (EM_CASES; the inputs are the main fixtures)
    python oracle/make_golden.py --pars     # tests/golden/pars/*.npz: Fitch parsimony scores per branch + stepwise-addition trees (PARS_CASES)
    python oracle/make_golden.py --repair   # tests/golden/repair/*.npz: the branch kernel's NaN / Inf repair on a zeroed vector (REPAIR_CASES)

Synthetic alignments (codon, deep DNA tree, morphological, multifurcating) are generated here with fixed seeds.
"""
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import iqtree_b200  # noqa: E402
from iqtree_b200 import modelinputs as mi  # noqa: E402

HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
PARTIAL_BUDGET = 700 * 1024      # bytes of full partial_lh vectors kept per case

# the fixed example.phy topology of SURVEY Appendix C (IQ-TREE parsimony + NNI tree, seed 1)
EXAMPLE_TREE = "(FL-1-103:0.267966,(OSH-1-103:0.178481,CEU-1-103:0.192186):0.027962,(((TH-1-103:0.218789,((CAb1-103:0.132888,SU-1-103:0.295666):0.196127,(SI-1-103:0.139616,LU-1-103:0.203234):0.072703):0.012189):0.064091,CAa1-103:0.214885):0.496927,(((((((((((((RP-56-175:0.102676,A-14-133:0.154942):0.093931,((SA-60-179:0.170357,(BR-60-179:0.155192,HI-60-179:0.173803):0.035563):0.021249,YBN-56-175:0.075975):0.020526):0.035804,((PR-57-176:0.186295,SCR-58-177:0.112827):0.061885,((CPL-58-177:0.087689,LTU-59-178:0.050462):0.033910,TSP-58-177:0.072531):0.073590):0.041691):0.028147,(AN-56-175:0.147272,ANA-56-175:0.145316):0.058985):0.039938,PY-61-180:0.057538):0.090353,(MGI-58-176:0.090519,((ZE-48-166:0.162543,EAT-48-166:0.130523):0.009334,YSA-46-164:0.116990):0.069717):0.073030):0.049071,(TH-52-170:0.119428,LA-68-186:0.105581):0.163502):0.354227,(OE-36-161:0.167399,EME-43-168:0.185672):0.278778):0.053616,LRE-46-171:0.108692):0.045904,MO-29-154:0.142153):0.017619,GR-854-978:0.162413):0.175662,(((NE-2-109:0.364887,IFE-2-109:0.277987):0.130234,((HS-9-115:0.114813,RVI-7-114:0.105338):0.004759,IFE-8-115:0.298023):0.152597):0.112666,RVI-5-112:0.328546):0.227140):0.062625,((P6-2-107:0.253399,(SP-1-106:0.121438,PA-1-105:0.821678):0.216915):0.056940,HO-1-106:0.092991):0.424643):0.424592):0.090374);"


def read_phylip_names(path):
    names = []
    with open(path) as f:
        n = int(f.readline().split()[0])
        for line in f:
            if line.strip():
                names.append(line.split()[0])
            if len(names) == n:
                break
    return names


def write_phylip(path, names, seqs):
    with open(path, "w") as f:
        f.write("%d %d\n" % (len(names), len(seqs[0])))
        for n, s in zip(names, seqs):
            f.write("%s  %s\n" % (n, s))


def synth_dna(tmp, ntaxa, nsite, seed, tag, mean_len=0.1):
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, seed, names, mean_len=mean_len)
    st = mi.simulate_patterns(nwk, names, nsite, 4, seed)
    lut = np.array(list("ACGT"))
    seqs = ["".join(lut[st[i]]) for i in range(ntaxa)]
    # sprinkle gaps and ambiguity codes so that states 4..18 occur (alignment.cpp:877-888)
    rng = np.random.default_rng(seed + 7)
    seqs = [list(s) for s in seqs]
    for _ in range(max(4, ntaxa * nsite // 60)):
        seqs[rng.integers(ntaxa)][rng.integers(nsite)] = "-NRYWSKMBHDV"[rng.integers(12)]
    seqs = ["".join(s) for s in seqs]
    aln, tre = os.path.join(tmp, tag + ".phy"), os.path.join(tmp, tag + ".nwk")
    write_phylip(aln, names, seqs)
    open(tre, "w").write(nwk + "\n")
    return aln, tre


def synth_codon(tmp, ntaxa, ncodon, seed, tag):
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, seed, names, mean_len=0.15)
    bases = "TCAG"      # NCBI order is irrelevant here; any sense codons do
    stops = {"TAA", "TAG", "TGA"}
    sense = [a + b + c for a in bases for b in bases for c in bases if a + b + c not in stops]
    st = mi.simulate_patterns(nwk, names, ncodon, len(sense), seed)
    seqs = ["".join(sense[x] for x in st[i]) for i in range(ntaxa)]
    aln, tre = os.path.join(tmp, tag + ".phy"), os.path.join(tmp, tag + ".nwk")
    write_phylip(aln, names, seqs)
    open(tre, "w").write(nwk + "\n")
    return aln, tre


def synth_morph(tmp, ntaxa, nsite, nstates, seed, tag, variable_only=False):
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, seed, names, mean_len=0.2)
    st = mi.simulate_patterns(nwk, names, nsite * 3, nstates, seed)
    if variable_only:
        keep = [j for j in range(st.shape[1]) if len(set(st[:, j])) > 1]
        st = st[:, keep]
    st = st[:, :nsite]
    # make sure every state occurs somewhere so that num_states is what we asked for
    for s in range(nstates):
        st[s % ntaxa, s] = s
        st[(s + 1) % ntaxa, s] = (s + 1) % nstates
    seqs = ["".join(str(int(x)) for x in st[i]) for i in range(ntaxa)]
    aln, tre = os.path.join(tmp, tag + ".phy"), os.path.join(tmp, tag + ".nwk")
    write_phylip(aln, names, seqs)
    open(tre, "w").write(nwk + "\n")
    return aln, tre


def site_freq_file(tmp, nsite, nstates, seed, tag, nclass=9):
    """-fs input (ModelFactory::readSiteFreq, model/modelfactory.cpp:597-690): one line per site, `site f_1 .. f_S`; sites draw
    their frequency vector from `nclass` random profiles"""
    rng = np.random.default_rng(seed)
    prof = rng.dirichlet(np.full(nstates, 0.8), size=nclass)
    prof = np.clip(prof, 2e-3, None)
    prof /= prof.sum(axis=1, keepdims=True)
    path = os.path.join(tmp, tag + ".sitefreq")
    with open(path, "w") as f:
        for s in range(nsite):
            f.write("%d %s\n" % (s + 1, " ".join("%.6f" % x for x in prof[rng.integers(nclass)])))
    return path


def random_tree_file(tmp, names, seed, tag, mean_len=0.1):
    tre = os.path.join(tmp, tag + ".nwk")
    open(tre, "w").write(mi.random_tree_newick(len(names), seed, names, mean_len=mean_len) + "\n")
    return tre


def multifurcating_tree_file(tmp, names, seed, tag):
    """Collapse a few internal branches of a random binary tree into polytomies (degree-4 and degree-5 nodes)."""
    rng = np.random.default_rng(seed)
    n = len(names)
    order = list(rng.permutation(n))
    nm = [names[i] for i in order]

    def bl():
        return 0.02 + float(rng.exponential(0.08))

    # ((a,b,c),(d,e),(f,g,h,i), rest...) style: groups of 3 and 4 become polytomies
    groups, i = [], 0
    sizes = [3, 2, 4, 2, 3]
    k = 0
    while i < n - 3:
        s = min(sizes[k % len(sizes)], n - 3 - i)
        if s <= 0:
            break
        groups.append(nm[i:i + s]); i += s; k += 1
    rest = nm[i:]
    parts = []
    for g in groups:
        parts.append(g[0] + ":%.6f" % bl() if len(g) == 1 else "(" + ",".join("%s:%.6f" % (x, bl()) for x in g) + "):%.6f" % bl())
    # chain the groups as a caterpillar ending in the remaining taxa
    cur = "(" + ",".join("%s:%.6f" % (x, bl()) for x in rest) + ")"
    # top level: a 4-way multifurcation at the 'root' plus the chained groups
    top = rest[:]
    tree = "(" + ",".join(["%s:%.6f" % (x, bl()) for x in top] + parts) + ");"
    tre = os.path.join(tmp, tag + ".nwk")
    open(tre, "w").write(tree + "\n")
    return tre


def run_harness(outdir, args, cwd):
    cmd = [HARNESS, "dump", outdir] + args
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout[-4000:])
        raise RuntimeError("ref_harness failed: " + " ".join(cmd))
    return r.stdout


def rd(d, name, dtype):
    return np.fromfile(os.path.join(d, name), dtype=dtype)


def pack(case, d):
    meta = json.load(open(os.path.join(d, "meta.json")))
    P, U, T, B = meta["nptn"], meta["nunobs"], meta["ntaxa"], meta["block"]
    NP = P + U
    z = {"meta": json.dumps(meta)}
    z["tip_states"] = rd(d, "tip_states.u8", np.uint8).reshape(T, NP)
    z["const_char"] = rd(d, "const_char.i32", np.int32)
    z["site_pattern"] = rd(d, "site_pattern.i32", np.int32)
    z["taxa"] = np.array(open(os.path.join(d, "taxa.txt")).read().split("\n")[:T])
    for k in ("eval", "evec", "inv_evec", "rates", "props", "mixprops", "state_freq", "tip_partial_lh", "ptn_freq",
              "ptn_invar", "pattern_lh", "pattern_lh_scaled", "pattern_lh_cat", "branch_derv"):
        z[k] = rd(d, k + ".f64", np.float64)
    if meta["family"] == "sitemodel":
        z["ptn_model"] = rd(d, "ptn_model.i32", np.int32)
    z["traversal"] = rd(d, "traversal.i32", np.int32).reshape(-1, 2)
    for k in ("tree_adj", "tree_adj_oab1", "tree_adj_oabfull"):
        z[k] = open(os.path.join(d, k + ".txt")).read()
    z["tree0"] = open(os.path.join(d, "tree0.nwk")).read()
    theta = rd(d, "theta_root.f64", np.float64).reshape(NP, B)
    z["theta_root_rowsum"] = theta.sum(axis=1)
    if theta.nbytes <= 300 * 1024:
        z["theta_root"] = theta
    idx = [l.split() for l in open(os.path.join(d, "partials.txt")) if l.strip()]
    pf = np.array([[int(a), int(b)] for a, b, _ in idx], dtype=np.int32).reshape(-1, 2)
    z["partial_branch"] = pf
    z["partial_scale_factor"] = np.array([float(c) for _, _, c in idx])
    n = len(idx)
    sc = np.zeros((n, NP), dtype=np.int16)
    rs = np.zeros((n, NP))
    vec_bytes = NP * B * 8
    nkeep = 1 if vec_bytes > (1 << 20) else max(3, PARTIAL_BUDGET // max(vec_bytes, 1))
    keep = sorted(set(np.linspace(0, n - 1, min(n, nkeep)).astype(int).tolist())) if n else []
    full = []
    for i, (a, b, _) in enumerate(idx):
        lh = rd(d, "partial_%s_%s.f64" % (a, b), np.float64).reshape(NP, B)
        sc[i] = rd(d, "partial_%s_%s.i16" % (a, b), np.int16)
        rs[i] = lh.sum(axis=1)
        if i in keep:
            full.append(lh)
    z["partial_scale_num"] = sc
    z["partial_rowsum"] = rs
    z["partial_full_index"] = np.array(keep, dtype=np.int32)
    z["partial_full"] = np.stack(full) if full else np.zeros((0, NP, B))
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, case + ".npz")
    np.savez_compressed(path, **z)
    print("%-28s lnl %.10f  P=%d U=%d B=%d  partials=%d (full %d)  scaled_ptns=%d  %.0f KB" % (
        case, meta["lnl"], P, U, B, n, len(keep), int((sc > 0).any(axis=0).sum()), os.path.getsize(path) / 1024))


AUX_CASES = ["dna_example_gtr_g4", "dna_example_gtr_i_g4", "dna_deep_scaling", "aa_lg_g4", "aa_lg_c20_g4", "codon_mg_g4"]
AUX_SAMPLES = 24


NNI_CASES = [("dna_example_gtr_g4", 1), ("dna_example_gtr_i_g4", 0), ("aa_lg_g4", 1), ("dna_deep_scaling", 1)]


def pack_nni(case, nni5, d):
    m = json.load(open(os.path.join(d, "metannni.json")))
    z = {"meta": json.dumps(m), "ids": rd(d, "nni_ids.i32", np.int32).reshape(-1, 6), "vals": rd(d, "nni_vals.f64", np.float64).reshape(-1, 2, 6)}
    os.makedirs(os.path.join(OUT, "nni"), exist_ok=True)
    path = os.path.join(OUT, "nni", "%s.nni%d.npz" % (case, nni5))
    np.savez_compressed(path, **z)
    print("%-28s nni5=%d: %d branches, lnl %.10f (after: %.10f), best delta %.4f  %.0f KB" % (
        case, nni5, m["nbranch"], m["lnl"], m["lnl_after"], z["vals"][:, :, 0].max() - m["lnl"], os.path.getsize(path) / 1024))


EM_CASES = ["dna_example_hky_r3", "aa_lg_r4", "dna_deep_gtr_r4"]


def pack_em(case, d):
    m = json.load(open(os.path.join(d, "metaem.json")))
    v = rd(d, "em_rates_props.f64", np.float64).reshape(4, m["ncat"])
    z = {"meta": json.dumps(m), "rates_before": v[0], "props_before": v[1], "rates": v[2], "props": v[3]}
    os.makedirs(os.path.join(OUT, "em"), exist_ok=True)
    path = os.path.join(OUT, "em", case + ".npz")
    np.savez_compressed(path, **z)
    print("%-28s em: lnl %.6f -> %.6f rates %s props %s" % (case, m["lnl_before"], m["lnl_after"], np.round(v[2], 5), np.round(v[3], 5)))


PARS_CASES = [("dna_example_gtr_g4", 1), ("dna_example_gtr_g4", 7), ("dna_deep_scaling", 1), ("aa_lg_g4", 1), ("aa_lg_g4", 5), ("codon_gy", 1),
              ("morph_mk_g4", 1), ("bin_jc2_g4", 3)]


def pack_pars(case, seed, d):
    m = json.load(open(os.path.join(d, "metapars.json")))
    nt, ns = m["ntaxa"], m["nsites_informative"]
    z = {"meta": json.dumps(m), "sites": rd(d, "pars_sites.u8", np.uint8).reshape(nt, ns),
         "branches": rd(d, "pars_branches.i32", np.int32).reshape(-1, 4), "order": rd(d, "pars_order.i32", np.int32),
         "tree": open(os.path.join(d, "pars_tree.txt")).read(), "stepwise_tree": open(os.path.join(d, "pars_stepwise_tree.txt")).read()}
    os.makedirs(os.path.join(OUT, "pars"), exist_ok=True)
    path = os.path.join(OUT, "pars", "%s.seed%d.npz" % (case, seed))
    np.savez_compressed(path, **z)
    print("%-28s seed %d: %d taxa x %d informative sites, score %d, stepwise %d (check %d)  %.0f KB" % (
        case, seed, nt, ns, m["score"], m["stepwise_score"], m["stepwise_check"], os.path.getsize(path) / 1024))


REPAIR_CASES = ["dna_example_gtr_g4", "aa_lg_g4", "dna_deep_scaling", "codon_gy"]


def pack_repair(case, d):
    m = json.load(open(os.path.join(d, "metarepair.json")))
    z = {"meta": json.dumps(m), "pattern_lh_repaired": rd(d, "pattern_lh_repaired.f64", np.float64)}
    os.makedirs(os.path.join(OUT, "repair"), exist_ok=True)
    path = os.path.join(OUT, "repair", case + ".npz")
    np.savez_compressed(path, **z)
    print("%-28s repair: lnl %.10f -> %.10f, zeroed %s of vector %s" % (case, m["lnl"], m["lnl_repaired"], m["zeroed"], m["vector"]))


def pack_aux(case, d):
    m = json.load(open(os.path.join(d, "metax.json")))
    P, ns = m["nptn"], m["nsamples"]
    z = {"meta": json.dumps(m)}
    z["boot_samples"] = rd(d, "boot_samples.f32", np.float32).reshape(ns, P)
    z["rell"] = rd(d, "rell.f64", np.float64)
    z["ptn_lh_full"] = rd(d, "ptn_lh_full.f64", np.float64)
    if m["nrate_cat"]:
        z["pattern_rates"] = rd(d, "pattern_rates.f64", np.float64)
        z["pattern_cat"] = rd(d, "pattern_cat.i32", np.int32)
    if m["has_em"]:
        z["lhcat_mixture"] = rd(d, "lhcat_mixture.f64", np.float64).reshape(P, m["nmix"])
        z["em_props_before"] = rd(d, "em_props_before.f64", np.float64)
        z["em_props"] = rd(d, "em_props.f64", np.float64)
    os.makedirs(os.path.join(OUT, "aux"), exist_ok=True)
    path = os.path.join(OUT, "aux", case + ".npz")
    np.savez_compressed(path, **z)
    print("%-28s aux: lnl %.10f rell[0] %.6f rates %d em %d (em lnl %.10f)  %.0f KB" % (
        case, m["lnl"], z["rell"][0], m["nrate_cat"], m["has_em"], m["em_lnl"], os.path.getsize(path) / 1024))


def build_cases(tmp, ex, prot):
    """harness arguments of every golden case; `ex` / `prot` are the two alignments that come from the reference's example
    directories (the tests rebuild them from the fixtures), everything else is generated here with fixed seeds"""
    ex_tree = os.path.join(tmp, "example.nwk")
    open(ex_tree, "w").write(EXAMPLE_TREE + "\n")
    prot_tree = random_tree_file(tmp, read_phylip_names(prot), 3, "prot", mean_len=0.15)
    deep_aln, deep_tree = synth_dna(tmp, 320, 90, 11, "deep", mean_len=0.35)
    cod_aln, cod_tree = synth_codon(tmp, 8, 60, 5, "codon")
    mo_aln, mo_tree = synth_morph(tmp, 12, 80, 5, 9, "morph")
    asc_aln, asc_tree = synth_morph(tmp, 10, 70, 3, 13, "morphasc", variable_only=True)
    bin_aln, bin_tree = synth_morph(tmp, 9, 64, 2, 21, "bin")
    mf_names = read_phylip_names(ex)
    mf_tree = multifurcating_tree_file(tmp, mf_names, 4, "multif")

    ex_fs = site_freq_file(tmp, 384, 4, 31, "exfs")
    prot_fs = site_freq_file(tmp, 269, 20, 32, "protfs")
    deep_fs = site_freq_file(tmp, 90, 4, 33, "deepfs")
    return {
        "dna_sitefreq_gtr_g4": ["-s", ex, "-m", "GTR+G4", "-fs", ex_fs, "-te", ex_tree],
        "dna_sitefreq_deep_scaling": ["-s", deep_aln, "-m", "HKY+G4", "-fs", deep_fs, "-te", deep_tree],
        "aa_sitefreq_lg_g4": ["-s", prot, "-m", "LG+G4", "-fs", prot_fs, "-te", prot_tree],
        "aa_sitefreq_lg_generic": ["-s", prot, "-m", "LG", "-fs", prot_fs, "-te", prot_tree, "-fastlk"],
        "dna_example_gtr_g4": ["-s", ex, "-m", "GTR+G4", "-te", ex_tree],
        "dna_example_gtr_i_g4": ["-s", ex, "-m", "GTR+I+G4", "-te", ex_tree],
        "dna_example_hky_r3": ["-s", ex, "-m", "HKY+R3", "-te", ex_tree],
        "dna_example_jc": ["-s", ex, "-m", "JC", "-te", ex_tree],
        "aa_lg_r4": ["-s", prot, "-m", "LG+R4", "-te", prot_tree],
        "dna_deep_gtr_r4": ["-s", deep_aln, "-m", "GTR+R4", "-te", deep_tree],
        "dna_example_generic": ["-s", ex, "-m", "GTR+G4", "-te", ex_tree, "-fastlk"],
        "dna_example_multifurc": ["-s", ex, "-m", "GTR+G4", "-te", mf_tree],
        "dna_deep_scaling": ["-s", deep_aln, "-m", "GTR+G4", "-te", deep_tree],
        "dna_deep_scaling_invar": ["-s", deep_aln, "-m", "HKY+I+G4", "-te", deep_tree],
        "aa_lg_g4": ["-s", prot, "-m", "LG+G4", "-te", prot_tree],
        "aa_lg4x": ["-s", prot, "-m", "LG4X", "-te", prot_tree],
        "aa_lg_c20_g4": ["-s", prot, "-m", "LG+C20+G4", "-te", prot_tree],
        "codon_gy": ["-s", cod_aln, "-st", "CODON", "-m", "GY", "-te", cod_tree],
        "codon_mg_g4": ["-s", cod_aln, "-st", "CODON", "-m", "MG+G4", "-te", cod_tree],
        "morph_mk_g4": ["-s", mo_aln, "-st", "MORPH", "-m", "MK+G4", "-te", mo_tree],
        "morph_mk_asc": ["-s", asc_aln, "-st", "MORPH", "-m", "MK+ASC", "-te", asc_tree],
        "bin_jc2_g4": ["-s", bin_aln, "-st", "BIN", "-m", "JC2+G4", "-te", bin_tree],
    }


def main():
    if not os.path.exists(HARNESS):
        raise SystemExit("build the reference first: make -f oracle/Makefile.ref -j8")
    tmp = tempfile.mkdtemp(prefix="lk_golden_")
    ex = os.path.join(tmp, "example.phy")
    shutil.copy(os.path.join(REF, "example", "example.phy"), ex)
    prot = os.path.join(tmp, "prot.phy")
    shutil.copy(os.path.join(REF, "test_scripts", "test_data", "prot_M126_27_269.phy"), prot)
    cases = build_cases(tmp, ex, prot)
    if sys.argv[1:2] == ["--nni"]:
        for case, nni5 in NNI_CASES:
            d = os.path.join(tmp, "outn_" + case)
            cmd = [HARNESS, "dumpnni", d, str(nni5)] + cases[case] + ["-nt", "1", "-seed", "1", "-pre", os.path.join(tmp, "runn_" + case)]
            r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                print("%-28s nni FAILED:\n%s" % (case, r.stdout[-1500:]))
                continue
            pack_nni(case, nni5, d)
        shutil.rmtree(tmp, ignore_errors=True)
        return
    if sys.argv[1:2] == ["--em"]:
        for case in EM_CASES:
            d = os.path.join(tmp, "oute_" + case)
            cmd = [HARNESS, "dumpem", d] + cases[case] + ["-nt", "1", "-seed", "1", "-pre", os.path.join(tmp, "rune_" + case)]
            r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                print("%-28s em FAILED:\n%s" % (case, r.stdout[-1500:]))
                continue
            pack_em(case, d)
        shutil.rmtree(tmp, ignore_errors=True)
        return
    if sys.argv[1:2] == ["--repair"]:
        for case in REPAIR_CASES:
            d = os.path.join(tmp, "outr_" + case)
            cmd = [HARNESS, "dumprepair", d] + cases[case] + ["-nt", "1", "-seed", "1", "-pre", os.path.join(tmp, "runr_" + case)]
            r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                print("%-28s repair FAILED:\n%s" % (case, r.stdout[-1500:]))
                continue
            pack_repair(case, d)
        shutil.rmtree(tmp, ignore_errors=True)
        return
    if sys.argv[1:2] == ["--pars"]:
        for case, seed in PARS_CASES:
            d = os.path.join(tmp, "outp_%s_%d" % (case, seed))
            cmd = [HARNESS, "dumppars", d, str(seed)] + cases[case] + ["-nt", "1", "-seed", "1", "-pre", os.path.join(tmp, "runp_" + case)]
            r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                print("%-28s pars FAILED:\n%s" % (case, r.stdout[-1500:]))
                continue
            pack_pars(case, seed, d)
        shutil.rmtree(tmp, ignore_errors=True)
        return
    if sys.argv[1:2] == ["--aux"]:
        for case in (sys.argv[2:] or AUX_CASES):
            d = os.path.join(tmp, "outx_" + case)
            cmd = [HARNESS, "dumpx", d, str(AUX_SAMPLES)] + cases[case] + ["-nt", "1", "-seed", "1", "-pre", os.path.join(tmp, "runx_" + case)]
            r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                print("%-28s aux FAILED:\n%s" % (case, r.stdout[-1500:]))
                continue
            pack_aux(case, d)
        shutil.rmtree(tmp, ignore_errors=True)
        return
    want = sys.argv[1:] or list(cases)
    for case in want:
        d = os.path.join(tmp, "out_" + case)
        try:
            run_harness(d, cases[case] + ["-nt", "1", "-seed", "1", "-pre", os.path.join(tmp, "run_" + case)], tmp)
            pack(case, d)
        except Exception as ex_:      # keep going: a case the reference rejects is reported, not fatal
            print("%-28s FAILED: %s" % (case, ex_))
    shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
