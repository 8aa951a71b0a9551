"""The drop-in, demonstrated: the reference's OWN code (IQ-TREE 1.4.3 compiled from its unmodified sources plus the ~160-line
integration/iqtree-1.4.3-lkcuda.patch and integration/phylotree_cuda.cpp) runs with `-cuda` on the engine behind its four kernel pointers
and its buffer allocator, and reproduces what the unmodified reference computed with its AVX kernels:

  * computeLikelihood, computePatternLikelihood, computePatternLhCat, computeLikelihoodDerv / FromBuffer on every branch,
    optimizeAllBranches(1) and to convergence, every partial_lh / scale_num vector - for all model families of tests/golden, in BOTH
    buffer modes (the default LM_PER_NODE with its re-orientation, and LM_ALL_BRANCH);
  * PhyloTree::getBestNNIForBran on every internal branch, RateFree::optimizeWithEM;
  * the complete command line: ModelFinder (-m TEST) + tree search (-n 2) give the CPU run's model, likelihood and topology.

`-m gpu` tests run oracle/_ref/ref_harness_cuda (links iq-tree_b200/lib/liblktree_cuda.so).  The CPU tests run the same objects
linked against the scalar C restatement of the ABI (ref_harness_oracle): they cover the binding's host logic where there is no GPU.
"""
import json
import os
import re
import shutil
import subprocess

import numpy as np
import pytest

import dropin_util as du
import golden_util as gu

DROPIN_CASES = [c for c in gu.golden_cases() if "sitefreq" not in c]     # -fs stays on the CPU kernels in this binding
FAST_CPU_CASES = ["dna_example_gtr_g4", "dna_example_gtr_i_g4", "dna_example_multifurc", "dna_deep_scaling", "aa_lg4x", "morph_mk_asc",
                  "dna_example_generic", "bin_jc2_g4"]
COMMON = ["-nt", "1", "-seed", "1"]


def _need(path):
    if not os.path.exists(path):
        pytest.skip("%s not built (make -f oracle/Makefile.ref cuda)" % os.path.relpath(path, du.ROOT))


@pytest.fixture(scope="module")
def work(tmp_path_factory):
    return str(tmp_path_factory.mktemp("dropin"))


def _dump(harness, case, work, extra, tag):
    d = os.path.join(work, "%s_%s" % (tag, case))
    du.run(harness, ["dump", d] + du.case_args(work)[case] + COMMON + ["-pre", os.path.join(work, "run_%s_%s" % (tag, case))] + extra, work)
    return d


# ------------------------------------------------------------------ CPU: the binding's host logic on the scalar engine ------------------------------------------------------------------

def test_patch_is_current():
    """the committed patch is what integration/make_patch.py generates from the reference sources (where they are available)"""
    if not os.path.isdir("/root/reference"):
        pytest.skip("needs /root/reference")
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_patch", os.path.join(du.ROOT, "integration", "make_patch.py"))
    mp = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mp)
    import difflib
    out = []
    for name in sorted(mp.EDITS):
        a, b = mp.patched(name)
        out.extend(difflib.unified_diff(a.splitlines(True), b.splitlines(True), "a/" + name, "b/" + name, n=3))
    cur = open(os.path.join(du.ROOT, "integration", "iqtree-1.4.3-lkcuda.patch"), "rb").read()
    assert "".join(out).encode("latin-1") == cur
    assert cur.count(b"\n+") < 120, "the binding is meant to stay a small patch"


@pytest.mark.parametrize("case", FAST_CPU_CASES)
@pytest.mark.parametrize("mode", ["-lhmemsave", "-nolhmemsave"])
def test_binding_on_scalar_engine(case, mode, work):
    _need(du.HARNESS_ORACLE)
    if mode == "-nolhmemsave" and case not in ("dna_example_gtr_g4", "dna_deep_scaling"):
        pytest.skip("LM_ALL_BRANCH: two cases suffice on the CPU")
    d = _dump(du.HARNESS_ORACLE, case, work, ["-cuda", mode], "o" + mode[1:3])
    if mode == "-lhmemsave":
        du.compare_dump(case, d, "cuda")
    else:       # other buffer mode: other set of valid vectors; values are compared through lnL / derivatives / optimisation
        m = json.load(open(os.path.join(d, "meta.json")))
        gm = gu.Golden(case).meta
        assert m["kernel"] == "cuda" and m["lh_mem_save"] == 1
        for k in ("lnl", "oab1_lnl", "oabfull_lnl"):
            assert abs(m[k] - gm[k]) <= gu.LNL_RTOL * abs(gm[k]), (k, m[k], gm[k])


# ------------------------------------------------------------------ GPU: the reference on the product engine ------------------------------------------------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("case", DROPIN_CASES)
def test_reference_on_engine(case, work, cuda_lib):
    """LM_PER_NODE (the reference's default): everything the fixture holds, same tolerances as the engine's own parity tests"""
    assert os.path.exists(du.HARNESS_CUDA), "oracle/_ref/ref_harness_cuda missing: run __graft_entry__.build() where /root/reference exists"
    d = _dump(du.HARNESS_CUDA, case, work, ["-cuda"], "g")
    du.compare_dump(case, d, "cuda")


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["dna_example_gtr_g4", "dna_deep_scaling_invar", "aa_lg_c20_g4", "codon_mg_g4", "dna_example_multifurc"])
def test_reference_on_engine_all_branch(case, work, cuda_lib):
    """LM_ALL_BRANCH (-nolhmemsave): one vector per direction, no re-orientation"""
    d = _dump(du.HARNESS_CUDA, case, work, ["-cuda", "-nolhmemsave"], "ga")
    m = json.load(open(os.path.join(d, "meta.json")))
    g = gu.Golden(case)
    gm = g.meta
    assert m["kernel"] == "cuda" and m["lh_mem_save"] == 1
    for k in ("lnl", "oab1_lnl", "oabfull_lnl"):
        assert abs(m[k] - gm[k]) <= gu.LNL_RTOL * abs(gm[k]), (k, m[k], gm[k])
    np.testing.assert_allclose(du.rd(d, "pattern_lh_scaled.f64", np.float64), g["pattern_lh_scaled"], rtol=gu.PTN_RTOL, atol=1e-300)
    dv, gdv = du.rd(d, "branch_derv.f64", np.float64).reshape(-1, 3), g["branch_derv"].reshape(-1, 3)
    np.testing.assert_allclose(dv[:, 2], gdv[:, 2], rtol=gu.LNL_RTOL, atol=0)
    a = gu.parse_adjacency(open(os.path.join(d, "tree_adj_oabfull.txt")).read())
    b = gu.parse_adjacency(g["tree_adj_oabfull"])
    assert max(abs(a[k] - b[k]) for k in a) <= gu.BRLEN_ATOL


def _check_nni(harness, case, nni5, work):
    z = np.load(os.path.join(gu.GOLDEN_DIR, "nni", "%s.nni%d.npz" % (case, nni5)))
    gm = json.loads(str(z["meta"]))
    d = os.path.join(work, "nni_%s_%d" % (case, nni5))
    du.run(harness, ["dumpnni", d, str(nni5)] + du.case_args(work)[case] + COMMON + ["-pre", os.path.join(work, "runn_" + case), "-cuda"], work)
    m = json.load(open(os.path.join(d, "metannni.json")))
    assert m["nbranch"] == gm["nbranch"]
    assert abs(m["lnl"] - gm["lnl"]) <= gu.LNL_RTOL * abs(gm["lnl"]) and abs(m["lnl_after"] - gm["lnl_after"]) <= gu.LNL_RTOL * abs(gm["lnl"])
    assert np.array_equal(du.rd(d, "nni_ids.i32", np.int32).reshape(-1, 6), z["ids"])
    v, gv = du.rd(d, "nni_vals.f64", np.float64).reshape(-1, 2, 6), z["vals"]
    np.testing.assert_allclose(v[:, :, 0], gv[:, :, 0], rtol=gu.LNL_RTOL, atol=0)
    assert np.abs(v[:, :, 1:] - gv[:, :, 1:]).max() <= gu.BRLEN_ATOL


@pytest.mark.gpu
@pytest.mark.parametrize("case,nni5", [("dna_example_gtr_g4", 1), ("dna_example_gtr_i_g4", 0), ("aa_lg_g4", 1), ("dna_deep_scaling", 1)])
def test_reference_nni_on_engine(case, nni5, work, cuda_lib):
    """the reference's getBestNNIForBran (phylotree.cpp:2873-3066) on every internal branch: the six nni_partial_lh vectors, neighbour
    replacement, re-orientation before the swap (LM_PER_NODE)"""
    _check_nni(du.HARNESS_CUDA, case, nni5, work)


def test_binding_nni_on_scalar_engine(work):
    _need(du.HARNESS_ORACLE)
    _check_nni(du.HARNESS_ORACLE, "dna_example_gtr_i_g4", 0, work)


def _check_em(harness, case, work):
    z = np.load(os.path.join(gu.GOLDEN_DIR, "em", case + ".npz"))
    gm = json.loads(str(z["meta"]))
    d = os.path.join(work, "em_" + case)
    du.run(harness, ["dumpem", d] + du.case_args(work)[case] + COMMON + ["-pre", os.path.join(work, "rune_" + case), "-cuda"], work)
    m = json.load(open(os.path.join(d, "metaem.json")))
    v = du.rd(d, "em_rates_props.f64", np.float64).reshape(4, gm["ncat"])
    assert abs(m["lnl_before"] - gm["lnl_before"]) <= gu.LNL_RTOL * abs(gm["lnl_before"])
    assert abs(m["lnl_after"] - gm["lnl_after"]) <= 1e-8 * abs(gm["lnl_after"]), (m["lnl_after"], gm["lnl_after"])
    np.testing.assert_allclose(v[2], z["rates"], rtol=1e-6)
    np.testing.assert_allclose(v[3], z["props"], rtol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["dna_example_hky_r3", "aa_lg_r4"])
def test_reference_ratefree_em_on_engine(case, work, cuda_lib):
    """RateFree::optimizeWithEM (model/ratefree.cpp:450-578) unchanged: computePatternLhCat + one temporary PhyloTree (own engine) per category"""
    _check_em(du.HARNESS_CUDA, case, work)


def test_binding_ratefree_em_on_scalar_engine(work):
    _need(du.HARNESS_ORACLE)
    _check_em(du.HARNESS_ORACLE, "dna_example_hky_r3", work)


def _search(harness, work, tag, extra, args):
    pre = os.path.join(work, "search_" + tag)
    out = du.run(harness, ["iqtree", "-s", os.path.join(work, "example.phy")] + args + ["-nt", "1", "-seed", "1", "-pre", pre] + extra, work, timeout=3000)
    rep = open(pre + ".iqtree").read()
    model = re.search(r"Model of substitution: (\S+)", rep).group(1)
    lnl = float(re.search(r"Log-likelihood of the tree: (-?[0-9.]+)", rep).group(1))
    tree = re.sub(r":[0-9.eE+-]+", "", open(pre + ".treefile").read().strip())
    return model, lnl, tree, out


@pytest.mark.gpu
def test_full_command_line_on_engine(work, cuda_lib):
    """`iqtree -s example.phy -m TEST -n 2`: ModelFinder (88 models: testModel, phylotesting.cpp:1161), parsimony start trees, NNI search
    (IQTree::optimizeNNI, iqtree.cpp:2125) and ModelFactory::optimizeParameters (model/modelfactory.cpp:937) of the reference run on the
    engine and give the model, likelihood and topology of the reference's own CPU run on the same box"""
    du.case_args(work)
    assert os.path.exists(du.HARNESS_REF) and os.path.exists(du.HARNESS_CUDA)
    cpu = _search(du.HARNESS_REF, work, "cpu", [], ["-m", "TEST", "-n", "2"])
    gpu = _search(du.HARNESS_CUDA, work, "gpu", ["-cuda"], ["-m", "TEST", "-n", "2"])
    assert "Kernel:" in cpu[3]
    assert gpu[0] == cpu[0], (gpu[0], cpu[0])
    assert abs(gpu[1] - cpu[1]) <= 2e-3, (gpu[1], cpu[1])           # the report prints 4 decimals; optimiser paths differ in the last digits
    assert gpu[2] == cpu[2], "topology differs"


def _mixture_em(harness, work, tag, extra):
    """`-m MIX{HKY,GTR}+G4 -te BIONJ`: components with free parameters make ModelMixture::optimizeParameters take optimizeWithEM
    (model/modelmixture.cpp:1418-1532): computePatternLhCat(WSL_MIXTURE), posterior weights -> ptn_freq of a temporary one-component tree that
    aliases the arena pointers, BFGS on that tree, new weights.  Returns lnL, the component table of the report and the tree."""
    pre = os.path.join(work, "mixem_" + tag)
    out = du.run(harness, ["iqtree", "-s", os.path.join(work, "example.phy"), "-m", "MIX{HKY,GTR}+G4", "-te", "BIONJ", "-nt", "1", "-seed", "1", "-pre", pre] + extra,
                 work, timeout=3000)
    rep = open(pre + ".iqtree").read()
    assert "Mixture model of substitution: MIX{HKY+F,GTR+F}+G4" in rep
    lnl = float(re.search(r"BEST SCORE FOUND : (-?[0-9.]+)", out).group(1))
    comps = re.findall(r"^\s+[12]\s+(?:HKY|GTR)\s+([0-9.]+)\s+([0-9.]+)\s+(\S+)", rep, flags=re.M)
    assert len(comps) == 2
    return lnl, comps, open(pre + ".treefile").read().strip()


def _check_mixture_em(cpu, eng):
    assert abs(eng[0] - cpu[0]) <= 2e-3, (eng[0], cpu[0])               # the report prints 4 decimals
    for (r0, w0, par0), (r1, w1, par1) in zip(cpu[1], eng[1]):
        assert abs(float(w0) - float(w1)) <= 2e-4, (w0, w1)             # EM weights, printed with 4 decimals
        a = [float(x) for x in re.findall(r"[0-9]+\.[0-9]+(?:e[+-]?[0-9]+)?", par0)]
        b = [float(x) for x in re.findall(r"[0-9]+\.[0-9]+(?:e[+-]?[0-9]+)?", par1)]
        assert len(a) == len(b) and all(abs(x - y) <= 2e-3 * max(1.0, abs(x)) for x, y in zip(a, b)), (par0, par1)
    la = gu.parse_newick_lengths(cpu[2]) if hasattr(gu, "parse_newick_lengths") else None
    if la is None:      # branch lengths of the two reports, in file order
        la = [float(x) for x in re.findall(r":([0-9.eE+-]+)", cpu[2])]
        lb = [float(x) for x in re.findall(r":([0-9.eE+-]+)", eng[2])]
        assert len(la) == len(lb) and max(abs(x - y) for x, y in zip(la, lb)) <= 1e-4
    assert re.sub(r":[0-9.eE+-]+", "", cpu[2]) == re.sub(r":[0-9.eE+-]+", "", eng[2])


def test_reference_mixture_em_on_scalar_engine(work):
    """ModelMixture::optimizeWithEM is the reference's own code: with the binding it runs unchanged on the engine ABI,
    including the temporary tree that borrows central_partial_lh (modelmixture.cpp:1427-1429)"""
    _need(du.HARNESS_ORACLE)
    if not os.path.exists(du.HARNESS_REF):
        pytest.skip("unmodified reference harness not built")
    du.case_args(work)
    _check_mixture_em(_mixture_em(du.HARNESS_REF, work, "cpu", []), _mixture_em(du.HARNESS_ORACLE, work, "eng", ["-cuda"]))


@pytest.mark.gpu
def test_reference_mixture_em_on_engine(work, cuda_lib):
    du.case_args(work)
    assert os.path.exists(du.HARNESS_REF) and os.path.exists(du.HARNESS_CUDA)
    _check_mixture_em(_mixture_em(du.HARNESS_REF, work, "cpug", []), _mixture_em(du.HARNESS_CUDA, work, "gpu", ["-cuda"]))


def test_reference_invar_free_rate_em_is_broken_upstream(work):
    """+I+R (model/ratefreeinvar.cpp) cannot be mirrored: the UNMODIFIED reference aborts at its own default start
    (ratefree.cpp:535: Assertion `fabs(sum_prop-1.0) < MIN_PROP')"""
    if not os.path.exists(du.HARNESS_REF):
        pytest.skip("unmodified reference harness not built")
    du.case_args(work)
    r = subprocess.run([du.HARNESS_REF, "iqtree", "-s", os.path.join(work, "example.phy"), "-m", "GTR+I+R3", "-te", "BIONJ", "-nt", "1", "-seed", "1",
                        "-pre", os.path.join(work, "ir_cpu")], cwd=work, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode != 0 and "sum_prop" in r.stdout


def test_binding_short_search_on_scalar_engine(work):
    """a short fixed-model search through the complete command line (CPU, scalar engine): topology and likelihood of the reference's CPU run"""
    _need(du.HARNESS_ORACLE)
    if not os.path.exists(du.HARNESS_REF):
        pytest.skip("unmodified reference harness not built")
    du.case_args(work)
    cpu = _search(du.HARNESS_REF, work, "cpu_s", [], ["-m", "HKY+G4", "-n", "1"])
    eng = _search(du.HARNESS_ORACLE, work, "eng_s", ["-cuda"], ["-m", "HKY+G4", "-n", "1"])
    assert abs(eng[1] - cpu[1]) <= 2e-3 and eng[2] == cpu[2]


def test_reference_codon_site_frequencies_are_broken_upstream(tmp_path):
    """Site-specific frequency models with more than 32 states (the reference's generic site-model kernels take any state
    count, and 64-state codon data is the only alphabet beyond 32) have no parity target: the UNMODIFIED reference corrupts its heap on
    `-st CODON -m GY -fs <file>` before the first likelihood is computed"""
    if not os.path.exists(du.HARNESS_REF):
        pytest.skip("unmodified reference harness not built")
    import sys
    sys.path.insert(0, os.path.join(du.ROOT, "oracle"))
    import make_golden as mg
    tmp = str(tmp_path)
    aln, tre = mg.synth_codon(tmp, 8, 60, 5, "codon")
    rng = np.random.default_rng(34)
    prof = np.clip(rng.dirichlet(np.full(64, 0.8), size=7), 2e-3, None)
    prof /= prof.sum(axis=1, keepdims=True)
    fs = os.path.join(tmp, "codfs.sitefreq")
    with open(fs, "w") as f:
        for s in range(60):     # codon sites are addressed as nucleotide ranges
            f.write("%d-%d %s\n" % (3 * s + 1, 3 * s + 3, " ".join("%.6f" % x for x in prof[rng.integers(7)])))
    r = subprocess.run([du.HARNESS_REF, "dump", os.path.join(tmp, "out"), "-s", aln, "-st", "CODON", "-m", "GY", "-fs", fs, "-te", tre, "-nt", "1", "-seed", "1",
                        "-pre", os.path.join(tmp, "run")], cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode != 0 and not os.path.exists(os.path.join(tmp, "out", "meta.json"))
