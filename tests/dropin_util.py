"""Helpers of tests/test_dropin.py: run the reference WITH the engine bound behind its kernel pointers
(oracle/_ref/ref_harness_cuda = the unmodified IQ-TREE 1.4.3 sources + integration/iqtree-1.4.3-lkcuda.patch +
integration/phylotree_cuda.cpp, built by `make -f oracle/Makefile.ref cuda`) on the inputs of the golden fixtures and
compare what IT dumps with what the unmodified reference dumped (tests/golden/*.npz).

Nothing here reads /root/reference: the two alignments that originally came from the reference's example directories
are rebuilt from the fixtures (pattern matrix + site -> pattern map), the synthetic ones are regenerated with the
fixed seeds of oracle/make_golden.py.
"""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import make_golden as mg  # noqa: E402

import golden_util as gu  # noqa: E402

REF_DIR = os.path.join(ROOT, "oracle", "_ref")
HARNESS_CUDA = os.path.join(REF_DIR, "ref_harness_cuda")        # patched reference + liblktree_cuda.so (the product)
HARNESS_ORACLE = os.path.join(REF_DIR, "ref_harness_oracle")    # the same objects + the CPU restatement behind the same C ABI
HARNESS_REF = os.path.join(REF_DIR, "ref_harness")              # the unmodified reference

DNA = "ACGT"
AA = "ARNDCQEGHILKMFPSTWYV"
# state = bit mask + 3 (A = 1, C = 2, G = 4, T = 8; alignment.cpp:870-888)
DNA_AMBIG = {1 + 4 + 3: "R", 2 + 8 + 3: "Y", 1 + 8 + 3: "W", 2 + 4 + 3: "S", 1 + 2 + 3: "M", 4 + 8 + 3: "K",
             2 + 4 + 8 + 3: "B", 1 + 2 + 8 + 3: "H", 1 + 4 + 8 + 3: "D", 1 + 2 + 4 + 3: "V"}


def alignment_from_golden(case, path):
    """PHYLIP file with the sites of the fixture's alignment in their original order (DNA / protein)"""
    g = gu.Golden(case)
    m = g.meta
    P = m["nptn"]
    st = g["tip_states"][:, :P][:, g["site_pattern"]]
    if m["seq_type"] == 0:
        lut = {i: c for i, c in enumerate(DNA)}
        lut.update(DNA_AMBIG)
    elif m["seq_type"] == 1:
        lut = {i: c for i, c in enumerate(AA)}
        lut.update({20: "B", 21: "Z", 22: "J"})
    else:
        raise ValueError("alignment_from_golden: DNA / protein only")
    lut[m["state_unknown"]] = "-"
    names = [str(x) for x in g["taxa"]]
    mg.write_phylip(path, names, ["".join(lut[int(x)] for x in row) for row in st])
    return path


_cases = {}


def case_args(tmp):
    """harness arguments of every golden case, files under `tmp`"""
    if tmp not in _cases:
        ex = alignment_from_golden("dna_example_gtr_g4", os.path.join(tmp, "example.phy"))
        prot = alignment_from_golden("aa_lg_g4", os.path.join(tmp, "prot.phy"))
        _cases[tmp] = mg.build_cases(tmp, ex, prot)
    return _cases[tmp]


def run(harness, args, cwd, timeout=1200):
    r = subprocess.run([harness] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
    assert r.returncode == 0, "%s %s failed (%d):\n%s" % (os.path.basename(harness), " ".join(args[:3]), r.returncode, r.stdout[-3000:])
    return r.stdout


def rd(d, name, dtype):
    return np.fromfile(os.path.join(d, name), dtype=dtype)


def compare_dump(case, d, expect_kernel):
    """a `ref_harness* dump` directory against the fixture the unmodified reference produced: same checks and tolerances as
    golden_util (lnL 1e-9, per-pattern 1e-10, branch lengths 1e-6, scale_num / traversal / set of computed partials exact)"""
    g = gu.Golden(case)
    gm = g.meta
    m = json.load(open(os.path.join(d, "meta.json")))
    assert m["kernel"] == expect_kernel, m["kernel"]
    P, U, T, B = gm["nptn"], gm["nunobs"], gm["ntaxa"], gm["block"]
    NP = P + U
    for k in ("nstates", "ncat", "nmix", "family", "block", "nptn", "nunobs", "ntaxa", "nsite", "root_branch", "derv_branch"):
        assert m[k] == gm[k], (k, m[k], gm[k])
    # identical inputs: pattern order, pattern matrix, frequencies
    assert np.array_equal(rd(d, "tip_states.u8", np.uint8).reshape(T, NP), g["tip_states"])
    assert np.array_equal(rd(d, "site_pattern.i32", np.int32), g["site_pattern"])
    assert np.array_equal(rd(d, "ptn_freq.f64", np.float64), g["ptn_freq"])
    np.testing.assert_allclose(rd(d, "ptn_invar.f64", np.float64), g["ptn_invar"], rtol=1e-13, atol=0)
    # likelihood
    assert abs(m["lnl"] - gm["lnl"]) <= gu.LNL_RTOL * abs(gm["lnl"]), (m["lnl"], gm["lnl"])
    np.testing.assert_allclose(rd(d, "pattern_lh.f64", np.float64), g["pattern_lh"], rtol=gu.PTN_RTOL, atol=1e-300)
    np.testing.assert_allclose(rd(d, "pattern_lh_scaled.f64", np.float64), g["pattern_lh_scaled"], rtol=gu.PTN_RTOL, atol=1e-300)
    cat, gcat = rd(d, "pattern_lh_cat.f64", np.float64), g["pattern_lh_cat"]
    assert cat.shape == gcat.shape
    np.testing.assert_allclose(cat, gcat, rtol=1e-9, atol=1e-12 * np.abs(gcat).max())
    # the SAME set of partial vectors is valid after the call sequence (LM_PER_NODE re-orientation included), with the same contents
    idx = [l.split() for l in open(os.path.join(d, "partials.txt")) if l.strip()]
    pb = np.array([[int(a), int(b)] for a, b, _ in idx], dtype=np.int32).reshape(-1, 2)
    assert np.array_equal(pb, g["partial_branch"]), "set / order of computed partial vectors differs"
    sf = np.array([float(c) for _, _, c in idx])
    np.testing.assert_allclose(sf, g["partial_scale_factor"], rtol=1e-12, atol=1e-9)
    full = {int(i): k for k, i in enumerate(g["partial_full_index"])}
    rtol = gu.partial_rtol(g)
    for i, (a, b, _) in enumerate(idx):
        sc = rd(d, "partial_%s_%s.i16" % (a, b), np.int16)
        assert np.array_equal(sc, g["partial_scale_num"][i]), "scale_num of vector %s->%s" % (a, b)
        lh = rd(d, "partial_%s_%s.f64" % (a, b), np.float64).reshape(NP, B)
        ref_rs = g["partial_rowsum"][i]
        if i in full:
            ref = g["partial_full"][full[i]]
            scale = np.abs(ref).max(axis=1, keepdims=True) + 1e-300
            assert (np.abs(lh - ref) / scale).max() <= rtol, "partial_lh of vector %s->%s" % (a, b)
        else:
            amax = np.abs(lh).max(axis=1) * B + 1e-300
            assert (np.abs(lh.sum(axis=1) - ref_rs) / amax).max() <= rtol, "partial_lh row sums of vector %s->%s" % (a, b)
    # derivatives and lnL-from-buffer on every branch, in the reference's traversal order
    assert np.array_equal(rd(d, "traversal.i32", np.int32).reshape(-1, 2), g["traversal"])
    dv, gdv = rd(d, "branch_derv.f64", np.float64).reshape(-1, 3), g["branch_derv"].reshape(-1, 3)
    np.testing.assert_allclose(dv[:, 2], gdv[:, 2], rtol=gu.LNL_RTOL, atol=0)
    for col in (0, 1):
        np.testing.assert_allclose(dv[:, col], gdv[:, col], rtol=1e-7, atol=1e-7 * np.abs(gdv[:, col]).max())
    assert abs(m["df_root"] - gm["df_root"]) <= 1e-7 * max(1.0, np.abs(gdv[:, 0]).max())
    assert abs(m["ddf_root"] - gm["ddf_root"]) <= 1e-7 * abs(gm["ddf_root"])
    th = rd(d, "theta_root.f64", np.float64).reshape(NP, B)
    grs = g["theta_root_rowsum"]
    assert (np.abs(th.sum(axis=1) - grs) / (np.abs(th).max(axis=1) * B + 1e-300)).max() <= rtol
    # optimizeAllBranches: one pass and to convergence
    assert abs(m["oab1_lnl"] - gm["oab1_lnl"]) <= gu.LNL_RTOL * abs(gm["oab1_lnl"]), (m["oab1_lnl"], gm["oab1_lnl"])
    assert abs(m["oabfull_lnl"] - gm["oabfull_lnl"]) <= gu.LNL_RTOL * abs(gm["oabfull_lnl"]), (m["oabfull_lnl"], gm["oabfull_lnl"])
    for key, tol in (("tree_adj_oab1", gu.BRLEN_ATOL), ("tree_adj_oabfull", gu.BRLEN_ATOL)):
        a = gu.parse_adjacency(open(os.path.join(d, key + ".txt")).read())
        b = gu.parse_adjacency(g[key])
        assert a.keys() == b.keys()
        worst = max(abs(a[k] - b[k]) for k in a)
        assert worst <= tol, (key, worst)
    return m
