"""Shared parity checks: drive a PhyloTree (CUDA product or CPU oracle) through the same call sequence the
reference harness (oracle/ref_harness.cpp) ran when it produced tests/golden/*.npz, and compare.

Tolerances (BASELINE.json north_star): lnL 1e-9 relative, per-pattern log-likelihood 1e-10 relative,
optimised branch lengths 1e-6 absolute; integer bookkeeping (scale_num, root branch, traversal order, the set of
partials recomputed) exact.
"""
import glob
import json
import os

import numpy as np

import iqtree_b200
from iqtree_b200 import PhyloTree

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FAMILY = {"plain": 0, "mixrate": 1, "mixture": 2, "sitemodel": 3}

LNL_RTOL = 1e-9
PTN_RTOL = 1e-10
BRLEN_ATOL = 1e-6
PARTIAL_RTOL = 1e-9     # partial_lh entries, relative to the largest entry of the pattern's block


def partial_rtol(g):
    """The 64-state codon eigen-systems (stop codons kept at frequency 1e-4) make the eigen-space
    dot products cancel by ~4 digits: a mere change of summation order (the reference's own AVX vs generic
    kernels) moves partial_lh entries by ~1e-10 of the block maximum, while lnL stays within 1e-13."""
    return 1e-7 if g.meta["nstates"] >= 64 else PARTIAL_RTOL


def golden_cases():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


class Golden:
    def __init__(self, case):
        self.case = case
        self.z = np.load(os.path.join(GOLDEN_DIR, case + ".npz"))
        self.meta = json.loads(str(self.z["meta"]))

    def __getitem__(self, k):
        return self.z[k]

    def has(self, k):
        return k in self.z.files


def parse_adjacency(text):
    """-> {(a, b): length} for every directed pair"""
    out = {}
    for line in str(text).split("\n"):
        t = line.split()
        if not t or t[0] != "node":
            continue
        nid, deg = int(t[1]), int(t[4])
        for k in range(deg):
            out[(nid, int(t[5 + 2 * k]))] = float(t[6 + 2 * k])
    return out


def build_tree(g, lib, generic=None, devices=None, shard=None, device_newton=False):
    m = g.meta
    P, U = m["nptn"], m["nunobs"]
    t = PhyloTree(lib=lib)
    st = g["tip_states"]
    t.setAlignment(st[:, :P], g["ptn_freq"], names=[str(x) for x in g["taxa"]], seq_type=m["seq_type"],
                   nstates=m["nstates"], state_unknown=m["state_unknown"])
    t.setConstChar(g["const_char"])
    if U:
        t.setUnobservedPatterns(st[0, P:])
    t.readAdjacency(str(g["tree_adj"]))
    t.setModel(FAMILY[m["family"]], m["ncat"], m["nmix"], g["eval"], g["evec"], g["inv_evec"], g["rates"], g["props"],
               g["mixprops"], g["state_freq"], m["p_invar"])
    if m["family"] == "sitemodel":
        t.setSiteModels(g["ptn_model"])
    if generic is None:
        generic = m["kernel"] == "generic"
    t.setParams(m["min_branch_length"], m["max_branch_length"], device_newton=device_newton, generic_kernel=generic)
    if devices is not None:
        t.setDevices(devices)
    if shard is not None:
        t.setProcessShard(*shard)
    return t


def rel_close(a, b, rtol, what):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, "%s: shape %s vs %s" % (what, a.shape, b.shape)
    err = np.abs(a - b) / np.maximum(np.abs(b), 1e-300)
    bad = ~(err <= rtol)
    assert not bad.any(), "%s: max rel err %.3e (> %.1e) at %s: %r vs %r" % (
        what, np.nanmax(err), rtol, np.argwhere(bad)[0], a[bad][0], b[bad][0])


def block_close(a, b, rtol, what):
    """entries compared relative to the largest magnitude of their pattern's block"""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, "%s: shape %s vs %s" % (what, a.shape, b.shape)
    scale = np.maximum(np.abs(b).max(axis=-1, keepdims=True), 1e-300)
    err = np.abs(a - b) / scale
    assert np.all(err <= rtol), "%s: max block-relative err %.3e (> %.1e)" % (what, err.max(), rtol)


def check_inputs(t, g):
    """host logic: tip table and ptn_invar (phylotreesse.cpp:359-569)"""
    if g.meta["family"] != "sitemodel":      # site-specific models: per-pattern tip vectors are formed inside the kernels
        rel_close(t.tipPartialLh(), g["tip_partial_lh"], 1e-12, "tip_partial_lh")
    inv = t.ptnInvar()
    np.testing.assert_allclose(inv, g["ptn_invar"], rtol=1e-13, atol=0, err_msg="ptn_invar")


def check_likelihood(t, g):
    m = g.meta
    t.clearAllPartialLH()
    lnl = t.computeLikelihood()
    assert abs(lnl - m["lnl"]) <= LNL_RTOL * abs(m["lnl"]), "lnL %.12f vs reference %.12f" % (lnl, m["lnl"])
    assert list(t.rootBranch()) == list(m["root_branch"]), "root branch differs"
    lnl2, plh = t.computeLikelihood(pattern_lh=True)
    # the second evaluation finds every partial computed and runs the plain branch kernel, the first one may have evaluated the branch in
    # the tail of the traversal launch (lkcuda_partial_batch_lnl): same sums, other summation order
    assert abs(lnl2 - lnl) <= 1e-13 * abs(lnl)
    rel_close(plh, g["pattern_lh_scaled"], PTN_RTOL, "pattern_lh (scale corrected)")
    return lnl


def check_partials(t, g):
    """every partial the reference computed for the root branch: same set, scale_num exact, values close"""
    pb = g["partial_branch"]
    sc_ref = g["partial_scale_num"]
    rs_ref = g["partial_rowsum"]
    full_idx = {int(i): k for k, i in enumerate(g["partial_full_index"])}
    adj = parse_adjacency(g["tree_adj"])
    computed_ref = set((int(a), int(b)) for a, b in pb)
    ncomputed = 0
    for (a, b) in adj:
        vec, comp = t.branchVector(a, b)
        if vec >= 0 and comp:
            ncomputed += 1
            assert (a, b) in computed_ref, "partial %d->%d computed here but not by the reference" % (a, b)
    assert ncomputed == len(computed_ref), "set of computed partials differs (%d vs %d)" % (ncomputed, len(computed_ref))
    for i, (a, b) in enumerate(pb):
        vec, comp = t.branchVector(int(a), int(b))
        lh, sc = t.fetchVector(vec)
        assert np.array_equal(sc, sc_ref[i]), "scale_num of partial %d->%d differs" % (a, b)
        sf = t.vectorScaleFactor(vec)
        ref_sf = float(g["partial_scale_factor"][i])
        assert abs(sf - ref_sf) <= 1e-9 * max(1.0, abs(ref_sf)), "lh_scale_factor of %d->%d: %r vs %r" % (a, b, sf, ref_sf)
        if i in full_idx:
            block_close(lh, g["partial_full"][full_idx[i]], partial_rtol(g), "partial_lh %d->%d" % (a, b))
        else:
            scale = np.maximum(np.abs(lh).max(axis=1), 1e-300) * lh.shape[1]
            assert np.all(np.abs(lh.sum(axis=1) - rs_ref[i]) <= partial_rtol(g) * scale), "partial_lh row sums %d->%d" % (a, b)


def check_pattern_lh_cat(t, g):
    cat = t.computePatternLhCat(iqtree_b200.WSL_RATECAT)
    # the reference compacts the per-category sums in place at the front of _pattern_lh_cat (phylotree.cpp:1147-1157)
    ref = g["pattern_lh_cat"].ravel()[: cat.size].reshape(cat.shape)
    block_close(cat, ref, 0.1 * partial_rtol(g), "pattern_lh_cat")


def check_derivatives(t, g):
    m = g.meta
    t.clearAllPartialLH()
    t.computeLikelihood()
    a, b = m["derv_branch"]
    df, ddf = t.computeLikelihoodDerv(a, b, reset_theta=True)
    drt = partial_rtol(g)      # df/ddf weight the same cancelling sums by eval and eval^2
    tol = drt * max(abs(m["ddf_root"]), abs(m["df_root"]), 1.0)
    assert abs(df - m["df_root"]) <= tol and abs(ddf - m["ddf_root"]) <= tol, \
        "root-branch df/ddf %r %r vs %r %r" % (df, ddf, m["df_root"], m["ddf_root"])
    th = t.fetchTheta()
    if g.has("theta_root"):
        block_close(th, g["theta_root"], partial_rtol(g), "theta_all")
    else:
        scale = np.maximum(np.abs(th).max(axis=1), 1e-300) * th.shape[1]
        assert np.all(np.abs(th.sum(axis=1) - g["theta_root_rowsum"]) <= partial_rtol(g) * scale), "theta_all row sums"
    trav = t.getTraversal()
    assert np.array_equal(trav, g["traversal"]), "branch traversal order differs"
    ref = g["branch_derv"].reshape(-1, 3)
    for j, (n1, n2) in enumerate(trav):
        df, ddf = t.computeLikelihoodDerv(int(n1), int(n2), reset_theta=True)
        fb = t.computeLikelihoodFromBuffer()
        tol = drt * max(abs(ref[j, 1]), abs(ref[j, 0]), 1.0)
        assert abs(df - ref[j, 0]) <= tol, "df on branch %d-%d: %r vs %r" % (n1, n2, df, ref[j, 0])
        assert abs(ddf - ref[j, 1]) <= tol, "ddf on branch %d-%d: %r vs %r" % (n1, n2, ddf, ref[j, 1])
        assert abs(fb - ref[j, 2]) <= LNL_RTOL * abs(ref[j, 2]), "lnL from buffer on %d-%d: %r vs %r" % (n1, n2, fb, ref[j, 2])


def check_branch_lengths(t, adj_text, what):
    ref = parse_adjacency(adj_text)
    worst = 0.0
    for (a, b), ln in ref.items():
        worst = max(worst, abs(t.getBranchLength(a, b) - ln))
    assert worst <= BRLEN_ATOL, "%s: branch lengths differ by %.3e (> %.1e)" % (what, worst, BRLEN_ATOL)


def check_optimize(t, g, full=True):
    m = g.meta
    t.clearAllPartialLH()
    t.computeLikelihood()
    l1 = t.optimizeAllBranches(1)
    assert abs(l1 - m["oab1_lnl"]) <= LNL_RTOL * abs(m["oab1_lnl"]), "optimizeAllBranches(1): %.10f vs %.10f" % (l1, m["oab1_lnl"])
    check_branch_lengths(t, g["tree_adj_oab1"], "optimizeAllBranches(1)")
    if full:
        lf = t.optimizeAllBranches()
        assert abs(lf - m["oabfull_lnl"]) <= LNL_RTOL * abs(m["oabfull_lnl"]), "optimizeAllBranches(): %.10f vs %.10f" % (lf, m["oabfull_lnl"])
        check_branch_lengths(t, g["tree_adj_oabfull"], "optimizeAllBranches()")


def run_all_checks(t, g, full_opt=True):
    check_inputs(t, g)
    t.initializeAllPartialLh()
    check_likelihood(t, g)
    check_partials(t, g)
    check_pattern_lh_cat(t, g)
    check_derivatives(t, g)
    check_optimize(t, g, full=full_opt)
