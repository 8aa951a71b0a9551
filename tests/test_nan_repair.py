"""The NaN / Inf repair of the branch log-likelihood (phylokernel.h:848-866; LKCUDA_R_NREPAIRED): a pattern whose likelihood is 0 gets
log(0) = -inf, the reference then sets _pattern_lh to 4 * LOG_SCALING_THRESHOLD = log(2^-1024) and re-sums the tree likelihood.

Fixtures (tests/golden/repair, `ref_harness dumprepair`): the UNMODIFIED reference with three patterns' rows of the root branch's
vector zeroed on the host before its own computeLikelihoodBranch ran.  The same is done here through the C ABI (fetch / zero / store the
vector, lkcuda_branch_lnl): the tip-end kernels (DNA: branch_tip_b16_kernel, others: branch_kernel<TIP>), and - against the CPU
restatement - an internal branch (branch_kernel<!TIP>) and the lnL-from-buffer kernel.
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

import golden_util as gu

LOG_THR = -177.44567822334599
CASES = ["dna_example_gtr_g4", "aa_lg_g4", "dna_deep_scaling", "codon_gy"]


def _tip(taxon):
    return -1 - taxon


def _run_repair(lib, case):
    from iqtree_b200 import phylotree as pt
    z = np.load(os.path.join(gu.GOLDEN_DIR, "repair", case + ".npz"))
    m = json.loads(str(z["meta"]))
    g = gu.Golden(case)
    t = gu.build_tree(g, lib)
    t.initializeAllPartialLh()
    t.clearAllPartialLH()
    lnl = t.computeLikelihood()
    assert abs(lnl - m["lnl"]) <= gu.LNL_RTOL * abs(m["lnl"])
    leaf, inner = m["vector"]
    assert list(t.rootBranch()) == [leaf, inner] or list(t.rootBranch()) == [inner, leaf]
    vec, comp = t.branchVector(leaf, inner)
    assert vec >= 0 and comp
    lh, sc = t.fetchVector(vec)
    P, U = g.meta["nptn"], g.meta["nunobs"]
    e = t.engine(0)
    full = np.zeros(((P + U), lh.shape[1]))
    full[:P] = lh
    for p in m["zeroed"]:
        full[p] = 0.0
    rc = lib.lkcuda_store_vector(e, vec, full.ctypes.data_as(pt.c_double_p), None)
    assert rc == 0, lib.lkcuda_last_error(e)
    r = (C.c_double * 8)()
    rc = lib.lkcuda_branch_lnl(e, _tip(leaf), vec, t.getBranchLength(leaf, inner), r)
    assert rc == 0, lib.lkcuda_last_error(e)
    assert r[4] == 3.0, "LKCUDA_R_NREPAIRED = %r" % r[4]
    assert abs(r[0] - m["lnl_repaired"]) <= gu.LNL_RTOL * abs(m["lnl_repaired"]), (r[0], m["lnl_repaired"])
    plh = np.zeros(P)
    assert lib.lkcuda_fetch_pattern_lh(e, plh.ctypes.data_as(pt.c_double_p)) == 0
    for p in m["zeroed"]:
        assert plh[p] == 4.0 * LOG_THR
    gu.rel_close(plh, z["pattern_lh_repaired"], gu.PTN_RTOL, "_pattern_lh after the repair")
    return t, m


@pytest.mark.parametrize("case", CASES)
def test_oracle_repair_matches_reference(case, oracle_lib):
    t, _ = _run_repair(oracle_lib, case)
    t.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_cuda_repair_matches_reference(case, cuda_lib):
    t, _ = _run_repair(cuda_lib, case)
    t.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["dna_example_gtr_g4", "aa_lg_g4", "aa_lg_c20_g4"])
def test_cuda_repair_internal_branch_and_from_buffer(case, cuda_lib, oracle_lib):
    """an internal branch (both ends vectors) and lnL-from-buffer: CUDA against the CPU restatement pinned above"""
    from iqtree_b200 import phylotree as pt
    g = gu.Golden(case)
    out = []
    for lib in (oracle_lib, cuda_lib):
        t = gu.build_tree(g, lib)
        t.initializeAllPartialLh()
        t.clearAllPartialLH()
        t.computeLikelihood()
        trav = t.getTraversal()
        nleaf = g.meta["ntaxa"]
        a, b = [(int(x), int(y)) for x, y in trav if x >= nleaf and y >= nleaf][len(trav) // 5]
        t.computeLikelihoodBranch(a, b)                 # both directions of the branch are valid now
        va, ca = t.branchVector(a, b)
        vb, cb = t.branchVector(b, a)
        assert ca and cb
        P, U = g.meta["nptn"], g.meta["nunobs"]
        lh, _ = t.fetchVector(va)
        full = np.zeros((P + U, lh.shape[1]))
        full[:P] = lh
        zero = [1, P // 3, P - 2]
        full[zero] = 0.0
        e = t.engine(0)
        assert lib.lkcuda_store_vector(e, va, full.ctypes.data_as(pt.c_double_p), None) == 0
        r = (C.c_double * 8)()
        ln = t.getBranchLength(a, b)
        assert lib.lkcuda_branch_lnl(e, vb, va, ln, r) == 0, lib.lkcuda_last_error(e)
        res = [r[0], r[4]]
        assert lib.lkcuda_theta_build(e, vb, va) == 0
        assert lib.lkcuda_lnl_from_buffer(e, ln, r) == 0
        res += [r[0], r[4]]
        plh = np.zeros(P)
        assert lib.lkcuda_fetch_pattern_lh(e, plh.ctypes.data_as(pt.c_double_p)) == 0
        out.append((res, plh, zero))
        t.close()
    (ro, po, zero), (rc, pc, _) = out
    assert ro[1] == rc[1] == 3.0 and ro[3] == rc[3] == 3.0
    assert abs(rc[0] - ro[0]) <= gu.LNL_RTOL * abs(ro[0]) and abs(rc[2] - ro[2]) <= gu.LNL_RTOL * abs(ro[2])
    assert all(pc[p] == 4.0 * LOG_THR for p in zero)
    gu.rel_close(pc, po, gu.PTN_RTOL, "_pattern_lh after the repair (from buffer)")
