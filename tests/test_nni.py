"""the NNI evaluation of the tree search (PhyloTree::getBestNNIForBran, phylotree.cpp:2873-3066) runs on the engine
unchanged: single-node partial updates into the nni_partial_lh vectors, theta, at most 10 Newton steps per branch, lnL from buffer.
Golden values: the UNMODIFIED reference's getBestNNIForBran on every internal branch (tests/golden/nni/*.npz, oracle/make_golden.py --nni).
Tolerances: lnL 1e-9 relative, optimised branch lengths 1e-6 (north_star), swapped subtrees and branch order exact."""
import glob
import json
import os

import numpy as np
import pytest

import golden_util as gu

NNI_DIR = os.path.join(gu.GOLDEN_DIR, "nni")
NNI_FILES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(NNI_DIR, "*.npz")))


def run_nni_checks(t, g, z, m, max_branches=None):
    t.initializeAllPartialLh()
    lnl = t.computeLikelihood()
    assert abs(lnl - m["lnl"]) <= gu.LNL_RTOL * abs(m["lnl"])
    ids, vals = z["ids"], z["vals"]
    trav = [(int(a), int(b)) for a, b in t.getTraversal()]
    leaves = set(int(l.split()[1]) for l in str(g["tree_adj"]).split("\n") if l.startswith("node ") and l.split()[2] == "1")
    internal = [(a, b) for a, b in trav if a not in leaves and b not in leaves]
    assert internal == [(int(r[0]), int(r[1])) for r in ids], "internal-branch order differs"
    step = 1 if max_branches is None else max(1, len(ids) // max_branches)
    for j in range(0, len(ids), step):
        mv = t.getBestNNIForBran(int(ids[j, 0]), int(ids[j, 1]), nni5=bool(m["nni5"]))
        for c in range(2):
            loglh, lens, n1, n2 = mv[c]
            assert (n1, n2) == (int(ids[j, 2 + 2 * c]), int(ids[j, 3 + 2 * c])), "NNI %d of branch %d swaps other subtrees" % (c, j)
            assert abs(loglh - vals[j, c, 0]) <= gu.LNL_RTOL * abs(vals[j, c, 0]), "branch %d NNI %d: %.10f vs %.10f" % (j, c, loglh, vals[j, c, 0])
            assert np.all(np.abs(lens - vals[j, c, 1:]) <= gu.BRLEN_ATOL), "branch %d NNI %d lengths %r vs %r" % (j, c, lens, vals[j, c, 1:])
    # the tree is left as it was
    adj = gu.parse_adjacency(g["tree_adj"])
    for (a, b), l in adj.items():
        assert t.getBranchLength(a, b) == l
    t.clearAllPartialLH()
    l2 = t.computeLikelihood()
    assert abs(l2 - m["lnl_after"]) <= gu.LNL_RTOL * abs(m["lnl_after"])


def _load(name):
    z = np.load(os.path.join(NNI_DIR, name + ".npz"))
    return z, json.loads(str(z["meta"])), name.split(".nni")[0]


@pytest.mark.parametrize("name", NNI_FILES)
def test_oracle_nni_matches_reference(name, oracle_lib):
    z, m, case = _load(name)
    g = gu.Golden(case)
    t = gu.build_tree(g, oracle_lib)
    run_nni_checks(t, g, z, m, max_branches=24)
    assert t.nniQueuedStats() == (0, 0)
    t.close()


@pytest.mark.parametrize("name", [n for n in NNI_FILES if n.startswith("dna_example_gtr_i_g4") or n.startswith("aa_lg_g4")])
def test_oracle_nni_queued_form(name, oracle_lib):
    """the host logic of the queued form (device length slots: lengths referenced by later ops, Newton starts and the score) on the scalar
    engine, which executes queued work at once"""
    z, m, case = _load(name)
    g = gu.Golden(case)
    t = gu.build_tree(g, oracle_lib, device_newton=True)
    run_nni_checks(t, g, z, m, max_branches=16)
    calls, fallbacks = t.nniQueuedStats()
    assert calls > 0 and fallbacks <= calls // 4
    t.close()


@pytest.mark.gpu
@pytest.mark.parametrize("nshard,device_newton", [(1, False), (1, True), (2, False)])
@pytest.mark.parametrize("name", NNI_FILES)
def test_cuda_nni_matches_reference(name, nshard, device_newton, cuda_lib):
    """device_newton: every branch optimisation is ONE persistent launch (newton_b16_kernel for DNA+G4, newton_generic_kernel elsewhere)"""
    z, m, case = _load(name)
    g = gu.Golden(case)
    t = gu.build_tree(g, cuda_lib, devices=[0] * nshard, device_newton=device_newton)
    run_nni_checks(t, g, z, m, max_branches=None if nshard == 1 else 12)
    calls, fallbacks = t.nniQueuedStats()
    if nshard == 1 and device_newton:
        assert calls > 0 and fallbacks <= calls // 4, (calls, fallbacks)      # the queued form is what ran
    else:
        assert calls == 0
    t.close()
