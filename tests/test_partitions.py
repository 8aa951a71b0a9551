"""Partition models (BASELINE configs[4]): PhyloSuperTree::computeLikelihood sums the partition trees' lnL, calling
them concurrently from OpenMP threads (phylosupertree.cpp:969-976).  Engines must be independent (own stream, re-entrant ABI)."""
import threading

import pytest

import golden_util as gu

CASES = ["dna_example_gtr_g4", "aa_lg_g4", "dna_example_hky_r3", "aa_lg4x", "dna_example_gtr_i_g4", "codon_gy", "morph_mk_asc", "dna_deep_scaling"]


def _concurrent_partitions(lib, rounds):
    parts = []
    for c in CASES:
        g = gu.Golden(c)
        t = gu.build_tree(g, lib)
        t.initializeAllPartialLh()
        parts.append((g, t))
    errors, lnl = [], [[None] * len(parts) for _ in range(rounds)]

    def work(i):
        g, t = parts[i]
        try:
            for r in range(rounds):
                t.clearAllPartialLH()
                lnl[r][i] = t.computeLikelihood()
                trav = t.getTraversal()
                a, b = int(trav[len(trav) // 2][0]), int(trav[len(trav) // 2][1])
                t.computeLikelihoodDerv(a, b)
        except Exception as ex:      # noqa: BLE001
            errors.append((CASES[i], repr(ex)))

    th = [threading.Thread(target=work, args=(i,)) for i in range(len(parts))]
    for x in th:
        x.start()
    for x in th:
        x.join()
    assert not errors, errors
    for r in range(rounds):
        for i, (g, t) in enumerate(parts):
            ref = g.meta["lnl"]
            assert abs(lnl[r][i] - ref) <= gu.LNL_RTOL * abs(ref), (CASES[i], lnl[r][i], ref)
    total = sum(lnl[0])
    assert abs(total - sum(g.meta["lnl"] for g, _ in parts)) <= gu.LNL_RTOL * abs(total)
    for _, t in parts:
        t.close()


def test_concurrent_partitions_oracle(oracle_lib):
    _concurrent_partitions(oracle_lib, 1)


@pytest.mark.gpu
def test_concurrent_partitions_cuda(cuda_lib):
    _concurrent_partitions(cuda_lib, 5)


# ---- proportional partition model: PhyloSuperTreePlen::computeFuncDerv (phylosupertreeplen.cpp:530-578) over partition engines ----

def _make_linked_partitions(lib, devices):
    """four edge-linked partitions (2 DNA GTR+G4, 2 protein LG-like +G4) on ONE topology, each its own tree handle / engine"""
    import numpy as np
    import iqtree_b200
    from iqtree_b200 import modelinputs as mi
    ntaxa = 12
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, 3, names)
    parts = []
    for i, (S, nptn) in enumerate([(4, 900), (20, 300), (4, 500), (20, 200)]):
        st = mi.simulate_patterns(nwk, names, nptn, S, 20 + i)
        mdl = mi.gtr_g4_inputs(20 + i, ncat=4) if S == 4 else mi.random_reversible_inputs(S, 20 + i, ncat=4)
        t = iqtree_b200.PhyloTree(lib=lib)
        t.setAlignment(st, np.ones(nptn), names=names, seq_type=0 if S == 4 else 1, nstates=S, state_unknown=18 if S == 4 else 23)
        t.readTreeString(nwk)
        t.setModel(0, 4, 1, mdl["eval"], mdl["evec"], mdl["inv_evec"], mdl["rates"], mdl["props"], None, mdl["state_freq"], 0.0)
        t.setParams(1e-6, 100.0)
        if devices:
            t.setDevices([devices[i % len(devices)]])
        t.initializeAllPartialLh()
        parts.append(t)
    return parts


def _super_newton(parts, rates, a, b, super_len):
    """the super-tree's branch (a, b) under a proportional model: partition lengths = part_rate * super length; df = sum rate * df_part,
    ddf = sum rate^2 * ddf_part (phylosupertreeplen.cpp:566-568); a few plain Newton steps on the summed derivatives"""
    trace = []
    x = super_len
    for it in range(6):
        df = ddf = 0.0
        for t, r in zip(parts, rates):
            t.setBranchLength(a, b, x * r)
            d1, d2 = t.computeLikelihoodDerv(a, b, reset_theta=(it == 0))
            df += r * d1
            ddf += r * r * d2
        lnl = sum(t.computeLikelihoodFromBuffer() for t in parts)
        trace.append((x, lnl, df, ddf))
        if ddf >= 0.0:
            break
        x = min(max(x - df / ddf, 1e-6), 10.0)
    return trace


def _check_linked(parts_a, parts_b):
    rates = [0.6, 1.7, 1.0, 0.8]
    trav = parts_a[0].getTraversal()
    a, b = int(trav[len(trav) // 2][0]), int(trav[len(trav) // 2][1])
    for ps in (parts_a, parts_b):
        for t in ps:
            t.clearAllPartialLH()
            t.computeLikelihood()
    ta, tb = _super_newton(parts_a, rates, a, b, 0.07), _super_newton(parts_b, rates, a, b, 0.07)
    assert len(ta) == len(tb) and len(ta) >= 3
    for (xa, la, dfa, ddfa), (xb, lb, dfb, ddfb) in zip(ta, tb):
        assert abs(xa - xb) <= 1e-9 * max(1.0, abs(xb))
        assert abs(la - lb) <= gu.LNL_RTOL * abs(lb)
        assert abs(dfa - dfb) <= 1e-8 * max(1.0, abs(dfb), abs(ddfb) * 1e-3) and abs(ddfa - ddfb) <= 1e-8 * max(1.0, abs(ddfb))
    # the summed first derivative is the derivative of the summed lnL with respect to the super length
    x0, h = ta[0][0], 1e-5

    def total(x):
        s = 0.0
        for t, r in zip(parts_a, rates):
            t.setBranchLength(a, b, x * r)
            s += t.computeLikelihoodBranch(a, b)
        return s
    fd = (total(x0 + h) - total(x0 - h)) / (2 * h)
    assert abs(fd - ta[0][2]) <= 1e-4 * max(1.0, abs(fd)), (fd, ta[0][2])
    assert ta[-1][1] >= ta[0][1] - 1e-9      # Newton on the summed derivatives does not lose likelihood here
    for t in parts_a + parts_b:
        t.close()


def test_proportional_partitions_oracle(oracle_lib):
    _check_linked(_make_linked_partitions(oracle_lib, None), _make_linked_partitions(oracle_lib, None))


@pytest.mark.gpu
def test_proportional_partitions_cuda_over_devices(cuda_lib, oracle_lib):
    """partition engines spread over every GPU of the box (second axis: whole partitions -> GPUs); the host adds
    part_rate * df and part_rate^2 * ddf exactly as PhyloSuperTreePlen::computeFuncDerv does - no collective"""
    import torch
    ndev = max(1, min(torch.cuda.device_count(), 4))
    _check_linked(_make_linked_partitions(cuda_lib, list(range(ndev))), _make_linked_partitions(oracle_lib, None))
