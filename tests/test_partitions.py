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
