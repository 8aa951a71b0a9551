"""Fitch parsimony (PhyloTree::computeParsimony / computeParsimonyBranch / fixNegativeBranch / computeParsimonyTree)
against golden values the UNMODIFIED reference produced (tests/golden/pars/*.npz, written by `python oracle/make_golden.py --pars`
through oracle/_ref/ref_harness dumppars).

Everything is integer work: scores, per-branch substitution counts and the stepwise-addition tree (topology AND stored neighbour order,
i.e. every tie broken as the reference breaks it) must be identical; the Jukes-Cantor corrected branch lengths of fixNegativeBranch are
one log() of an integer ratio and are compared to 1e-12 relative.
"""
import ctypes as C
import glob
import json
import os

import numpy as np
import pytest

import golden_util as gu
import iqtree_b200
from iqtree_b200 import phylotree as pt

PARS_DIR = os.path.join(gu.GOLDEN_DIR, "pars")
PARS_CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(PARS_DIR, "*.npz")))


def load(case):
    z = np.load(os.path.join(PARS_DIR, case + ".npz"))
    return z, json.loads(str(z["meta"]))


def parse_adjacency(text):
    """-> (root id, {node id: (is_leaf, name, [(neighbour id, length), ...])})"""
    lines = [l.split() for l in text.strip().splitlines()]
    root = int(lines[0][1])
    nodes = {}
    for t in lines[1:]:
        deg = int(t[4])
        nodes[int(t[1])] = (int(t[2]), t[3], [(int(t[5 + 2 * k]), float(t[6 + 2 * k])) for k in range(deg)])
    return root, nodes


def names_of(adj_text, ntaxa):
    _, nodes = parse_adjacency(adj_text)
    names = [None] * ntaxa
    for nid, (leaf, name, _) in nodes.items():
        if leaf:
            names[nid] = name
    assert all(n is not None for n in names)
    return names


def build(z, m, lib):
    t = iqtree_b200.PhyloTree(lib=lib)
    nt = m["ntaxa"]
    # the likelihood alignment is not needed for parsimony: one dummy pattern
    t.setAlignment(np.zeros((nt, 1), dtype=np.uint8), np.ones(1), names=names_of(str(z["tree"]), nt), seq_type=m["seq_type"],
                   nstates=m["nstates"], state_unknown=m["state_unknown"])
    t.setParams(m["min_branch_length"], 100.0)
    t.setParsimonySites(z["sites"], m["aln_nsite"])
    return t


def check_fixed_tree(t, z, m):
    t.readAdjacency(str(z["tree"]))
    assert t.computeParsimony() == m["score"]
    br = z["branches"]
    mine = t.getBranches()
    assert np.array_equal(mine, br[:, :2]), "MTree::getBranches order differs"
    for n1, n2, score, subst in br:
        assert t.computeParsimonyBranch(int(n1), int(n2)) == (int(score), int(subst)), (n1, n2)
    # fixNegativeBranch(true): lengths from the substitution counts (phylotree.cpp:2669-2675)
    assert t.fixNegativeBranch(True) == len(br)
    S, nsite = m["nstates"], m["aln_nsite"]
    zc = S / (S - 1.0)
    for n1, n2, score, subst in br:
        bl = (subst if subst > 0 else 1) / nsite
        x = 1.0 - zc * bl
        if x > 0:
            bl = -np.log(x) / zc
        bl = max(bl, m["min_branch_length"])
        assert abs(t.getBranchLength(int(n1), int(n2)) - bl) <= 1e-12 * bl


def check_stepwise(t, z, m):
    score = t.computeParsimonyTree(z["order"])
    assert score == m["stepwise_score"], (score, m["stepwise_score"])
    r0, ref = parse_adjacency(str(z["stepwise_tree"]))
    r1, got = parse_adjacency(t.getAdjacency())
    assert r0 == r1 and set(ref) == set(got)
    for nid in ref:
        assert ref[nid][0] == got[nid][0] and ref[nid][1] == got[nid][1], nid
        assert [k for k, _ in ref[nid][2]] == [k for k, _ in got[nid][2]], "neighbour order of node %d differs" % nid
        for (_, a), (_, b) in zip(ref[nid][2], got[nid][2]):
            assert abs(a - b) <= 1e-12 * abs(a), (nid, a, b)
    assert t.computeParsimony() == m["stepwise_check"]


@pytest.mark.parametrize("case", PARS_CASES)
def test_oracle_parsimony_matches_reference(case, oracle_lib):
    z, m = load(case)
    t = build(z, m, oracle_lib)
    check_fixed_tree(t, z, m)
    check_stepwise(t, z, m)
    t.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", PARS_CASES)
def test_cuda_parsimony_matches_reference(case, cuda_lib):
    z, m = load(case)
    t = build(z, m, cuda_lib)
    check_fixed_tree(t, z, m)
    check_stepwise(t, z, m)
    launches, _ = t.parsLaunches()
    assert launches > 0
    t.close()


def _run_ops(lib, e, ops):
    arr = (C.c_int32 * (5 * len(ops)))(*[v for op in ops for v in op])
    counts = (C.c_uint32 * (2 * len(ops)))()
    rc = lib.lkpars_run(e, arr, len(ops), counts)
    return rc, np.array(counts, dtype=np.uint32).reshape(-1, 2)


def _random_engine(lib, rng, S, ntaxa, nsites, nvec):
    nmask = S + 2
    mask = np.array([1 << s for s in range(S)] + [3, (1 << S) - 1 if S < 64 else 2 ** 64 - 1], dtype=np.uint64)
    sites = rng.integers(0, nmask, size=(ntaxa, nsites)).astype(np.uint8)
    e = C.c_void_p()
    rc = lib.lkpars_create(0, S, ntaxa, nsites, sites.ctypes.data_as(pt.c_uint8_p), mask.ctypes.data_as(pt.c_uint64_p), nmask, nvec, C.byref(e))
    assert rc == 0, lib.lkpars_last_error(None)
    return e, sites, mask


@pytest.mark.gpu
@pytest.mark.parametrize("S,nsites", [(4, 1), (4, 31), (4, 4097), (20, 333), (64, 200), (5, 1000), (2, 65), (4, 0)])
def test_cuda_engine_matches_oracle_engine(S, nsites, cuda_lib, oracle_lib):
    """lkpars_run on random ops (dependent chains, insertions with and without a stored result, a read-only tail): counts and every
    vector bit-identical between the CUDA engine and the CPU restatement, for word-boundary site counts and all kernel variants"""
    rng = np.random.default_rng(S * 1000 + nsites)
    ntaxa, nvec = 9, 24
    ops = []
    avail = list(range(ntaxa))
    for d in range(ntaxa, ntaxa + nvec - 4):
        a, b = rng.choice(avail, 2, replace=False)
        ops.append((0, d, int(a), int(b), -1))
        avail.append(d)
    ops.append((2, ntaxa + nvec - 1, 3, ntaxa + 2, 5))                  # insertion that keeps its Fitch set
    avail.append(ntaxa + nvec - 1)
    for _ in range(40):                                                 # the read-only tail
        a, b, c = rng.choice(avail, 3, replace=False)
        ops.append((int(rng.integers(1, 3)), -1, int(a), int(b), int(c)))
    res = []
    for lib in (cuda_lib, oracle_lib):
        e, _, _ = _random_engine(lib, np.random.default_rng(7), S, ntaxa, nsites, nvec)
        rc, counts = _run_ops(lib, e, ops)
        assert rc == 0, lib.lkpars_last_error(e)
        nw = C.c_int64()
        lib.lkpars_words(e, C.byref(nw))
        vecs = np.zeros((ntaxa + nvec, S, nw.value), dtype=np.uint32)
        for v in list(range(ntaxa)) + [op[1] for op in ops if op[1] >= 0]:
            assert lib.lkpars_fetch_vector(e, v, vecs[v].ctypes.data_as(C.POINTER(C.c_uint32))) == 0
        res.append((counts, vecs))
        lib.lkpars_destroy(e)
    assert np.array_equal(res[0][0], res[1][0])
    assert np.array_equal(res[0][1], res[1][1])


@pytest.mark.gpu
def test_cuda_engine_long_chain(cuda_lib, oracle_lib):
    """more dependent ops than one chain launch holds (4096): split launches, groups never straddle the split"""
    rng = np.random.default_rng(11)
    ntaxa, nvec = 6, 4500
    ops, avail = [], list(range(ntaxa))
    for d in range(ntaxa, ntaxa + nvec):
        a, b = rng.choice(avail[-40:], 2, replace=False)
        ops.append((0, d, int(a), int(b), -1))
        avail.append(d)
    ops += [(1, -1, ntaxa + nvec - 1, 0, -1), (2, -1, ntaxa + nvec - 2, 1, 2)]
    res = []
    for lib in (cuda_lib, oracle_lib):
        e, _, _ = _random_engine(lib, np.random.default_rng(3), 4, ntaxa, 777, nvec)
        rc, counts = _run_ops(lib, e, ops)
        assert rc == 0, lib.lkpars_last_error(e)
        res.append(counts)
        lib.lkpars_destroy(e)
    assert np.array_equal(res[0], res[1])


def test_engine_argument_checks(oracle_lib):
    lib = oracle_lib
    e = C.c_void_p()
    mask = np.array([1, 2, 4, 8, 15], dtype=np.uint64)
    sites = np.zeros((3, 5), dtype=np.uint8)
    args = (sites.ctypes.data_as(pt.c_uint8_p), mask.ctypes.data_as(pt.c_uint64_p))
    assert lib.lkpars_create(0, 1, 3, 5, args[0], args[1], 5, 2, C.byref(e)) != 0          # nstates < 2
    assert lib.lkpars_create(0, 4, 3, 5, args[0], args[1], 3, 2, C.byref(e)) != 0          # fewer masks than states
    assert lib.lkpars_create(0, 4, 3, 5, args[0], args[1], 5, 2, C.byref(e)) == 0
    for bad in [(0, 1, 0, 1, -1), (0, 3, 0, 9, -1), (0, 3, 3, 1, -1), (2, -1, 0, 1, 7), (5, 3, 0, 1, -1)]:
        rc, _ = _run_ops(lib, e, [bad])
        assert rc != 0, bad
    rc, counts = _run_ops(lib, e, [(0, 3, 0, 1, -1), (1, -1, 3, 2, -1)])
    assert rc == 0 and counts[0, 0] == 0 and counts[1, 0] == 0                               # identical tips: no union site
    lib.lkpars_destroy(e)


def test_library_exports_lkpars_symbols():
    from iqtree_b200 import _build
    lib = C.CDLL(_build.build_cuda())
    for name in pt.LKPARS_SYMBOLS:
        assert hasattr(lib, name), name


@pytest.mark.gpu
@pytest.mark.parametrize("S,seq_type,unknown,ntaxa,nsites", [(4, 0, 18, 70, 6000), (20, 1, 23, 40, 2500)])
def test_cuda_stepwise_addition_matches_oracle_at_size(S, seq_type, unknown, ntaxa, nsites, cuda_lib, oracle_lib):
    """a whole stepwise-addition start tree on random data with ambiguity codes, larger than the reference fixtures (hundreds of ops per
    launch, several grid rows per SM in the wide kernel): the CUDA engine and the CPU restatement must build the same tree"""
    rng = np.random.default_rng(ntaxa * 7 + S)
    base = rng.integers(0, S, size=nsites)
    sites = np.empty((ntaxa, nsites), dtype=np.uint8)
    for t in range(ntaxa):                     # noisy copies of one sequence, so that the tree has signal and ties
        row = base.copy()
        flip = rng.random(nsites) < 0.15
        row[flip] = rng.integers(0, S, size=int(flip.sum()))
        amb = rng.random(nsites) < 0.02
        row[amb] = rng.integers(S, unknown + 1, size=int(amb.sum())) if seq_type == 0 else unknown
        sites[t] = row
    order = rng.permutation(ntaxa).astype(np.int32)
    names = ["t%d" % i for i in range(ntaxa)]
    res = []
    for lib in (cuda_lib, oracle_lib):
        t = iqtree_b200.PhyloTree(lib=lib)
        t.setAlignment(np.zeros((ntaxa, 1), dtype=np.uint8), np.ones(1), names=names, seq_type=seq_type, nstates=S, state_unknown=unknown)
        t.setParsimonySites(sites, nsites)
        score = t.computeParsimonyTree(order)
        res.append((score, t.getAdjacency(), t.computeParsimony()))
        t.close()
    assert res[0] == res[1]
    assert res[0][0] == res[0][2]
