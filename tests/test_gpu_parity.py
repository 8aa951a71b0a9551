"""GPU suite (-m gpu): the CUDA engine, driven through the C ABI, against the golden fixtures of the unmodified
reference AND against the CPU oracle on the same inputs."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

import numpy as np
import pytest

import golden_util as gu
import iqtree_b200
from iqtree_b200 import modelinputs as mi

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", gu.golden_cases())
def test_cuda_matches_reference(case, cuda_lib):
    g = gu.Golden(case)
    t = gu.build_tree(g, cuda_lib)
    gu.run_all_checks(t, g, full_opt=g.meta["block"] <= 256)
    assert t.launchCount() > 0
    t.close()


@pytest.mark.parametrize("case", ["dna_example_gtr_g4", "dna_deep_scaling", "aa_lg_g4", "morph_mk_asc", "codon_gy", "aa_lg4x", "aa_lg_c20_g4", "bin_jc2_g4", "dna_example_hky_r3"])
def test_cuda_device_newton(case, cuda_lib):
    """optimizeAllBranches with the Newton loop driven inside the engine gives the same lengths"""
    g = gu.Golden(case)
    t = gu.build_tree(g, cuda_lib, device_newton=True)
    t.initializeAllPartialLh()
    gu.check_optimize(t, g, full=True)
    t.close()


@pytest.mark.parametrize("case", ["dna_example_gtr_g4", "aa_lg_g4", "codon_gy", "dna_example_hky_r3"])
def test_generic_kernel_flag(case, cuda_lib):
    """the -fastlk semantics (zero guard, _pattern_lh_cat filled by every branch call) on non-generic fixtures"""
    g = gu.Golden(case)
    t = gu.build_tree(g, cuda_lib, generic=True)
    t.initializeAllPartialLh()
    gu.check_likelihood(t, g)
    t.close()


def _synthetic(lib, ntaxa, nptn, nstates, ncat, seed, state_unknown):
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, seed, names)
    st = mi.simulate_patterns(nwk, names, nptn, nstates, seed)
    rng = np.random.default_rng(seed)
    st[rng.integers(0, ntaxa, nptn // 50), rng.integers(0, nptn, nptn // 50)] = state_unknown
    freq = rng.integers(1, 4, nptn).astype(np.float64)
    mdl = mi.gtr_g4_inputs(seed, ncat=ncat) if nstates == 4 else mi.random_reversible_inputs(nstates, seed, ncat=ncat)
    t = iqtree_b200.PhyloTree(lib=lib)
    t.setAlignment(st, freq, names=names, seq_type=0 if nstates == 4 else (1 if nstates == 20 else 4), nstates=nstates,
                   state_unknown=state_unknown)
    t.readTreeString(nwk)
    t.setModel(0, ncat, 1, mdl["eval"], mdl["evec"], mdl["inv_evec"], mdl["rates"], mdl["props"], None, mdl["state_freq"], 0.0)
    return t


@pytest.mark.parametrize("shape", [(24, 5000, 4, 4, 18), (150, 2000, 4, 4, 18), (12, 700, 20, 4, 23), (10, 300, 64, 1, 64),
                                   (9, 257, 4, 8, 18), (7, 100, 7, 3, 7), (6, 33, 4, 1, 18), (5, 1, 4, 4, 18)])
def test_cuda_matches_oracle_synthetic(shape, cuda_lib, oracle_lib):
    """same seeded inputs through the CUDA engine and through the CPU oracle: lnL, pattern lh, scale_num, df/ddf, OAB"""
    ntaxa, nptn, S, C, SU = shape
    a = _synthetic(cuda_lib, ntaxa, nptn, S, C, 3, SU)
    b = _synthetic(oracle_lib, ntaxa, nptn, S, C, 3, SU)
    for t in (a, b):
        t.initializeAllPartialLh()
    la, pa = a.computeLikelihood(pattern_lh=True)
    lb, pb = b.computeLikelihood(pattern_lh=True)
    assert abs(la - lb) <= gu.LNL_RTOL * abs(lb)
    gu.rel_close(pa, pb, gu.PTN_RTOL, "pattern_lh")
    assert a.rootBranch() == b.rootBranch()
    trav = a.getTraversal()
    assert np.array_equal(trav, b.getTraversal())
    for (n1, n2) in trav[:: max(1, len(trav) // 12)]:
        va, ca = a.branchVector(int(n1), int(n2))
        vb, cb = b.branchVector(int(n1), int(n2))
        assert va == vb and ca == cb
        da, db = a.computeLikelihoodDerv(int(n1), int(n2)), b.computeLikelihoodDerv(int(n1), int(n2))
        tol = (1e-7 if S >= 64 else 1e-9) * max(abs(db[0]), abs(db[1]), 1.0)
        assert abs(da[0] - db[0]) <= tol and abs(da[1] - db[1]) <= tol
        if va >= 0 and ca:
            (lha, sca), (lhb, scb) = a.fetchVector(va), b.fetchVector(vb)
            assert np.array_equal(sca, scb)
            gu.block_close(lha, lhb, 1e-7 if S >= 64 else 1e-9, "partial_lh")
    oa, ob = a.optimizeAllBranches(2), b.optimizeAllBranches(2)
    assert abs(oa - ob) <= gu.LNL_RTOL * abs(ob)
    for (n1, n2) in trav:
        assert abs(a.getBranchLength(int(n1), int(n2)) - b.getBranchLength(int(n1), int(n2))) <= gu.BRLEN_ATOL
    a.close(); b.close()


@pytest.mark.parametrize("host_sum", [False, True])
@pytest.mark.parametrize("case,nshard", [("dna_example_gtr_i_g4", 2), ("dna_example_gtr_g4", 4), ("morph_mk_asc", 3), ("aa_lg_g4", 2),
                                         ("aa_sitefreq_lg_g4", 3), ("dna_sitefreq_deep_scaling", 2)])
def test_shards_one_device(case, nshard, host_sum, cuda_lib, monkeypatch):
    """pattern sharding: N engines on the same device (each on its own stream) give the sums of one.  Default: the reduction kernels
    exchange and add their result blocks among themselves through the peer mailboxes (fused all-reduce, lkcuda_attach_peers);
    LKTREE_NO_PEERS=1: host-side sum of the shards."""
    if host_sum:
        monkeypatch.setenv("LKTREE_NO_PEERS", "1")
    g = gu.Golden(case)
    t = gu.build_tree(g, cuda_lib, devices=[0] * nshard)
    t.initializeAllPartialLh()
    assert t.peersAttached() == (not host_sum)
    gu.check_likelihood(t, g)
    gu.check_derivatives(t, g)
    gu.check_optimize(t, g, full=False)
    t.close()


def test_wait_result_matches_read_result(cuda_lib):
    """lkcuda_wait_result (host-mapped block + sequence number, no copy) returns what lkcuda_read_result copies back"""
    import ctypes
    g = gu.Golden("dna_example_gtr_g4")
    t = gu.build_tree(g, cuda_lib)
    t.initializeAllPartialLh()
    lnl = t.computeLikelihood()
    e = t.engine(0)
    a = (ctypes.c_double * 8)(); b = (ctypes.c_double * 8)()
    for _ in range(3):
        assert cuda_lib.lkcuda_wait_result(e, a) == 0
        assert cuda_lib.lkcuda_read_result(e, b) == 0
        assert list(a) == list(b) and a[0] == lnl
    t.close()


def test_peer_attach_argument_checks(cuda_lib):
    import ctypes
    g = gu.Golden("dna_example_gtr_g4")
    t = gu.build_tree(g, cuda_lib)
    t.initializeAllPartialLh()
    e = t.engine(0)
    assert cuda_lib.lkcuda_attach_peers(e, 0, 2, None, None) != 0          # no mailbox yet
    h = (ctypes.c_uint8 * 64)(); p = ctypes.c_void_p()
    assert cuda_lib.lkcuda_peer_handle(e, h, ctypes.byref(p)) == 0 and p.value
    assert any(h)
    assert cuda_lib.lkcuda_attach_peers(e, 0, 17, None, None) != 0         # more than LKCUDA_MAX_PEERS
    assert cuda_lib.lkcuda_attach_peers(e, 2, 2, None, None) != 0          # rank out of range
    assert cuda_lib.lkcuda_attach_peers(e, 0, 2, None, None) != 0          # no addresses
    assert cuda_lib.lkcuda_attach_peers(e, 0, 1, None, None) == 0          # a world of one: nothing to exchange
    assert abs(t.computeLikelihood() - g.meta["lnl"]) <= gu.LNL_RTOL * abs(g.meta["lnl"])
    t.close()


@pytest.mark.parametrize("maxslots", [0, 1, 2])
@pytest.mark.parametrize("case", ["dna_example_gtr_g4", "dna_deep_scaling", "dna_example_hky_r3", "dna_example_jc"])
def test_traversal_kernel_slot_spill(case, maxslots, cuda_lib, monkeypatch):
    """whole-traversal kernel with the shared-memory slots capped: spilled results are re-read from the arena"""
    monkeypatch.setenv("LKCUDA_TRV_MAXSLOTS", str(maxslots))
    g = gu.Golden(case)
    t = gu.build_tree(g, cuda_lib)
    t.initializeAllPartialLh()
    gu.check_likelihood(t, g)
    gu.check_partials(t, g)
    gu.check_derivatives(t, g)
    t.close()


@pytest.mark.parametrize("maxslots", [-1, 0, 1])
@pytest.mark.parametrize("case", ["aa_lg_g4", "aa_lg4x"])
def test_protein_mma_kernel(case, maxslots, cuda_lib, monkeypatch):
    """protein partials on the FP64 tensor cores (mma.sync m8n8k4), with all / no / one shared-memory slot"""
    monkeypatch.setenv("LKCUDA_TRV_MAXSLOTS", str(maxslots))
    g = gu.Golden(case)
    t = gu.build_tree(g, cuda_lib)
    t.initializeAllPartialLh()
    gu.check_likelihood(t, g)
    gu.check_partials(t, g)
    gu.check_derivatives(t, g)
    gu.check_optimize(t, g, full=False)
    t.close()


def test_protein_generic_kernel_still_covered(cuda_lib, monkeypatch):
    """LKCUDA_NO_MMA=1 sends 20-state data through the generic kernel (the path every other state count takes)"""
    monkeypatch.setenv("LKCUDA_NO_MMA", "1")
    g = gu.Golden("aa_lg_g4")
    t = gu.build_tree(g, cuda_lib)
    t.initializeAllPartialLh()
    gu.check_likelihood(t, g)
    gu.check_partials(t, g)
    t.close()


def test_protein_mma_matches_oracle_large(cuda_lib, oracle_lib):
    """more patterns than one tile per SM, ragged last tile, unknown states"""
    a = _synthetic(cuda_lib, 30, 64 * 148 + 37, 20, 4, 5, 23)
    b = _synthetic(oracle_lib, 30, 64 * 148 + 37, 20, 4, 5, 23)
    for t in (a, b):
        t.initializeAllPartialLh()
    la, pa = a.computeLikelihood(pattern_lh=True)
    lb, pb = b.computeLikelihood(pattern_lh=True)
    assert abs(la - lb) <= gu.LNL_RTOL * abs(lb)
    gu.rel_close(pa, pb, gu.PTN_RTOL, "pattern_lh")
    trav = a.getTraversal()
    for (n1, n2) in trav[::7]:
        va, ca = a.branchVector(int(n1), int(n2))
        vb, cb = b.branchVector(int(n1), int(n2))
        assert va == vb and ca == cb
        if va >= 0 and ca:
            (lha, sca), (lhb, scb) = a.fetchVector(va), b.fetchVector(vb)
            assert np.array_equal(sca, scb)
            gu.block_close(lha, lhb, 1e-9, "partial_lh")
    a.close(); b.close()


@pytest.mark.parametrize("maxslots", [-1, 0])
@pytest.mark.parametrize("case", ["codon_gy", "codon_mg_g4", "aa_lg_c20_g4"])
def test_codon_mixture_mma_kernel(case, maxslots, cuda_lib, monkeypatch):
    """64-state codon and wide-block protein mixtures on the FP64 tensor cores, one (node, block) pair per ring stage"""
    monkeypatch.setenv("LKCUDA_TRV_MAXSLOTS", str(maxslots))
    g = gu.Golden(case)
    t = gu.build_tree(g, cuda_lib)
    t.initializeAllPartialLh()
    gu.check_likelihood(t, g)
    gu.check_partials(t, g)
    gu.check_derivatives(t, g)
    gu.check_optimize(t, g, full=False)
    t.close()


@pytest.mark.parametrize("case", ["codon_gy", "aa_lg_c20_g4"])
def test_codon_mixture_generic_kernel_still_covered(case, cuda_lib, monkeypatch):
    monkeypatch.setenv("LKCUDA_NO_MMA", "1")
    g = gu.Golden(case)
    t = gu.build_tree(g, cuda_lib)
    t.initializeAllPartialLh()
    gu.check_likelihood(t, g)
    gu.check_partials(t, g)
    t.close()


@pytest.mark.parametrize("shape", [(12, 64 * 148 + 37, 64, 1, 64), (8, 1500, 64, 4, 64), (10, 3000, 20, 8, 23)])
def test_mmak_matches_oracle_large(shape, cuda_lib, oracle_lib):
    """more patterns than one tile per SM, ragged last tile, unknown states, several blocks per pattern"""
    ntaxa, nptn, S, C, SU = shape
    a = _synthetic(cuda_lib, ntaxa, nptn, S, C, 5, SU)
    b = _synthetic(oracle_lib, ntaxa, nptn, S, C, 5, SU)
    for t in (a, b):
        t.initializeAllPartialLh()
    la, pa = a.computeLikelihood(pattern_lh=True)
    lb, pb = b.computeLikelihood(pattern_lh=True)
    assert abs(la - lb) <= gu.LNL_RTOL * abs(lb)
    gu.rel_close(pa, pb, gu.PTN_RTOL, "pattern_lh")
    trav = a.getTraversal()
    for (n1, n2) in trav[::3]:
        va, ca = a.branchVector(int(n1), int(n2))
        vb, cb = b.branchVector(int(n1), int(n2))
        assert va == vb and ca == cb
        if va >= 0 and ca:
            (lha, sca), (lhb, scb) = a.fetchVector(va), b.fetchVector(vb)
            assert np.array_equal(sca, scb)
            gu.block_close(lha, lhb, 1e-7 if S >= 64 else 1e-9, "partial_lh")
    a.close(); b.close()


@pytest.mark.parametrize("nchunks", [1, 3, 8])
def test_pipelined_tip_state_upload(nchunks, cuda_lib, oracle_lib):
    """lkcuda_set_tip_states_async: the traversal kernel polls the per-chunk counter; results equal the plain upload's and the
    oracle's; a second upload with different states is picked up; an out-of-range state is reported at the next read-back"""
    import ctypes
    import torch
    ntaxa, nptn = 40, 148 * 352 * 2 + 77
    a = _synthetic(cuda_lib, ntaxa, nptn, 4, 4, 11, 18)
    b = _synthetic(oracle_lib, ntaxa, nptn, 4, 4, 11, 18)
    for t in (a, b):
        t.initializeAllPartialLh()
    l0 = a.computeLikelihood()
    lb = b.computeLikelihood()
    assert abs(l0 - lb) <= gu.LNL_RTOL * abs(lb)
    names = ["t%d" % i for i in range(ntaxa)]
    st = mi.simulate_patterns(mi.random_tree_newick(ntaxa, 11, names), names, nptn, 4, 11)
    rng = np.random.default_rng(11)
    st[rng.integers(0, ntaxa, nptn // 50), rng.integers(0, nptn, nptn // 50)] = 18
    pinned = torch.from_numpy(np.ascontiguousarray(st)).pin_memory()
    eng = a.engine(0)
    u8p = ctypes.POINTER(ctypes.c_uint8)
    for rep in range(3):
        assert cuda_lib.lkcuda_set_tip_states_async(eng, ctypes.cast(pinned.data_ptr(), u8p), nchunks) == 0
        a.clearAllPartialLH()
        l1, p1 = a.computeLikelihood(pattern_lh=True)
        assert l1 == l0
    # different data: every state rotated
    st2 = np.where(st < 4, (st + 1) % 4, st).astype(np.uint8)
    pinned2 = torch.from_numpy(np.ascontiguousarray(st2)).pin_memory()
    assert cuda_lib.lkcuda_set_tip_states_async(eng, ctypes.cast(pinned2.data_ptr(), u8p), nchunks) == 0
    a.clearAllPartialLH()
    l2 = a.computeLikelihood()
    assert l2 != l0
    assert cuda_lib.lkcuda_set_tip_states(eng, ctypes.cast(pinned2.data_ptr(), u8p)) == 0
    a.clearAllPartialLH()
    assert a.computeLikelihood() == l2
    # range check of the pipelined upload
    bad = st.copy(); bad[3, nptn - 5] = 19
    pinned3 = torch.from_numpy(np.ascontiguousarray(bad)).pin_memory()
    assert cuda_lib.lkcuda_set_tip_states_async(eng, ctypes.cast(pinned3.data_ptr(), u8p), nchunks) == 0
    assert cuda_lib.lkcuda_synchronize(eng) == 1      # LKCUDA_ERR_ARG
    assert b"out of range" in cuda_lib.lkcuda_last_error(eng)
    a.close(); b.close()


def _subset_tree(g, lib, npat, devices=None):
    """the fixture's tree and model on its first `npat` patterns only"""
    from iqtree_b200 import PhyloTree
    m = g.meta
    t = PhyloTree(lib=lib)
    st = g["tip_states"]
    t.setAlignment(st[:, :npat], g["ptn_freq"][:npat], names=[str(x) for x in g["taxa"]], seq_type=m["seq_type"],
                   nstates=m["nstates"], state_unknown=m["state_unknown"])
    t.setConstChar(g["const_char"][:npat])
    t.readAdjacency(str(g["tree_adj"]))
    t.setModel(gu.FAMILY[m["family"]], m["ncat"], m["nmix"], g["eval"], g["evec"], g["inv_evec"], g["rates"], g["props"],
               g["mixprops"], g["state_freq"], m["p_invar"])
    t.setParams(m["min_branch_length"], m["max_branch_length"])
    if devices is not None:
        t.setDevices(devices)
    return t


@pytest.mark.parametrize("npat,nshard", [(1, 1), (3, 5), (33, 2), (257, 4)])
def test_ragged_and_empty_shards(npat, nshard, cuda_lib, oracle_lib):
    """edge sizes: a single pattern, fewer patterns than shards (EMPTY shards still take part in the fused all-reduce), sizes that are
    not multiples of the kernels' tiles; checked against the oracle on the same truncated alignment"""
    g = gu.Golden("dna_example_gtr_i_g4")
    a = _subset_tree(g, cuda_lib, npat, devices=[0] * nshard)
    b = _subset_tree(g, oracle_lib, npat)
    a.initializeAllPartialLh(); b.initializeAllPartialLh()
    la, lb = a.computeLikelihood(), b.computeLikelihood()
    assert abs(la - lb) <= gu.LNL_RTOL * abs(lb)
    _, pa = a.computeLikelihood(pattern_lh=True)
    _, pb = b.computeLikelihood(pattern_lh=True)
    gu.rel_close(pa, pb, gu.PTN_RTOL, "pattern_lh")
    for (n1, n2) in a.getTraversal()[::7]:
        da = a.computeLikelihoodDerv(int(n1), int(n2), reset_theta=True)
        db = b.computeLikelihoodDerv(int(n1), int(n2), reset_theta=True)
        tol = 1e-9 * max(abs(db[0]), abs(db[1]), 1.0)
        assert abs(da[0] - db[0]) <= tol and abs(da[1] - db[1]) <= tol
        fa, fb = a.computeLikelihoodFromBuffer(), b.computeLikelihoodFromBuffer()
        assert abs(fa - fb) <= gu.LNL_RTOL * abs(fb)
    oa, ob = a.optimizeAllBranches(1), b.optimizeAllBranches(1)
    assert abs(oa - ob) <= gu.LNL_RTOL * abs(ob)
    a.close(); b.close()


@pytest.mark.parametrize("nshard", [1, 3])
def test_async_pattern_lh_fetch(nshard, cuda_lib):
    """lktree_compute_likelihood_async_fetch + lktree_wait_pattern_lh deliver what the synchronous computeLikelihood(pattern_lh) does,
    also when df/ddf work is queued behind the fetch and when fetches follow each other"""
    import ctypes as C
    g = gu.Golden("dna_deep_scaling")
    t = gu.build_tree(g, cuda_lib, devices=[0] * nshard)
    t.initializeAllPartialLh()
    lnl, plh = t.computeLikelihood(pattern_lh=True)
    a, b = g.meta["derv_branch"]
    for _ in range(3):
        out = np.zeros_like(plh)
        v = C.c_double()
        t.clearAllPartialLH()
        assert cuda_lib.lktree_compute_likelihood_async_fetch(t.h, C.byref(v), out.ctypes.data_as(C.POINTER(C.c_double))) == 0
        df, ddf = t.computeLikelihoodDerv(a, b, reset_theta=True)
        assert cuda_lib.lktree_wait_pattern_lh(t.h) == 0
        assert v.value == lnl and np.array_equal(out, plh)
        assert abs(df - g.meta["df_root"]) <= 1e-7 * max(1.0, abs(g.meta["df_root"]))
    gu.rel_close(plh, g["pattern_lh_scaled"], gu.PTN_RTOL, "pattern_lh")
    t.close()


@pytest.mark.gpu
def test_fused_root_branch_variant(cuda_lib):
    """lkcuda_partial_batch_lnl can evaluate the root branch in the tail of the traversal launch (traverse_s4_kernel<true>, opt-in with
    LKCUDA_FUSED_LNL=1 because it measured slower, profiles/r2_s16_fused_root_lnl.txt): the DNA fixtures, the pipelined upload and the
    sharded runs once more with the variant switched on, in a child process (the switch is read once per process)"""
    import subprocess
    if os.environ.get("LKCUDA_FUSED_LNL"):
        pytest.skip("already inside the variant run")
    env = dict(os.environ, LKCUDA_FUSED_LNL="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-x", "-q", "-k",
                        "test_cuda_matches_reference or test_pipelined_tip_state_upload or test_shards_one_device or test_ragged_and_empty_shards"],
                       env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-3000:]
    assert " passed" in r.stdout


@pytest.mark.gpu
def test_traversal_writes_every_pattern_of_every_vector(cuda_lib, monkeypatch):
    """Coverage of the whole-traversal kernel on NaN-poisoned engines (tests/conftest.py sets LKCUDA_POISON): after ONE full traversal of a
    fresh engine no arena vector may hold a NaN row - a pattern no warp wrote, or one written through a ring stage that had already been
    refilled (the hole DESIGN 3.1 describes) - and the bits of all vectors must not depend on the run or on how many shared-memory slots
    the schedule got (slot, register and arena-read paths compute the same bits)."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import trv_probe
    big = trv_probe.shape(100, 250_000)
    runs = [trv_probe.probe(cuda_lib, *big, 4) for _ in range(3)]
    for r in runs:
        assert r["nan_rows"] == 0 and r["vectors"] == 98, r
        assert r["digest"] == runs[0]["digest"] and r["lnl"] == runs[0]["lnl"], (r, runs[0])
    small = trv_probe.shape(100, 60_011)
    ref = trv_probe.probe(cuda_lib, *small, 4)
    assert ref["nan_rows"] == 0
    for maxslots in (0, 1, 2):
        monkeypatch.setenv("LKCUDA_TRV_MAXSLOTS", str(maxslots))
        r = trv_probe.probe(cuda_lib, *small, 4)
        assert r["nan_rows"] == 0 and r["digest"] == ref["digest"] and r["lnl"] == ref["lnl"], (maxslots, r, ref)
