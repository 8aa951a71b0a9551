"""Caller-owned arenas (second allocator): two partition trees carve their vectors out of ONE shared arena with their
own strides, as PhyloSuperTreePlen::initializeAllPartialLh does (phylosupertreeplen.cpp:1742-1871); the engines borrow the memory
and never free it (ModelMixture::optimizeWithEM aliasing, model/modelmixture.cpp:1427-1429).  Runs against the oracle library on
the CPU (same ABI, host logic) and against the CUDA engine on the GPU."""
import ctypes as C

import numpy as np
import pytest

import golden_util as gu


def _shared_arena_roundtrip(lib):
    cases = ["dna_example_gtr_g4", "aa_lg_g4"]
    trees, layouts = [], []
    for c in cases:
        g = gu.Golden(c)
        t = gu.build_tree(g, lib)
        t.initializeAllPartialLh()
        vs, ss = C.c_int64(), C.c_int64()
        assert lib.lkcuda_arena_layout(t.engine(0), C.byref(vs), C.byref(ss)) == 0
        P = g.meta["nptn"] + g.meta["nunobs"]
        assert vs.value >= P * g.meta["block"] and ss.value >= P
        nvec = 2 * (len(g["taxa"]) - 1) * 2          # generous: every direction of every branch
        trees.append((g, t)); layouts.append((vs.value, ss.value, nvec))
    # one arena for both partitions: [part 0 vectors | part 1 vectors], strides padded differently from the engines' own
    strides = [(vs + 36, ss + 6) for vs, ss, _ in layouts]
    lh_bytes = sum(n * s[0] * 8 for (_, _, n), s in zip(layouts, strides))
    sc_bytes = sum(n * s[1] * 2 for (_, _, n), s in zip(layouts, strides))
    lh_base, sc_base = C.c_void_p(), C.c_void_p()
    assert lib.lkcuda_device_malloc(0, lh_bytes, C.byref(lh_base)) == 0
    assert lib.lkcuda_device_malloc(0, sc_bytes, C.byref(sc_base)) == 0
    lh_off = sc_off = 0
    for (g, t), (_, _, nvec), (vst, sst) in zip(trees, layouts, strides):
        assert lh_off % 32 == 0
        rc = lib.lkcuda_attach_arena(t.engine(0), lh_base.value + lh_off, vst, sc_base.value + sc_off, sst, nvec)
        assert rc == 0, lib.lkcuda_last_error(t.engine(0))
        lh_off += nvec * vst * 8; sc_off += nvec * sst * 2
        sc_off += sc_off % 2
    for g, t in trees:
        t.clearAllPartialLH()
        gu.check_likelihood(t, g)
        gu.check_partials(t, g)
        gu.check_derivatives(t, g)
        gu.check_optimize(t, g, full=False)
    # misuse is refused: misaligned base, short stride
    g, t = trees[0]
    t.clearAllPartialLH()
    before = t.computeLikelihood()
    assert lib.lkcuda_attach_arena(t.engine(0), lh_base.value + 8, strides[0][0], sc_base.value, strides[0][1], 4) == 1
    assert lib.lkcuda_attach_arena(t.engine(0), lh_base.value, 8, sc_base.value, strides[0][1], 4) == 1
    # detach: back on the engine's own arena, the borrowed memory survives the engines
    assert lib.lkcuda_attach_arena(t.engine(0), None, 0, None, 0, 0) == 0
    t.clearAllPartialLH()
    assert t.computeLikelihood() == before
    for _, t in trees:
        t.close()
    assert lib.lkcuda_device_free(0, lh_base) == 0
    assert lib.lkcuda_device_free(0, sc_base) == 0


def test_shared_arena_oracle(oracle_lib):
    _shared_arena_roundtrip(oracle_lib)


@pytest.mark.gpu
def test_shared_arena_cuda(cuda_lib):
    _shared_arena_roundtrip(cuda_lib)


@pytest.mark.gpu
def test_page_locked_host_arrays(cuda_lib):
    """lkcuda_host_register / lkcuda_host_unregister (the binding page-locks ptn_freq / ptn_invar / _pattern_lh): same results from a
    registered array, idempotent, and unregistering an unknown pointer is not an error"""
    import ctypes as C
    g = gu.Golden("dna_example_gtr_i_g4")
    t = gu.build_tree(g, cuda_lib)
    t.initializeAllPartialLh()
    lnl = t.computeLikelihood()
    e = t.engine(0)
    P = g.meta["nptn"]
    freq = np.ascontiguousarray(g["ptn_freq"], dtype=np.float64).copy()
    invar = np.ascontiguousarray(t.ptnInvar()[:P]).copy()
    plh = np.zeros(P)
    dp = C.POINTER(C.c_double)
    for a in (freq, invar, plh):
        assert cuda_lib.lkcuda_host_register(e, a.ctypes.data, a.nbytes) == 0, cuda_lib.lkcuda_last_error(e)
    assert cuda_lib.lkcuda_host_register(e, freq.ctypes.data, freq.nbytes) == 0          # already registered: fine
    assert cuda_lib.lkcuda_set_pattern_weights(e, freq.ctypes.data_as(dp), invar.ctypes.data_as(dp)) == 0
    t.clearAllPartialLH()
    assert t.computeLikelihood() == lnl
    assert cuda_lib.lkcuda_fetch_pattern_lh(e, plh.ctypes.data_as(dp)) == 0
    assert np.all(np.isfinite(plh)) and np.all(plh < 0)
    # half the weight on every pattern: half the log-likelihood (no +ASC in this fixture)
    half = freq * 0.5
    assert cuda_lib.lkcuda_set_pattern_weights(e, half.ctypes.data_as(dp), invar.ctypes.data_as(dp)) == 0
    t.clearAllPartialLH()
    assert abs(t.computeLikelihood() - 0.5 * lnl) <= 1e-12 * abs(lnl)
    for a in (freq, invar, plh):
        assert cuda_lib.lkcuda_host_unregister(e, a.ctypes.data) == 0
    assert cuda_lib.lkcuda_host_unregister(e, half.ctypes.data) == 0                     # never registered: fine
    t.close()
