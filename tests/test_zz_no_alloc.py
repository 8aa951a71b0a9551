"""lkcuda_derv_noalloc (include/lkcuda.h): the df / ddf reduction for callers that must not allocate - static device scratch, caller-owned
device inputs and output - against the existing entry point lkcuda_derv on an engine's theta_all.  Collected last on purpose.
The GPU test passes alone in a fresh process and after the other tests: it keeps no state except the
library's static scratch, which every call re-initialises."""
import ctypes as C

import numpy as np
import pytest

import golden_util as gu

MAX_PATTERNS = 1 << 21      # LKCUDA_NOALLOC_MAX_PATTERNS
ERR_ARG = 1


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _model(g):
    return (np.ascontiguousarray(g["eval"], dtype=np.float64), np.ascontiguousarray(g["rates"], dtype=np.float64),
            np.ascontiguousarray(g["props"], dtype=np.float64))


def test_noalloc_refuses_sizes_beyond_the_static_scratch():
    """CPU: the argument checks come before any device work, so the product library refuses here too"""
    import iqtree_b200
    lib = iqtree_b200.load_library()
    one = np.ones(4)
    args = lambda nptn, S, K: (0, None, C.c_void_p(8), C.c_void_p(8), C.c_void_p(8), nptn, S, K, _dp(one), _dp(one), _dp(one), 0.1, C.c_void_p(8))
    assert lib.lkcuda_derv_noalloc(*args(MAX_PATTERNS + 1, 4, 4)) == ERR_ARG
    assert b"LKCUDA_NOALLOC_MAX_PATTERNS" in lib.lkcuda_last_error(None)
    assert lib.lkcuda_derv_noalloc(*args(10, 20, 20)) == ERR_ARG          # 400 > LKCUDA_NOALLOC_MAX_BLOCK
    assert lib.lkcuda_derv_noalloc(*args(10, 65, 1)) == ERR_ARG
    assert lib.lkcuda_derv_noalloc(*args(-1, 4, 4)) == ERR_ARG


@pytest.mark.parametrize("case", ["dna_example_gtr_i_g4", "aa_lg_g4"])
def test_noalloc_oracle_statement_matches_engine_derv(case, oracle_lib):
    """CPU: the scalar statement of the new entry point (host arrays) against the oracle engine's lkcuda_derv"""
    g = gu.Golden(case)
    t = gu.build_tree(g, oracle_lib)
    t.initializeAllPartialLh()
    t.computeLikelihood()
    a, b = g.meta["derv_branch"]
    t.computeLikelihoodDerv(a, b, reset_theta=True)
    e = t.engine(0)
    theta = np.ascontiguousarray(t.fetchTheta())
    P = g.meta["nptn"]
    freq, invar = np.ascontiguousarray(g["ptn_freq"], dtype=np.float64), np.ascontiguousarray(t.ptnInvar()[:P])
    ev, rates, props = _model(g)
    for ln in (t.getBranchLength(a, b), 0.37):
        res = (C.c_double * 8)()
        assert oracle_lib.lkcuda_derv(e, ln, res) == 0
        out = np.zeros(2)
        rc = oracle_lib.lkcuda_derv_noalloc(0, None, theta.ctypes.data, freq.ctypes.data, invar.ctypes.data, P, g.meta["nstates"], g.meta["ncat"],
                                            _dp(ev), _dp(rates), _dp(props), ln, out.ctypes.data)
        assert rc == 0
        tol = 1e-12 * max(abs(res[1]), abs(res[2]), 1.0)
        assert abs(out[0] - res[1]) <= tol and abs(out[1] - res[2]) <= tol, (out, res[1], res[2])
    t.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["dna_example_gtr_i_g4", "aa_lg_g4"])
def test_noalloc_matches_lkcuda_derv_twice_in_a_row(case, cuda_lib):
    import torch
    lib = cuda_lib
    g = gu.Golden(case)
    assert g.meta["nunobs"] == 0 and g.meta["nmix"] == 1
    t = gu.build_tree(g, lib)
    t.initializeAllPartialLh()
    t.computeLikelihood()
    a, b = g.meta["derv_branch"]
    t.computeLikelihoodDerv(a, b, reset_theta=True)      # theta_all of the branch is in the engine now
    e = t.engine(0)
    P = g.meta["nptn"]
    dev = torch.device("cuda:0")
    theta = torch.from_numpy(np.ascontiguousarray(t.fetchTheta())).to(dev)
    assert theta.shape == (P, g.meta["nstates"] * g.meta["ncat"])
    freq = torch.from_numpy(np.ascontiguousarray(g["ptn_freq"], dtype=np.float64)).to(dev)
    invar = torch.from_numpy(np.ascontiguousarray(t.ptnInvar()[:P])).to(dev)
    ev, rates, props = _model(g)
    out = torch.full((2,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    rtol = gu.partial_rtol(g)
    got = []
    # two calls in a row on the shared static scratch (ordered on one stream), at two lengths, then the first length again
    for ln in (t.getBranchLength(a, b), 0.37, t.getBranchLength(a, b)):
        res = (C.c_double * 8)()
        assert lib.lkcuda_derv(e, ln, res) == 0, lib.lkcuda_last_error(e)
        rc = lib.lkcuda_derv_noalloc(0, None, theta.data_ptr(), freq.data_ptr(), invar.data_ptr(), P, g.meta["nstates"], g.meta["ncat"],
                                     _dp(ev), _dp(rates), _dp(props), ln, out.data_ptr())
        assert rc == 0, lib.lkcuda_last_error(None)
        torch.cuda.synchronize()
        o = out.cpu().numpy().copy()
        tol = rtol * max(abs(res[1]), abs(res[2]), 1.0)
        assert abs(o[0] - res[1]) <= tol and abs(o[1] - res[2]) <= tol, (ln, o, res[1], res[2])
        got.append(o)
        out.fill_(float("nan"))
        torch.cuda.synchronize()
    assert np.array_equal(got[0], got[2]), "same inputs must give the same bits on every call"
    # against the golden df / ddf of the reference at the branch's own length
    m = g.meta
    tol = rtol * max(abs(m["ddf_root"]), abs(m["df_root"]), 1.0)
    if list(m["derv_branch"]) == [a, b] and "df_root" in m:
        assert abs(got[0][0] - m["df_root"]) <= tol and abs(got[0][1] - m["ddf_root"]) <= tol
    # a size beyond the static scratch is refused with a clear error, nothing is launched
    assert lib.lkcuda_derv_noalloc(0, None, theta.data_ptr(), freq.data_ptr(), invar.data_ptr(), MAX_PATTERNS + 1, g.meta["nstates"], g.meta["ncat"],
                                   _dp(ev), _dp(rates), _dp(props), 0.1, out.data_ptr()) == ERR_ARG
    assert b"LKCUDA_NOALLOC_MAX_PATTERNS" in lib.lkcuda_last_error(None)
    t.close()


@pytest.mark.gpu
def test_noalloc_many_blocks(cuda_lib):
    """more patterns than one block handles (several hundred blocks, a ragged last one) against a float64 numpy statement"""
    import torch
    rng = np.random.default_rng(5)
    P, S, K = 200_003, 4, 4
    theta = rng.random((P, S * K)) * 1e-3
    freq = rng.integers(1, 5, P).astype(np.float64)
    invar = rng.random(P) * 1e-4
    ev = -np.array([0.0, 0.9, 1.3, 2.1]); rates = np.array([0.1, 0.5, 1.0, 2.4]); props = np.full(4, 0.25)
    ln = 0.21
    cof = (rates[:, None] * ev[None, :]).reshape(-1)
    v0 = np.exp(cof * ln) * np.repeat(props, S); v1 = cof * v0; v2 = cof * v1
    lh = np.abs(theta @ v0 + invar); d1 = (theta @ v1) / lh; d2 = (theta @ v2) / lh
    df, ddf = float(np.sum(freq * d1)), float(np.sum(freq * d2 - freq * d1 * d1))
    dev = torch.device("cuda:0")
    th, fr, iv = (torch.from_numpy(x).to(dev) for x in (theta, freq, invar))
    out = torch.zeros(2, dtype=torch.float64, device=dev)
    for _ in range(2):
        out.fill_(float("nan"))
        rc = cuda_lib.lkcuda_derv_noalloc(0, None, th.data_ptr(), fr.data_ptr(), iv.data_ptr(), P, S, K, _dp(ev), _dp(rates), _dp(props), ln, out.data_ptr())
        assert rc == 0, cuda_lib.lkcuda_last_error(None)
        torch.cuda.synchronize()
        o = out.cpu().numpy()
        assert abs(o[0] - df) <= 1e-11 * abs(df) and abs(o[1] - ddf) <= 1e-11 * abs(ddf), (o, df, ddf)


def test_noalloc_oracle_edge_cases(oracle_lib):
    """CPU: empty input gives zeros; a pattern whose likelihood is 0 makes df NaN/Inf, which the entry point reports as df = ddf = 0 like
    the reference's derivative kernels (phylokernel.h:647-651)"""
    ev = np.array([0.0, -1.0, -2.0, -3.0]); rates = np.array([0.5, 1.5]); props = np.array([0.5, 0.5])
    out = np.full(2, 7.0)
    one = np.ones(8)
    rc = oracle_lib.lkcuda_derv_noalloc(0, None, one.ctypes.data, one.ctypes.data, one.ctypes.data, 0, 4, 2, _dp(ev), _dp(rates), _dp(props), 0.1, out.ctypes.data)
    assert rc == 0 and out[0] == 0.0 and out[1] == 0.0
    theta = np.zeros((3, 8)); theta[0] = 1e-3; theta[2] = 2e-3          # pattern 1: all zero -> lh = 0 -> 0 / 0
    freq = np.ones(3); invar = np.zeros(3)
    rc = oracle_lib.lkcuda_derv_noalloc(0, None, theta.ctypes.data, freq.ctypes.data, invar.ctypes.data, 3, 4, 2, _dp(ev), _dp(rates), _dp(props), 0.1, out.ctypes.data)
    assert rc == 0 and out[0] == 0.0 and out[1] == 0.0
    theta[1] = 5e-4
    rc = oracle_lib.lkcuda_derv_noalloc(0, None, theta.ctypes.data, freq.ctypes.data, invar.ctypes.data, 3, 4, 2, _dp(ev), _dp(rates), _dp(props), 0.1, out.ctypes.data)
    assert rc == 0 and np.isfinite(out).all() and out[0] < 0.0


@pytest.mark.gpu
def test_noalloc_edge_cases_on_the_device(cuda_lib):
    """the same edge cases through the CUDA entry point: no pattern -> zeros (no launch), a 0 / 0 pattern -> df = ddf = 0"""
    import torch
    dev = torch.device("cuda:0")
    ev = np.array([0.0, -1.0, -2.0, -3.0]); rates = np.array([0.5, 1.5]); props = np.array([0.5, 0.5])
    out = torch.full((2,), 7.0, dtype=torch.float64, device=dev)
    theta = torch.zeros((3, 8), dtype=torch.float64, device=dev); theta[0] = 1e-3; theta[2] = 2e-3
    freq = torch.ones(3, dtype=torch.float64, device=dev); invar = torch.zeros(3, dtype=torch.float64, device=dev)
    call = lambda n: cuda_lib.lkcuda_derv_noalloc(0, None, theta.data_ptr(), freq.data_ptr(), invar.data_ptr(), n, 4, 2, _dp(ev), _dp(rates), _dp(props), 0.1, out.data_ptr())
    assert call(0) == 0
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [0.0, 0.0]
    out.fill_(7.0)
    assert call(3) == 0
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [0.0, 0.0]
    theta[1] = 5e-4
    assert call(3) == 0
    torch.cuda.synchronize()
    o = out.cpu().numpy()
    assert np.isfinite(o).all() and o[0] < 0.0
