"""the consumers of _pattern_lh_cat and _pattern_lh (UFBoot RELL dot products, RateGamma::computePatternRates,
ModelMixture::optimizeWeights) against golden values the UNMODIFIED reference produced (tests/golden/aux/*.npz, written by
`python oracle/make_golden.py --aux` through oracle/_ref/ref_harness dumpx).

Tolerances:
  * RELL on the oracle (a lane-by-lane restatement of dotProductSIMD<float, Vec8f, 8>): bit-exact with the reference's float result.
  * RELL on the device: fp32 products, fp64 accumulation - MORE accurate than the reference's float chains; compared within the
    reference's own float accumulation error (2e-5 relative) and within 1e-6 of an fp64 dot product of the float inputs.
  * pattern rates: 1e-10 relative (per-pattern likelihood tolerance), categories exact.
  * EM weights / lnL: 1e-9 relative.
"""
import ctypes as C
import glob
import json
import os

import numpy as np
import pytest

import golden_util as gu

AUX_DIR = os.path.join(gu.GOLDEN_DIR, "aux")
AUX_CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(AUX_DIR, "*.npz")))


def load_aux(case):
    z = np.load(os.path.join(AUX_DIR, case + ".npz"))
    return z, json.loads(str(z["meta"]))


def check_rell(t, g, z, m, exact):
    lnl = gu.check_likelihood(t, g)
    assert list(t.rootBranch()) == list(m["root_branch"])
    t.setBootSamples(z["boot_samples"])
    rell = t.computeRell()
    ref = z["rell"]
    if exact:
        assert np.array_equal(rell.astype(np.float32), ref.astype(np.float32)), "RELL differs from dotProductSIMD<float,Vec8f,8>: %r vs %r" % (rell[:3], ref[:3])
    else:
        gu.rel_close(rell, ref, 2e-5, "RELL vs reference (float accumulation)")
        x = z["ptn_lh_full"].astype(np.float32).astype(np.float64)
        exact64 = z["boot_samples"].astype(np.float64) @ x
        gu.rel_close(rell, exact64, 1e-6, "RELL vs fp64 dot product of the float inputs")
    # the sum of the sample frequencies is the site count: a sample of all-original frequencies reproduces lnL (float precision)
    t.setBootSamples(g["ptn_freq"][None, :].astype(np.float32))
    r1 = t.computeRell()
    assert abs(r1[0] - lnl) <= 2e-5 * abs(lnl)


def check_rates(t, z, m):
    if not m["nrate_cat"] or "pattern_rates" not in z.files:
        return
    r, c = t.computePatternRates()
    gu.rel_close(r, z["pattern_rates"], 1e-10, "pattern rates")
    assert np.array_equal(c, z["pattern_cat"]), "pattern categories differ"


def check_em(t, g, z, m):
    if not m["has_em"]:
        return
    t.clearAllPartialLH()
    t.computeLikelihood()
    lnl, w = t.optimizeMixtureWeights(m["nmix"])
    gu.rel_close(w, z["em_props"], 1e-9, "mixture weights after ModelMixture::optimizeWeights")
    assert abs(lnl - m["em_lnl"]) <= gu.LNL_RTOL * abs(m["em_lnl"]), (lnl, m["em_lnl"])
    assert abs(w.sum() - 1.0) < 1e-9


@pytest.mark.parametrize("case", AUX_CASES)
def test_oracle_aux_matches_reference(case, oracle_lib):
    g = gu.Golden(case)
    z, m = load_aux(case)
    t = gu.build_tree(g, oracle_lib)
    t.initializeAllPartialLh()
    check_rell(t, g, z, m, exact=m["dot_kernel"] == "AVX")
    check_rates(t, z, m)
    check_em(t, g, z, m)
    t.close()


@pytest.mark.gpu
@pytest.mark.parametrize("nshard", [1, 3])
@pytest.mark.parametrize("case", AUX_CASES)
def test_cuda_aux_matches_reference(case, nshard, cuda_lib):
    g = gu.Golden(case)
    z, m = load_aux(case)
    t = gu.build_tree(g, cuda_lib, devices=[0] * nshard)
    t.initializeAllPartialLh()
    check_rell(t, g, z, m, exact=False)
    check_rates(t, z, m)
    check_em(t, g, z, m)
    t.close()


@pytest.mark.gpu
def test_cuda_rell_storage_kinds(cuda_lib, oracle_lib):
    """the RELL matrix is stored losslessly as uint8, uint16 or float depending on its entries: all three give the oracle's sums"""
    import ctypes
    g = gu.Golden("dna_example_gtr_g4")
    P = g.meta["nptn"]
    rng = np.random.default_rng(5)
    base = rng.integers(0, 4, size=(19, P)).astype(np.float32)
    mats = {1: base, 2: base.copy(), 4: base.copy()}
    mats[2][3, 7] = 300.0
    mats[4][5, 11] = 0.5
    for nbytes, mat in mats.items():
        res = []
        for lib in (cuda_lib, oracle_lib):
            t = gu.build_tree(g, lib)
            t.initializeAllPartialLh()
            t.computeLikelihood()
            t.setBootSamples(mat)
            if lib is cuda_lib:
                k = ctypes.c_int32()
                assert lib.lkcuda_boot_storage(t.engine(0), ctypes.byref(k), None) == 0 and k.value == nbytes
            res.append(t.computeRell())
            t.close()
        gu.rel_close(res[0], res[1], 2e-5, "RELL %d-byte storage vs oracle" % nbytes)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(2000, 21, 12), (5000, 60, 40)])
def test_cuda_em_matches_oracle_synthetic(shape, cuda_lib, oracle_lib):
    """lkcuda_lh_cat_collapse / lkcuda_lh_cat_em on random _pattern_lh_cat-like input through a real engine: mixture C-classes x G-cats"""
    # a mixture fixture gives real engines of the right block structure; the EM itself is checked above against the reference, here
    # CUDA and oracle run ModelMixture::optimizeWeights from perturbed weights
    g = gu.Golden("aa_lg_c20_g4")
    z, m = load_aux("aa_lg_c20_g4")
    out = []
    for lib in (cuda_lib, oracle_lib):
        t = gu.build_tree(g, lib)
        t.initializeAllPartialLh()
        t.computeLikelihood()
        a = t.optimizeMixtureWeights(m["nmix"])
        b = t.optimizeMixtureWeights(m["nmix"])     # a second round from the updated weights
        out.append((a, b))
        t.close()
    for k in range(2):
        assert abs(out[0][k][0] - out[1][k][0]) <= gu.LNL_RTOL * abs(out[1][k][0])
        gu.rel_close(out[0][k][1], out[1][k][1], 1e-9, "weights round %d" % k)


@pytest.mark.gpu
@pytest.mark.parametrize("case,wsl,include_invar", [("dna_example_gtr_i_g4", 2, 1), ("dna_example_gtr_i_g4", 2, 0), ("aa_lg_c20_g4", 3, 1)])
def test_cuda_posterior_weights_match_oracle(case, wsl, include_invar, cuda_lib, oracle_lib):
    """lkcuda_lh_cat_posterior_weights (the sub-tree weights of RateFree / ModelMixture::optimizeWithEM): the posterior column of one class
    becomes the ptn_freq of a second engine; its likelihood must agree between the CUDA engine and the CPU restatement"""
    g = gu.Golden(case)
    out = []
    for lib in (cuda_lib, oracle_lib):
        a, b = gu.build_tree(g, lib), gu.build_tree(g, lib)
        for t in (a, b):
            t.initializeAllPartialLh()
            t.computeLikelihood()
        a.computePatternLhCat(wsl)                  # fills _pattern_lh_cat on the device (the generic branch pass)
        n = C.c_int32()
        mixture = g.meta["family"] == "mixture"          # the arguments lk::PhyloTree::collapsePatternLhCat passes
        mode, outer = (1, g.meta["nmix"]) if mixture else (0, 1)
        assert lib.lkcuda_lh_cat_collapse(a.engine(0), mode, g.meta["ncat"], outer, C.byref(n)) == 0, lib.lkcuda_last_error(a.engine(0))
        vals = []
        for col in (0, n.value - 1):
            assert lib.lkcuda_lh_cat_posterior_weights(a.engine(0), col, include_invar, b.engine(0)) == 0, lib.lkcuda_last_error(a.engine(0))
            b.clearAllPartialLH()
            vals.append(b.computeLikelihood())
        assert lib.lkcuda_lh_cat_posterior_weights(a.engine(0), n.value, include_invar, b.engine(0)) != 0      # column out of range
        assert lib.lkcuda_lh_cat_posterior_weights(a.engine(0), 0, include_invar, a.engine(0)) != 0            # onto itself
        out.append(vals)
        a.close(); b.close()
    gu.rel_close(np.array(out[0]), np.array(out[1]), 1e-9, "lnL under posterior weights")
    assert out[0][0] != out[0][1]
