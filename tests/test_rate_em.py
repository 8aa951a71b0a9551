"""RateFree::optimizeWithEM (model/ratefree.cpp:450-578) against the UNMODIFIED reference's own result
(tests/golden/em/*.npz, `python oracle/make_golden.py --em` -> ref_harness dumpem).

The EM is a long chain of decisions (Brent steps on full-tree likelihoods of one-category tree copies whose pattern weights are
posteriors): every host statement is mirrored, so the final rates / proportions agree far below Brent's own 1e-3 tolerance.
Tolerances: lnL 1e-9 relative; rates and proportions 1e-7 relative (likelihood differences of 1e-12 move a parabolic step that much).
"""
import glob
import json
import os

import numpy as np
import pytest

import golden_util as gu

EM_DIR = os.path.join(gu.GOLDEN_DIR, "em")
EM_CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(EM_DIR, "*.npz")))


def run_case(case, lib, devices=None):
    g = gu.Golden(case)
    z = np.load(os.path.join(EM_DIR, case + ".npz"))
    m = json.loads(str(z["meta"]))
    t = gu.build_tree(g, lib, devices=devices) if devices else gu.build_tree(g, lib)
    t.initializeAllPartialLh()
    lnl0 = t.computeLikelihood()
    assert abs(lnl0 - m["lnl_before"]) <= gu.LNL_RTOL * abs(lnl0)
    lnl, rates, props = t.optimizeRateFreeEM(m["ncat"], copy_precision=m["num_precision"])
    assert abs(lnl - m["lnl_after"]) <= 1e-9 * abs(lnl), (lnl, m["lnl_after"])
    gu.rel_close(rates, z["rates"], 1e-7, "+R rates after RateFree::optimizeWithEM")
    gu.rel_close(props, z["props"], 1e-7, "+R proportions after RateFree::optimizeWithEM")
    assert lnl > lnl0
    # the tree itself is untouched (the EM works on copies): same likelihood at the new parameters after a full recomputation
    t.clearAllPartialLH()
    assert abs(t.computeLikelihood() - lnl) <= 1e-12 * abs(lnl)
    t.close()


@pytest.mark.parametrize("case", EM_CASES)
def test_oracle_rate_free_em_matches_reference(case, oracle_lib):
    run_case(case, oracle_lib)


@pytest.mark.gpu
@pytest.mark.parametrize("nshard", [1, 2])
@pytest.mark.parametrize("case", EM_CASES)
def test_cuda_rate_free_em_matches_reference(case, nshard, cuda_lib):
    run_case(case, cuda_lib, devices=[0] * nshard)
