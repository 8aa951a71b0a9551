"""CPU suite: the ring protocol of the whole-traversal kernels at SASS level.  A consumer warp must not hand a ring stage back
(mbarrier.arrive) while one of its shared-memory loads from the ring is still in flight: on sm_100a the arrive does not wait for it and the
producer's next bulk copy can replace the stage under the load (found in round 2 with LKCUDA_POISON).  tools/sass_release_check.py
walks the built library's SASS and reports every consumer arrive with an un-waited LDS scoreboard."""
import os
import shutil
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_no_shared_load_in_flight_when_a_stage_is_released():
    if shutil.which("cuobjdump") is None and not os.path.exists("/usr/local/cuda/bin/cuobjdump"):
        pytest.skip("cuobjdump not available")
    import sass_release_check as chk
    from iqtree_b200 import _build
    lib = _build.build_cuda()
    seen = 0
    for name, body in chk.functions(lib):
        if "traverse_" not in name:
            continue
        insns = chk.decode(body)
        arrives = sum(1 for _, t, _, _ in insns if "SYNCS.ARRIVE" in t and "A1T0" in t)
        bad = chk.check(insns)
        assert not bad, "%s: consumer arrive with an LDS in flight: %r" % (name, bad[:3])
        seen += arrives
    assert seen >= 7, "the ring kernels were not found in the library's SASS (%d consumer arrives)" % seen
