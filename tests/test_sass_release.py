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


def _insn(addr, text, wb=7, wait=0):
    """two cuobjdump lines of one instruction with the given write-scoreboard index (7 = none) and wait mask in its control word"""
    ctrl = (1 << 0) | (wb << 5) | (7 << 8) | (wait << 11)
    return ["        /*%04x*/                   %s ;   /* 0x0000000000000000 */" % (addr, text),
            "                                                                  /* 0x%016x */" % (ctrl << 41)]


def test_lint_flags_an_unwaited_shared_load():
    """the lint itself: an arrive right behind an LDS nobody waited for is reported, a wait on its scoreboard (by any instruction in
    between, or by the arrive) clears it, and a scoreboard re-used by a non-LDS instruction no longer counts"""
    import sass_release_check as chk
    lds = _insn(0x10, "LDS.64 R84, [R121+UR4]", wb=1)
    arrive = _insn(0x40, "@!P5 SYNCS.ARRIVE.TRANS64.A1T0 RZ, [R122+URZ+0x40], RZ")
    bad = chk.check(chk.decode(lds + _insn(0x20, "IMAD.MOV.U32 R1, RZ, RZ, R2") + arrive + [""]))
    assert len(bad) == 1 and bad[0][2] == ["0010"]
    assert not chk.check(chk.decode(lds + _insn(0x20, "LOP3.LUT R21, R87, R85, RZ, 0xfc, !PT", wait=0b000010) + arrive + [""]))
    assert not chk.check(chk.decode(lds + _insn(0x40, "@!P5 SYNCS.ARRIVE.TRANS64.A1T0 RZ, [R122+URZ+0x40], RZ", wait=0b000010) + [""]))
    assert not chk.check(chk.decode(lds + _insn(0x20, "LDG.E.64 R2, desc[UR18][R4.64]", wb=1) + arrive + [""]))
    # the producer's arrive.expect_tx (no A1T0) is not a stage release
    assert not chk.check(chk.decode(lds + _insn(0x40, "SYNCS.ARRIVE.TRANS64 RZ, [R33+URZ], R25") + [""]))
