#!/usr/bin/env python
"""Lint of the ring protocol of the whole-traversal kernels at SASS level ("stage release"): a consumer warp hands a ring
stage back with mbarrier.arrive (SYNCS.ARRIVE...A1T0) after reading it with ordinary shared-memory loads, and on sm_100a the arrive does
NOT wait for an LDS that is still in flight - ptxas emits no scoreboard wait on its own, and the producer's next bulk copy can replace the
stage under the load (measured, round 2: whole warps storing a node's result through the NEXT node's descriptor).  The kernels therefore
make the arrive depend on the registers loaded last.  This tool walks the SASS of a library in address order, tracks which scoreboards
carry an LDS that no later instruction has waited for, and lists every consumer arrive issued while one is pending.
usage: python tools/sass_release_check.py [liblktree_cuda.so] [kernel-name-substring ...]   (exit status 1 if anything is pending)"""
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INSN = re.compile(r"/\*([0-9a-f]{4,6})\*/\s+(.*?);\s*/\* (0x[0-9a-f]{16}) \*/")
WORD = re.compile(r"/\* (0x[0-9a-f]{16}) \*/")


def functions(path):
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    text = subprocess.run([exe, "-sass", path], capture_output=True, text=True, check=True).stdout
    name, body = None, []
    for line in text.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            if name:
                yield name, body
            name, body = m.group(1), []
        elif name:
            body.append(line)
    if name:
        yield name, body


def decode(lines):
    """-> [(address, text, write_barrier or None, wait_mask)]"""
    out = []
    for i, line in enumerate(lines[:-1]):
        m = INSN.search(line)
        m2 = WORD.search(lines[i + 1]) if m else None
        if not m or not m2 or INSN.search(lines[i + 1]):
            continue
        ctrl = (int(m2.group(1), 16) >> 41) & 0x7fffff
        wb = (ctrl >> 5) & 7
        out.append((m.group(1), m.group(2), None if wb == 7 else wb, (ctrl >> 11) & 0x3f))
    return out


def check(insns):
    """consumer arrives issued while an LDS result is outstanding: [(address, text, [pending LDS addresses])]"""
    pending, bad = {}, []
    for addr, text, wb, wait in insns:
        for b in range(6):
            if wait >> b & 1:
                pending.pop(b, None)
        if "SYNCS.ARRIVE" in text and "A1T0" in text and pending:
            bad.append((addr, text, sorted(pending.values())))
        if wb is not None:
            if re.search(r"\bLDS", text):
                pending[wb] = addr
            else:
                pending.pop(wb, None)       # the scoreboard now tracks another kind of instruction
    return bad


def main():
    args = sys.argv[1:]
    lib = args[0] if args and os.path.exists(args[0]) else os.path.join(ROOT, "iq-tree_b200", "lib", "liblktree_cuda.so")
    pats = [a for a in args if a != lib] or ["traverse_"]
    status = 0
    for name, body in functions(lib):
        if not any(p in name for p in pats):
            continue
        insns = decode(body)
        arrives = sum(1 for _, t, _, _ in insns if "SYNCS.ARRIVE" in t and "A1T0" in t)
        bad = check(insns)
        print("%s: %d instructions, %d consumer arrives, %d with an LDS in flight" % (name, len(insns), arrives, len(bad)))
        for addr, text, pend in bad:
            print("    %s %s   <- LDS at %s not waited for" % (addr, text, ", ".join(pend)))
            status = 1
    return status


if __name__ == "__main__":
    sys.exit(main())
