#!/usr/bin/env python
"""Multi-process check of the fused exchange (run under torchrun, one rank per GPU):
every rank holds a pattern shard of a golden fixture, the reduction kernels exchange their sums through the peer mailboxes
(lkcuda_attach_peers, CUDA IPC) and - with device_newton - the persistent Newton kernel exchanges once per evaluation INSIDE its loop.
Checks lnL, df/ddf, optimizeAllBranches and getBestNNIForBran against the reference's golden values on every rank.
usage: torchrun --nproc-per-node N tools/peer_newton_check.py [case]"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import iqtree_b200
import golden_util as gu


def main():
    case = sys.argv[1] if len(sys.argv) > 1 else "dna_example_gtr_g4"
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lib = iqtree_b200.load_library()
    g = gu.Golden(case)
    out = {}
    for device_newton in (False, True):
        t = gu.build_tree(g, lib, devices=[local], shard=(rank, world), device_newton=device_newton)
        t.initializeAllPartialLh()
        mine = torch.from_numpy(t.peerHandle()).to(dev)
        allh = [torch.zeros(64, dtype=torch.uint8, device=dev) for _ in range(world)]
        dist.all_gather(allh, mine)
        t.attachPeers(torch.stack(allh).cpu().numpy())
        lnl = t.computeLikelihood()
        assert abs(lnl - g.meta["lnl"]) <= gu.LNL_RTOL * abs(g.meta["lnl"]), (lnl, g.meta["lnl"])
        gu.check_optimize(t, g, full=True)
        nni = os.path.join(gu.GOLDEN_DIR, "nni", case + ".nni1.npz")
        nchecked = 0
        if os.path.exists(nni):
            z = np.load(nni)
            t2 = gu.build_tree(g, lib, devices=[local], shard=(rank, world), device_newton=device_newton)
            t.close()
            t = t2
            t.initializeAllPartialLh()
            mine = torch.from_numpy(t.peerHandle()).to(dev)
            dist.all_gather(allh, mine)
            t.attachPeers(torch.stack(allh).cpu().numpy())
            t.computeLikelihood()
            for j in range(0, len(z["ids"]), 5):
                mv = t.getBestNNIForBran(int(z["ids"][j, 0]), int(z["ids"][j, 1]), nni5=True)
                for c in range(2):
                    assert abs(mv[c][0] - z["vals"][j, c, 0]) <= gu.LNL_RTOL * abs(z["vals"][j, c, 0])
                    assert np.all(np.abs(mv[c][1] - z["vals"][j, c, 1:]) <= gu.BRLEN_ATOL)
                nchecked += 1
        out["device_newton=%s" % device_newton] = {"lnl": lnl, "nni_branches_checked": nchecked}
        t.close()
        dist.barrier()
    if rank == 0:
        print(json.dumps({"case": case, "world": world, "ok": True, **out}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
