"""GPU suite, needs >= 2 devices (skipped on a one-GPU box): one PROCESS per GPU, pattern shards, the scalar sums exchanged inside the
reduction kernels through the peer mailboxes (lkcuda_peer_handle / lkcuda_attach_peers over CUDA IPC) - the path bench.py
takes under torchrun.  The 64-byte handles travel through gloo; nothing else is exchanged on the host.  Also covers the persistent
device Newton kernel exchanging once per evaluation inside its loop."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, case, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import golden_util as gu
    import iqtree_b200

    dist.init_process_group("gloo", rank=rank, world_size=world)
    lib = iqtree_b200.load_library()
    g = gu.Golden(case)
    res = {}
    for device_newton in (False, True):
        t = gu.build_tree(g, lib, devices=[rank], shard=(rank, world), device_newton=device_newton)
        t.initializeAllPartialLh()
        mine = torch.from_numpy(t.peerHandle().copy())
        allh = [torch.zeros(64, dtype=torch.uint8) for _ in range(world)]
        dist.all_gather(allh, mine)
        t.attachPeers(torch.stack(allh).numpy())
        assert t.peersAttached()
        lnl = t.computeLikelihood()
        a, b = g.meta["derv_branch"]
        df, ddf = t.computeLikelihoodDerv(a, b, reset_theta=True)
        fb = t.computeLikelihoodFromBuffer()
        t.clearAllPartialLH()
        t.computeLikelihood()
        oab = t.optimizeAllBranches(1)
        lens = [t.getBranchLength(int(x), int(y)) for (x, y) in t.getTraversal()]
        res["dn%d" % int(device_newton)] = np.array([lnl, df, ddf, fb, oab] + lens)
        t.detachPeers()
        t.close()
        dist.barrier()
    np.savez(os.path.join(out_dir, "peer_rank%d.npz" % rank), **res)
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["dna_example_gtr_i_g4", "aa_lg_g4"])
def test_two_processes_two_gpus_peer_exchange(case, tmp_path, cuda_lib):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    import golden_util as gu
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), case, str(tmp_path)), nprocs=world, join=True)
    g = gu.Golden(case)
    m = g.meta
    r = [np.load(os.path.join(str(tmp_path), "peer_rank%d.npz" % i)) for i in range(world)]
    ref = gu.parse_adjacency(g["tree_adj_oab1"])
    for key in ("dn0", "dn1"):
        assert np.array_equal(r[0][key], r[1][key]), "the ranks must hold bit-identical sums and take identical Newton decisions"
        x = r[0][key]
        assert abs(x[0] - m["lnl"]) <= gu.LNL_RTOL * abs(m["lnl"])
        tol = gu.partial_rtol(g) * max(abs(m["ddf_root"]), abs(m["df_root"]), 1.0)
        assert abs(x[1] - m["df_root"]) <= tol and abs(x[2] - m["ddf_root"]) <= tol
        assert abs(x[3] - m["lnl"]) <= gu.LNL_RTOL * abs(m["lnl"])
        assert abs(x[4] - m["oab1_lnl"]) <= gu.LNL_RTOL * abs(m["oab1_lnl"])
        for (a, b), ln in zip(g["traversal"], x[5:]):
            assert abs(ln - ref[(int(a), int(b))]) <= gu.BRLEN_ATOL
