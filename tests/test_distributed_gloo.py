"""CPU suite: the one-process-per-GPU pattern-sharding path of the host mirror, world_size 2 over gloo.
Each rank owns half of the patterns (engine = the CPU oracle library here, the CUDA library on a GPU box) and the
8-double result block of every reduction is summed with one all-reduce — the only exchange of the path."""
import ctypes
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
    from oracle.oracle_lib import load_oracle_library

    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = gu.Golden(case)
    t = gu.build_tree(g, load_oracle_library(), shard=(rank, world))

    def allreduce(host, dev_ptr, n):
        if host is None:     # the oracle engine's "device" result block is host memory
            host = np.ctypeslib.as_array((ctypes.c_double * n).from_address(dev_ptr))
        x = torch.from_numpy(host.copy())
        dist.all_reduce(x)
        host[:] = x.numpy()

    t.setAllReduce(allreduce)
    t.initializeAllPartialLh()
    b, e = t.localPatterns()
    lnl, plh = t.computeLikelihood(pattern_lh=True)
    a, bb = g.meta["derv_branch"]
    df, ddf = t.computeLikelihoodDerv(a, bb, reset_theta=True)
    fb = t.computeLikelihoodFromBuffer()
    t.clearAllPartialLH()
    t.computeLikelihood()
    oab = t.optimizeAllBranches(1)
    lens = [t.getBranchLength(int(x), int(y)) for (x, y) in t.getTraversal()]
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), lnl=lnl, df=df, ddf=ddf, fb=fb, oab=oab, lens=np.array(lens),
             begin=b, end=e, plh=plh)
    t.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("case", ["dna_example_gtr_i_g4", "aa_lg4x", "morph_mk_asc"])
def test_two_ranks_gloo(case, tmp_path):
    import torch.multiprocessing as mp
    import golden_util as gu
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), case, str(tmp_path)), nprocs=world, join=True)
    g = gu.Golden(case)
    m = g.meta
    r = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % i)) for i in range(world)]
    # contiguous, disjoint, complete pattern ranges
    assert int(r[0]["begin"]) == 0 and int(r[0]["end"]) == int(r[1]["begin"]) and int(r[1]["end"]) == m["nptn"]
    for x in r:     # every rank holds the global sums
        assert abs(float(x["lnl"]) - m["lnl"]) <= gu.LNL_RTOL * abs(m["lnl"])
        tol = gu.partial_rtol(g) * max(abs(m["ddf_root"]), abs(m["df_root"]), 1.0)
        assert abs(float(x["df"]) - m["df_root"]) <= tol and abs(float(x["ddf"]) - m["ddf_root"]) <= tol
        assert abs(float(x["oab"]) - m["oab1_lnl"]) <= gu.LNL_RTOL * abs(m["oab1_lnl"])
    assert np.array_equal(r[0]["lens"], r[1]["lens"])           # identical Newton sequences on both ranks
    plh = np.concatenate([r[0]["plh"], r[1]["plh"]])
    gu.rel_close(plh, g["pattern_lh_scaled"], gu.PTN_RTOL, "pattern_lh over the two shards")
    ref = gu.parse_adjacency(g["tree_adj_oab1"])
    trav = g["traversal"]
    for (a, b), ln in zip(trav, r[0]["lens"]):
        assert abs(ln - ref[(int(a), int(b))]) <= gu.BRLEN_ATOL


def _em_worker(rank, world, port, case, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import json
    import torch
    import torch.distributed as dist
    import golden_util as gu
    from oracle.oracle_lib import load_oracle_library

    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = gu.Golden(case)
    m = json.loads(str(np.load(os.path.join(gu.GOLDEN_DIR, "em", case + ".npz"))["meta"]))
    t = gu.build_tree(g, load_oracle_library(), shard=(rank, world))

    def allreduce(host, dev_ptr, n):
        if host is None:
            host = np.ctypeslib.as_array((ctypes.c_double * n).from_address(dev_ptr))
        x = torch.from_numpy(host.copy())
        dist.all_reduce(x)
        host[:] = x.numpy()

    t.setAllReduce(allreduce)
    t.initializeAllPartialLh()
    t.computeLikelihood()
    lnl, rates, props = t.optimizeRateFreeEM(m["ncat"], copy_precision=m["num_precision"])
    np.savez(os.path.join(out_dir, "em_rank%d.npz" % rank), lnl=lnl, rates=rates, props=props)
    t.close()
    dist.barrier()
    dist.destroy_process_group()


def test_rate_free_em_two_ranks_gloo(tmp_path):
    """RateFree::optimizeWithEM with the patterns split over two ranks: the E-step sums and every likelihood of the one-category
    copies go through the all-reduce; both ranks take the same Brent decisions and reach the reference's rates"""
    import torch.multiprocessing as mp
    import golden_util as gu
    case, world = "dna_example_hky_r3", 2
    mp.spawn(_em_worker, args=(world, _free_port(), case, str(tmp_path)), nprocs=world, join=True)
    z = np.load(os.path.join(gu.GOLDEN_DIR, "em", case + ".npz"))
    m = __import__("json").loads(str(z["meta"]))
    r = [np.load(os.path.join(str(tmp_path), "em_rank%d.npz" % i)) for i in range(world)]
    assert np.array_equal(r[0]["rates"], r[1]["rates"]) and np.array_equal(r[0]["props"], r[1]["props"])
    for x in r:
        assert abs(float(x["lnl"]) - m["lnl_after"]) <= 1e-9 * abs(m["lnl_after"])
        gu.rel_close(x["rates"], z["rates"], 1e-7, "+R rates, two ranks")
        gu.rel_close(x["props"], z["props"], 1e-7, "+R proportions, two ranks")
