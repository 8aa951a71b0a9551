#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native likelihood engine (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (config.workload): BASELINE.json configs[1] — synthetic DNA, 100 taxa x 1,000,000 patterns, GTR+G4,
patterns sharded over the N ranks (strong scaling, one process per GPU under torchrun; only the scalar
lnL/df/ddf sums cross NVLink in one NCCL all-reduce per evaluation).

One STEP = one pass of the hot path over the whole alignment:
    clearAllPartialLH + computeLikelihood()            every internal node's partial-likelihood vector + branch lnL
  + computeLikelihoodDerv() on the root-leaf branch    theta_all build + df/ddf reduction
metric  = partial-LH updates/s, one update = one (pattern x rate category x internal node) unit
value   = inputs resident in HBM, device-timed (CUDA events, max over ranks)
e2e     = the same step through the C ABI with HOST buffers: the tip-state matrix goes up from pinned host memory
          (lkcuda_set_tip_states_async: chunked on a copy stream, overlapped with the traversal kernel), per-pattern
          log-likelihoods and the sums come back, all inside the timed region
roofline= the partial-likelihood level kernel (dominant): algorithmic bytes / CUDA-event time
cpu_baseline = the UNMODIFIED reference (oracle/_ref/ref_harness, AVX + OpenMP, all host cores) on a bounded sample
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NTAXA = int(os.environ.get("LK_BENCH_TAXA", 100))
NPTN = int(os.environ.get("LK_BENCH_PATTERNS", 1000000))
NCAT = 4
SEED = 1
UPLOAD_CHUNKS = int(os.environ.get("LK_BENCH_UPLOAD_CHUNKS", 0))     # 0 = about 2 MB per chunk, 2..16 chunks
CPU_SAMPLE_SITES = int(os.environ.get("LK_BENCH_CPU_SITES", 200000))
WORKLOAD = "synthetic DNA %d taxa x %d patterns, GTR+G4 (BASELINE configs[1])" % (NTAXA, NPTN)


def make_inputs(nptn):
    from iqtree_b200 import modelinputs as mi
    names = ["t%d" % i for i in range(NTAXA)]
    nwk = mi.random_tree_newick(NTAXA, SEED, names)
    states = mi.simulate_patterns(nwk, names, nptn, 4, SEED)
    model = mi.gtr_g4_inputs(SEED, ncat=NCAT)
    return names, nwk, states, model


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.stop_flag = False
        self.proc = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
                if self.stop_flag:
                    break
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            try:
                self.proc.terminate()
            except Exception:
                pass
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, val in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def measured_hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def write_sample_files(tmp, names, nwk, states, nsites):
    lut = np.array(list("ACGT"))
    aln = os.path.join(tmp, "sample.phy")
    with open(aln, "w") as f:
        f.write("%d %d\n" % (len(names), nsites))
        for i, n in enumerate(names):
            f.write("%s  %s\n" % (n, "".join(lut[states[i, :nsites]])))
    tre = os.path.join(tmp, "sample.nwk")
    open(tre, "w").write(nwk + "\n")
    return aln, tre


def run_reference(names, nwk, states, steps, warmup, threads=None):
    """time the unmodified reference's CPU kernels on a bounded sample (oracle/_ref/ref_harness time ...)"""
    harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    if not os.path.exists(harness):
        return None
    threads = threads or os.cpu_count() or 1
    nsites = min(CPU_SAMPLE_SITES, states.shape[1])
    tmp = tempfile.mkdtemp(prefix="lk_ref_")
    aln, tre = write_sample_files(tmp, names, nwk, states, nsites)
    cmd = [harness, "time", str(max(1, steps)), "0", "0", "-s", aln, "-m", "GTR+G4", "-te", tre, "-nt", str(threads), "-seed", "1",
           "-pre", os.path.join(tmp, "ref")]
    t0 = time.time()
    r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    wall = time.time() - t0
    out = None
    for line in r.stdout.split("\n"):
        if line.startswith('{"ref_time"'):
            out = json.loads(line)
    if out is None:
        sys.stderr.write(r.stdout[-2000:])
        return None
    out["wall_s"] = wall
    out["sample"] = "%d taxa x %d sites -> %d patterns, GTR+G4, %d timed steps (lnL + theta + df/ddf)" % (
        NTAXA, nsites, out["nptn"], out["lnl_reps"])
    out["updates_per_s"] = (NTAXA - 2) * out["nptn"] * NCAT / out["step_s_mean"]
    return out


def reference_arm(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return 0
    names, nwk, states, _ = make_inputs(min(CPU_SAMPLE_SITES, NPTN))
    ref = run_reference(names, nwk, states, args.steps + args.warmup, 0)
    if ref is None:
        emit({"impl": "reference", "unavailable": "oracle/_ref/ref_harness missing or failed"})
        return 0
    v = ref["updates_per_s"]
    line = {
        "impl": "reference", "metric": "partial-LH updates/s (patterns x cats x nodes)", "value": v, "unit": "updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ref["step_s_mean"] * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": ref["sample"], "kernel": "reference %s + OpenMP" % ref["kernel"]},
        "lnl_evals_per_s": 1.0 / ref["lnl_s_mean"],
        "cpu_baseline": {"value": v, "unit": "updates/s", "cores": ref["threads"], "kind": "reference", "sample": ref["sample"]},
        "e2e": {"value": v, "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


_REAL_STDOUT = None


def claim_stdout():
    """Everything libraries print to fd 1 (NCCL's version banner, torchrun notes) goes to stderr; the JSON line alone goes to stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist
    import iqtree_b200
    from iqtree_b200 import PhyloTree

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    lib = iqtree_b200.load_library()
    names, nwk, states, model = make_inputs(NPTN)
    freq = np.ones(NPTN)

    t = PhyloTree(lib=lib)
    t.setAlignment(states, freq, names=names, seq_type=0, nstates=4, state_unknown=18)
    t.readTreeString(nwk)
    t.setModel(0, NCAT, 1, model["eval"], model["evec"], model["inv_evec"], model["rates"], model["props"], None, model["state_freq"], 0.0)
    t.setParams(0.1 / NPTN if NPTN >= 100000 else 1e-6, 100.0)
    t.setDevices([local])
    t.setProcessShard(rank, world)
    # an explicit (non-default) stream: the engine, the CUDA events and NCCL's stream-ordering all refer to this one
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    t.setStream(stream.cuda_stream)
    t.initializeAllPartialLh()

    shard_check = None
    exchange = "none"
    if world > 1:
        # the engine keeps its 8-double result block inside this tensor; the NCCL variant of the exchange is one in-place sum of it
        resbuf = torch.zeros(8, dtype=torch.float64, device=dev)
        t.setResultBuffer(resbuf.data_ptr())

        def allreduce(host, dev_ptr, n):
            if host is not None:
                x = torch.from_numpy(host).to(dev)
                dist.all_reduce(x)
                host[:] = x.cpu().numpy()
                return
            assert dev_ptr == resbuf.data_ptr()
            dist.all_reduce(resbuf)

        # sanity of the sharded sum: local lnL values summed by hand == the all-reduced lnL
        t.clearAllPartialLH()
        loc = torch.tensor([t.computeLikelihood()], dtype=torch.float64, device=dev)
        dist.all_reduce(loc)
        # default exchange: fused into the reduction kernels over NVLink peer memory (lkcuda_attach_peers) - the mailbox handles
        # travel once through torch.distributed; LK_BENCH_EXCHANGE=nccl keeps one NCCL all-reduce per evaluation instead
        exchange = "nccl"
        if os.environ.get("LK_BENCH_EXCHANGE", "peer") == "peer":
            mine = torch.from_numpy(t.peerHandle()).to(dev)
            allh = [torch.zeros(64, dtype=torch.uint8, device=dev) for _ in range(world)]
            dist.all_gather(allh, mine)
            ok = torch.ones(1, dtype=torch.int32, device=dev)
            try:
                t.attachPeers(torch.stack(allh).cpu().numpy())
            except Exception as ex:
                sys.stderr.write("rank %d: peer attach failed (%s); falling back to NCCL\n" % (rank, ex))
                ok[0] = 0
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok[0]) == 1:
                exchange = "peer"
            else:       # CUDA IPC unavailable on some rank: every rank goes back to the NCCL exchange
                try:
                    t.detachPeers()
                except Exception:
                    pass
        if exchange == "nccl":
            t.setAllReduce(allreduce)
        t.clearAllPartialLH()
        glob = t.computeLikelihood()
        shard_check = bool(abs(glob - float(loc[0])) <= 1e-12 * abs(glob))
        if not shard_check:
            raise SystemExit("sharded lnL %r != sum of shard lnLs %r" % (glob, float(loc[0])))

    root_leaf, root_nei = 0, None
    trav = t.getTraversal()
    for a, b in trav:
        if a == 0 or b == 0:
            root_nei = int(b if a == 0 else a)
            break

    def step():
        t.clearAllPartialLH()
        lnl = t.computeLikelihood()
        df, ddf = t.computeLikelihoodDerv(root_leaf, root_nei, reset_theta=True)
        return lnl, df, ddf

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        res = step()
    t.resetCounters()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        res = step()
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = t.launchCount()

    # ---- the same steps once more with every launch bracketed by CUDA events: per-kernel device times (not the headline) ----
    t.resetCounters()
    t.setProfiling(True)
    barrier()
    for _ in range(args.steps):
        step()
    barrier()
    stats = t.kernelStats()

    # ---- lnL only (clearAllPartialLH + computeLikelihood): the full-tree traversal launch alone, for the roofline ----
    t.resetCounters()
    barrier()
    ev4, ev5 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev4.record(stream)
    for _ in range(args.steps):
        t.clearAllPartialLH()
        t.computeLikelihood()
    ev5.record(stream)
    barrier()
    ms_lnl = ev4.elapsed_time(ev5)
    stats_lnl = t.kernelStats()
    t.setProfiling(False)

    # ---- e2e: host buffers through the C ABI, copies inside the timed region --------------------------------
    b0, b1 = t.localPatterns()
    nloc = b1 - b0
    pinned_states = torch.from_numpy(np.ascontiguousarray(states[:, b0:b1])).pin_memory()
    pinned_plh = torch.empty(max(nloc, 1), dtype=torch.float64).pin_memory()
    eng = t.engine(0)
    c_u8p = ctypes.POINTER(ctypes.c_uint8)
    c_dp = ctypes.POINTER(ctypes.c_double)
    lnl_c = ctypes.c_double()

    nchunks = UPLOAD_CHUNKS if UPLOAD_CHUNKS > 0 else max(2, min(16, (NTAXA * nloc) // (2 << 20)))

    def e2e_step():
        # pipelined upload: the matrix goes up in `nchunks` pattern ranges on the engine's copy stream while the traversal kernel already works
        # on the ranges that have arrived (its producer warp polls a per-chunk counter); nothing is pre-loaded, every byte is re-sent
        rc = lib.lkcuda_set_tip_states_async(eng, ctypes.cast(pinned_states.data_ptr(), c_u8p), nchunks)
        if rc:
            raise RuntimeError(lib.lkcuda_last_error(eng).decode())
        t.clearAllPartialLH()
        # the per-pattern log-likelihoods travel to the pinned host buffer while the df/ddf part of the step runs
        rc = lib.lktree_compute_likelihood_async_fetch(t.h, ctypes.byref(lnl_c), ctypes.cast(pinned_plh.data_ptr(), c_dp))
        if rc:
            raise RuntimeError(lib.lktree_last_error(t.h).decode())
        d = t.computeLikelihoodDerv(root_leaf, root_nei, reset_theta=True)
        rc = lib.lktree_wait_pattern_lh(t.h)
        if rc:
            raise RuntimeError(lib.lktree_last_error(t.h).decode())
        return lnl_c.value, d[0], d[1]

    e2e_step()
    barrier()
    ev2, ev3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev2.record(stream)
    for _ in range(args.steps):
        res2 = e2e_step()
    ev3.record(stream)
    barrier()
    ms_e2e = ev2.elapsed_time(ev3)
    clocks = sampler.finish() if sampler else None
    if os.environ.get("LK_BENCH_E2E_TRACE") and rank == 0:       # host-side wall time of the three calls of an e2e step (diagnostic)
        acc = [0.0, 0.0, 0.0]
        for _ in range(args.steps):
            t0 = time.perf_counter()
            lib.lkcuda_set_tip_states_async(eng, ctypes.cast(pinned_states.data_ptr(), c_u8p), nchunks)
            t.clearAllPartialLH()
            t1 = time.perf_counter()
            lib.lktree_compute_likelihood(t.h, ctypes.byref(lnl_c), ctypes.cast(pinned_plh.data_ptr(), c_dp))
            t2 = time.perf_counter()
            t.computeLikelihoodDerv(root_leaf, root_nei, reset_theta=True)
            t3 = time.perf_counter()
            acc[0] += t1 - t0; acc[1] += t2 - t1; acc[2] += t3 - t2
        sys.stderr.write("e2e trace (ms per step): upload call %.4f, computeLikelihood+fetch %.4f, computeLikelihoodDerv %.4f, chunks %d\n" % (
            acc[0] / args.steps * 1e3, acc[1] / args.steps * 1e3, acc[2] / args.steps * 1e3, nchunks))

    tt = torch.tensor([ms, ms_e2e, ms_lnl], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_lnl = float(tt[0]), float(tt[1]), float(tt[2])

    if rank == 0:
        updates = (NTAXA - 2) * NPTN * NCAT
        value = updates * args.steps / (ms * 1e-3)
        e2e_value = updates * args.steps / (ms_e2e * 1e-3)
        peak, peak_src = measured_hbm_peak()
        n_p, b_p, ms_p = stats_lnl["partial"]      # the full-tree traversal launches of the lnL-only loop
        achieved = (b_p / 1e9) / (ms_p * 1e-3) if ms_p > 0 else 0.0
        nloc_all = t.localPatterns()[1] - t.localPatterns()[0]
        # what the traversal kernel is designed to move: every node's vector + scale_num written once, one byte per tip child read
        traffic_model = nloc_all * ((NTAXA - 2) * (4 * NCAT * 8 + 2) + NTAXA)
        traffic_ncu = 12.79e9 if (world == 1 and NTAXA == 100 and NPTN == 1000000) else None   # profiles/r1_trv_ncu_summary.txt
        kinds = {k: {"launches": v[0], "algorithmic_gb": v[1] / 1e9, "device_ms": v[2],
                     "gbps": (v[1] / 1e9) / (v[2] * 1e-3) if v[2] > 0 else None} for k, v in stats.items() if v[0]}
        line = {
            "metric": "partial-LH updates/s (patterns x cats x nodes)", "value": value, "unit": "updates/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "ntaxa": NTAXA, "patterns": NPTN, "ncat": NCAT, "sharding": "patterns x%d" % world, "exchange": exchange,
                       "step": "clearAllPartialLH + computeLikelihood + computeLikelihoodDerv(root branch, theta rebuilt)",
                       "l2": "inputs larger than L2: ~%.1f GB streamed per step per GPU vs 126 MB L2, no flush needed" % (
                           sum(v[1] for v in stats.values()) / 1e9 / max(args.steps, 1))},
            "steps_per_s": args.steps / (ms * 1e-3),
            "lnl_evals_per_s": args.steps / (ms_lnl * 1e-3), "ms_per_lnl_eval": ms_lnl / args.steps,
            "lnl": res[0], "df": res[1], "ddf": res[2], "shard_sum_check": shard_check,
            "e2e": {"value": e2e_value, "unit": "updates/s", "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(NTAXA * nloc), "d2h_bytes_per_step": int(8 * nloc + 3 * 64)},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "traverse_s4_kernel (whole post-order schedule in one launch; full-tree lnL launches only)",
                         "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak if peak else None, "traffic": traffic_ncu, "peak_source": peak_src,
                         "launches": int(n_p), "algorithmic_bytes_per_launch": b_p / max(n_p, 1), "avg_launch_ms": ms_p / max(n_p, 1),
                         "traffic_model_bytes_per_launch": traffic_model,
                         "hbm_gbps_on_model_traffic": (traffic_model / 1e9) / (ms_p / max(n_p, 1) * 1e-3) if ms_p > 0 else None,
                         "note": "achieved counts ALGORITHMIC bytes (SURVEY 8d: children read + result written per node); the kernel keeps "
                                 "children in shared memory/registers and only writes each node once, so DRAM traffic is ~0.38x algorithmic "
                                 "and frac can exceed 1"},
            "kernels": kinds,
            "clocks": clocks,
        }
        if not args.no_cpu_baseline and world == 1:      # the contract asks for it on rank 0 at N=1 only (at N>1 the other ranks' polling threads share the cores)
            ref = run_reference(names, nwk, states, 6, 0)
            if ref is not None:
                line["cpu_baseline"] = {"value": ref["updates_per_s"], "unit": "updates/s", "cores": ref["threads"], "kind": "reference",
                                        "sample": ref["sample"], "lnl_evals_per_s_on_sample": 1.0 / ref["lnl_s_mean"]}
            else:
                line["cpu_baseline"] = None
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    t.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
