#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native likelihood engine (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (config.workload): BASELINE.json configs[1] — synthetic DNA, 100 taxa x 1,000,000 DISTINCT patterns, GTR+G4, patterns sharded
over the N ranks (strong scaling, one process per GPU under torchrun; only the scalar lnL/df/ddf sums cross NVLink).  The patterns are
distinct columns in first-occurrence order, so the reference (which compresses identical columns) sees exactly the same 1M patterns.

One STEP = one pass of the hot path over the whole alignment:
    clearAllPartialLH + computeLikelihood()            every internal node's partial-likelihood vector + branch lnL
  + computeLikelihoodDerv() on the root-leaf branch    theta_all build + df/ddf reduction
metric  = partial-LH updates/s, one update = one (pattern x rate category x internal node) unit
value   = inputs resident in HBM, device-timed (CUDA events, max over ranks)
e2e     = the same step through the C ABI with HOST buffers: the tip-state matrix goes up from pinned host memory
          (lkcuda_set_tip_states_async: chunked on a copy stream, overlapped with the traversal kernel), per-pattern
          log-likelihoods and the sums come back, all inside the timed region
roofline= the dominant kernel (whole-traversal kernel): DRAM traffic the kernel is designed to move / CUDA-event time, against the
          measured HBM peak; the SURVEY 8d algorithmic-byte ratio is kept beside it (frac_algorithmic)
At N = 1 the line also carries
cpu_baseline  the UNMODIFIED reference (oracle/_ref/ref_harness, AVX + OpenMP, all host cores) on the SAME alignment file, full size
dropin        the reference's OWN computeLikelihood / computeLikelihoodDerv / optimizeAllBranches running on the engine
              (oracle/_ref/ref_harness_cuda = the reference sources + integration/, option -cuda) on that file
parity        lnL / df / ddf of `dropin` against `cpu_baseline`: same alignment, tree and model; the run FAILS when lnL differs by > 1e-9 rel
configs       BASELINE configs 3-5 (protein 200 x 200k LG+G4, codon 64 x 50k GY and MG+G4, LG+C20+G4 100 x 10k, 50-partition DNA+protein
              PartitionModel): reference on the host cores and reference-on-engine on the same files, device time per kernel kind, real
              flops by node case against the FP64 peak measured in this run (tools/fp64_peak), lnL parity
With --impl reference: the reference's CPU kernels on the full workload (rank 0 only).
"""
import argparse
import ctypes
import json
import os
import shutil
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
WITH_CONFIGS = int(os.environ.get("LK_BENCH_CONFIGS", 1))            # BASELINE configs 3-5 beside the headline (N = 1)
WITH_DROPIN = int(os.environ.get("LK_BENCH_DROPIN", 1))
CPU_STEPS = int(os.environ.get("LK_BENCH_CPU_STEPS", 4))
WORKLOAD = "synthetic DNA %d taxa x %d patterns, GTR+G4 (BASELINE configs[1])" % (NTAXA, NPTN)
METRIC = "partial-LH updates/s (patterns x cats x nodes)"
HARNESS_REF = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
HARNESS_CUDA = os.path.join(ROOT, "oracle", "_ref", "ref_harness_cuda")
HARNESS_FMA = os.path.join(ROOT, "oracle", "_ref", "ref_harness_fma")
AA = "ARNDCQEGHILKMFPSTWYV"
SENSE = [a + b + c for a in "ACGT" for b in "ACGT" for c in "ACGT" if a + b + c not in ("TAA", "TAG", "TGA")]


def make_inputs(nptn, ntaxa=None, nstates=4, seed=SEED):
    from iqtree_b200 import modelinputs as mi
    ntaxa = ntaxa or NTAXA
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, SEED, names)
    states = mi.simulate_distinct_patterns(nwk, names, nptn, nstates, seed)
    return names, nwk, states


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


def measured_fp64_peak():
    """FP64 denominators measured on THIS box in THIS run: tools/fp64_peak (DFMA loop, mma.sync DMMA loop; built by __graft_entry__.build())"""
    exe = os.path.join(ROOT, "tools", "fp64_peak")
    if not os.path.exists(exe):
        return None
    try:
        r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=120)
        for line in r.stdout.split("\n"):
            if line.startswith("{"):
                return json.loads(line)
    except Exception:
        pass
    return None


def measured_write_peak(gb=8.0):
    """Ceiling of a WRITE-ONLY HBM stream on THIS box in THIS run (tools/write_peak.cu: the traversal kernel's 256-bit streaming store from
    148 x 4 CTAs): the DNA traversal writes 99 % of its DRAM traffic, the copy figure of MEASURED_PEAKS.json is half reads"""
    exe = os.path.join(ROOT, "tools", "write_peak")
    if not os.path.exists(exe):
        return None
    try:
        r = subprocess.run([exe, str(gb)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=120)
        best = None
        for line in r.stdout.split("\n"):
            if line.startswith("{"):
                d = json.loads(line)
                if d["kernel"].startswith("st.v4.f64"):
                    best = d["best_gbps"] if best is None else max(best, d["best_gbps"])
        return best
    except Exception:
        return None


# ------------------------------------------------------------------ files for the reference binaries ------------------------------------------------------------------

def write_phylip(path, names, states, alphabet):
    if isinstance(alphabet, str):
        lut = np.frombuffer(alphabet.encode(), dtype=np.uint8)
        width = 1
    else:
        lut = np.array([s.encode() for s in alphabet], dtype="S%d" % len(alphabet[0]))
        width = len(alphabet[0])
    with open(path, "wb") as f:
        f.write(b"%d %d\n" % (len(names), states.shape[1] * width))
        for i, n in enumerate(names):
            f.write(n.encode() + b"  " + lut[states[i]].tobytes() + b"\n")


def run_harness(exe, args, reps, oab_reps=0, warmup=0, threads=None, cuda=False, cwd=None):
    """`ref_harness[_cuda] time`: the reference's own loop of clearAllPartialLH + computeLikelihood (+ theta / df / ddf, + optimizeAllBranches(1))"""
    if not os.path.exists(exe):
        return None
    threads = threads or os.cpu_count() or 1
    cmd = [exe, "time", str(max(1, reps + warmup)), "0", str(oab_reps)] + args + ["-nt", str(threads), "-seed", "1", "-pre", os.path.join(cwd, "run_%d" % int(cuda))]
    if cuda:
        cmd.append("-cuda")
    env = dict(os.environ, LK_REF_WARMUP=str(warmup))
    t0 = time.time()
    r = subprocess.run(cmd, cwd=cwd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    wall = time.time() - t0
    out = None
    for line in r.stdout.split("\n"):
        if line.startswith('{"ref_time"'):
            out = json.loads(line)
    if out is None:
        sys.stderr.write("%s failed:\n%s\n" % (" ".join(cmd[:6]), r.stdout[-2000:]))
        return None
    out["wall_s"] = wall
    return out


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


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


def headline_files(tmp, names, nwk, states):
    aln, tre = os.path.join(tmp, "dna.phy"), os.path.join(tmp, "dna.nwk")
    write_phylip(aln, names, states, "ACGT")
    open(tre, "w").write(nwk + "\n")
    return ["-s", aln, "-m", "GTR+G4", "-te", tre]


def reference_arm(args):
    """the reference's CPU kernels (AVX + OpenMP, all host cores) on the full workload; rank 0 only"""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return 0
    names, nwk, states = make_inputs(NPTN)
    tmp = tempfile.mkdtemp(prefix="lk_ref_")
    try:
        ref = run_harness(HARNESS_REF, headline_files(tmp, names, nwk, states), args.steps, warmup=args.warmup, cwd=tmp)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    if ref is None:
        emit({"impl": "reference", "unavailable": "oracle/_ref/ref_harness missing or failed"})
        return 0
    v = (NTAXA - 2) * ref["nptn"] * NCAT / ref["step_s_mean"]
    sample = "the full workload: %d taxa x %d patterns, GTR+G4, %d timed steps after %d warm-up (lnL + theta + df/ddf per step)" % (
        NTAXA, ref["nptn"], ref["lnl_reps"], args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ref["step_s_mean"] * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "ntaxa": NTAXA, "patterns": ref["nptn"], "ncat": NCAT,
                   "kernel": "reference %s + OpenMP, LM_PER_NODE buffers" % ref["kernel"]},
        "lnl_evals_per_s": 1.0 / ref["lnl_s_mean"], "lnl": ref["lnl"], "df": ref["df"], "ddf": ref["ddf"],
        "cpu_baseline": {"value": v, "unit": "updates/s", "cores": ref["threads"], "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ------------------------------------------------------------------ BASELINE configs 3-5 ------------------------------------------------------------------

def bind_to_gpu_numa_node(local):
    """one process per GPU: run on the CPUs of the GPU's NUMA node, so that the pinned staging buffers of the end-to-end step (first
    touch) and the threads that feed the copy engine are local to the GPU's PCIe root - with 8 ranks re-sending their slice of the
    alignment every step the upload is bound by host memory, not by the links.  Best effort: returns the node or None."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bdf).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


def config_list():
    return [
        dict(name="protein 200 taxa x 200k patterns, LG+G4", key="protein_lg_g4", ntaxa=200, nptn=200000, S=20, blocks=4, alphabet=AA, margs=["-m", "LG+G4"], reps=2, bound="fp64"),
        dict(name="codon 64 taxa x 50k patterns, GY (64-state layout)", key="codon_gy", ntaxa=64, nptn=50000, S=61, blocks=1, alphabet=SENSE, margs=["-st", "CODON", "-m", "GY"], reps=3, bound="fp64"),
        dict(name="codon 64 taxa x 50k patterns, MG+G4", key="codon_mg_g4", ntaxa=64, nptn=50000, S=61, blocks=4, alphabet=SENSE, margs=["-st", "CODON", "-m", "MG+G4"], reps=2, bound="fp64"),
        dict(name="protein mixture 100 taxa x 10k patterns, LG+C20+G4", key="lg_c20_g4", ntaxa=100, nptn=10000, S=20, blocks=80, alphabet=AA, margs=["-m", "LG+C20+G4"], reps=2, bound="fp64"),
    ]


def kernel_summary(res, fp64):
    """device time per lnL evaluation from the per-kind CUDA-event statistics of a `ref_harness_cuda time` run"""
    k = res.get("kernels") or {}
    reps = max(res.get("profiled_steps", 0) or res.get("lnl_reps", 1), 1)
    out = {}
    out["device_ms_per_step"] = sum(v["device_ms"] for v in k.values()) / reps
    out["step"] = "clearAllPartialLH + computeLikelihood + computeLikelihoodDerv on the root-leaf branch (theta rebuilt; a second, partial traversal towards that branch)"
    out["kernels"] = {kk: {"launches_per_step": v["launches"] / reps, "device_ms_per_step": v["device_ms"] / reps,
                           "algorithmic_gb_per_step": v["bytes"] / reps / 1e9} for kk, v in k.items()}
    if "partial" in k and k["partial"]["device_ms"] > 0:
        p = k["partial"]
        tf = p["flops"] / (p["device_ms"] * 1e-3) / 1e12
        out["partial_tflops"] = tf
        out["partial_flops_per_eval"] = p["flops"] / reps
        if fp64:
            peak = fp64.get("dmma_m8n8k4_tflops")
            out["roofline"] = {"bound": "fp64 tensor (mma.sync DMMA)", "achieved": tf, "peak": peak, "unit": "TFLOP/s",
                               "frac": tf / peak if peak else None, "peak_source": "tools/fp64_peak DMMA m8n8k4 loop, this run",
                               "flops": "real flops by node case from the op list (2S^2 per internal child + 2S^2 back to state space + S per product)"}
    return out


def run_config(cfg, tmp, fp64):
    names, nwk, states = make_inputs(cfg["nptn"], cfg["ntaxa"], cfg["S"], seed=SEED + 3)
    d = os.path.join(tmp, cfg["key"])
    os.makedirs(d, exist_ok=True)
    aln, tre = os.path.join(d, "a.phy"), os.path.join(d, "a.nwk")
    write_phylip(aln, names, states, cfg["alphabet"])
    open(tre, "w").write(nwk + "\n")
    fargs = ["-s", aln] + cfg["margs"] + ["-te", tre]
    cpu = run_harness(HARNESS_REF, fargs, cfg["reps"], cwd=d)
    gpu = run_harness(HARNESS_CUDA, fargs, 5, warmup=1, cuda=True, threads=1, cwd=d)
    shutil.rmtree(d, ignore_errors=True)
    out = {"config": cfg["name"], "ntaxa": cfg["ntaxa"], "patterns": cfg["nptn"], "blocks_per_pattern": cfg["blocks"]}
    units = (cfg["ntaxa"] - 2) * cfg["nptn"] * cfg["blocks"]
    out["updates_per_lnl_eval"] = units
    if cpu:
        out["reference"] = {"kernel": cpu["kernel"], "cores": cpu["threads"], "lnl_evals_per_s": 1.0 / cpu["lnl_s_mean"], "ms_per_lnl_eval": cpu["lnl_s_mean"] * 1e3,
                            "updates_per_s": units / cpu["lnl_s_mean"], "steps": cpu["lnl_reps"], "input": "same files, full size"}
    if gpu:
        ks = kernel_summary(gpu, fp64)
        out["engine"] = {"through": "the reference's own computeLikelihood on the engine (ref_harness_cuda -cuda)",
                         "ms_per_lnl_eval_host_wall": gpu["lnl_s_mean"] * 1e3, "lnl_evals_per_s": 1.0 / gpu["lnl_s_mean"],
                         "updates_per_s": units / gpu["lnl_s_mean"]}
        out["engine"].update(ks)
    if cpu and gpu:
        out["parity"] = {"lnl_reference": cpu["lnl"], "lnl_engine": gpu["lnl"], "lnl_rel": rel(gpu["lnl"], cpu["lnl"]),
                         "df_rel": rel(gpu["df"], cpu["df"]), "ddf_rel": rel(gpu["ddf"], cpu["ddf"]), "ok": rel(gpu["lnl"], cpu["lnl"]) <= 1e-9}
        out["speedup_lnl_eval"] = cpu["lnl_s_mean"] / gpu["lnl_s_mean"]
    return out


def run_partition_config(tmp, fp64):
    """BASELINE configs[4]: 25 DNA GTR+G4 + 25 protein LG+G4 partitions of 20k patterns, 100 taxa, unlinked PartitionModel (-sp):
    PhyloSuperTree::computeLikelihood (phylosupertree.cpp:955-980), one engine per partition under -cuda"""
    from iqtree_b200 import modelinputs as mi
    d = os.path.join(tmp, "part50")
    os.makedirs(d, exist_ok=True)
    ntaxa, nptn, npart = 100, 20000, 50
    names = ["t%d" % i for i in range(ntaxa)]
    nwk = mi.random_tree_newick(ntaxa, SEED, names)
    tre = os.path.join(d, "p.nwk")
    open(tre, "w").write(nwk + "\n")
    lines, cp = ["#nexus", "begin sets;"], []
    for i in range(npart):
        S = 4 if i % 2 == 0 else 20
        st = mi.simulate_distinct_patterns(nwk, names, nptn, S, 10 + i)
        write_phylip(os.path.join(d, "p%d.phy" % i), names, st, "ACGT" if S == 4 else AA)
        lines.append("  charset part%d = p%d.phy: 1-%d;" % (i, i, nptn))
        cp.append("%s:part%d" % ("GTR+G4" if S == 4 else "LG+G4", i))
    lines += ["  charpartition mine = " + ", ".join(cp) + ";", "end;"]
    nex = os.path.join(d, "p.nex")
    open(nex, "w").write("\n".join(lines) + "\n")
    fargs = ["-sp", nex, "-te", tre]
    cpu = run_harness(HARNESS_REF, fargs, 2, cwd=d)
    gpu = run_harness(HARNESS_CUDA, fargs, 5, warmup=1, cuda=True, threads=8, cwd=d)
    shutil.rmtree(d, ignore_errors=True)
    out = {"config": "50-partition PartitionModel: 25 DNA GTR+G4 + 25 protein LG+G4, 100 taxa x 20k patterns each (-sp)"}
    if cpu:
        u = cpu["updates_per_eval"]
        out["updates_per_lnl_eval"] = u
        out["reference"] = {"cores": cpu["threads"], "ms_per_lnl_eval": cpu["lnl_s_mean"] * 1e3, "lnl_evals_per_s": 1.0 / cpu["lnl_s_mean"], "updates_per_s": u / cpu["lnl_s_mean"]}
    if gpu:
        u = gpu["updates_per_eval"]
        out["engine"] = {"through": "the reference's own PhyloSuperTree::computeLikelihood, one engine per partition (ref_harness_cuda -cuda, 8 host threads)",
                         "ms_per_lnl_eval_host_wall": gpu["lnl_s_mean"] * 1e3, "lnl_evals_per_s": 1.0 / gpu["lnl_s_mean"], "updates_per_s": u / gpu["lnl_s_mean"]}
    if cpu and gpu:
        out["parity"] = {"lnl_reference": cpu["lnl"], "lnl_engine": gpu["lnl"], "lnl_rel": rel(gpu["lnl"], cpu["lnl"]), "ok": rel(gpu["lnl"], cpu["lnl"]) <= 1e-9}
        out["speedup_lnl_eval"] = cpu["lnl_s_mean"] / gpu["lnl_s_mean"]
    return out


# ------------------------------------------------------------------ the engine arm ------------------------------------------------------------------

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
    from iqtree_b200 import modelinputs as mi

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    lib = iqtree_b200.load_library()
    names, nwk, states = make_inputs(NPTN)
    model = mi.gtr_g4_inputs(SEED, ncat=NCAT)
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
    res = t.benchSteps(root_leaf, root_nei, args.steps)      # the same three calls per step, issued by the C++ driver like a C++ host would
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

    # ---- optimizeAllBranches(1): one Newton-Raphson pass over all 2n-3 branches, host wall time ----
    oab_ms = None
    if world == 1:
        t.clearAllPartialLH()
        t.computeLikelihood()
        t0 = time.perf_counter()
        oab_lnl = t.optimizeAllBranches(1)
        torch.cuda.synchronize()
        oab_ms = (time.perf_counter() - t0) * 1e3


    tt = torch.tensor([ms, ms_e2e, ms_lnl], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_lnl = float(tt[0]), float(tt[1]), float(tt[2])
    nloc_all = t.localPatterns()[1] - t.localPatterns()[0]
    t.close()
    if world > 1:
        dist.barrier()

    failed = None
    if rank == 0:
        updates = (NTAXA - 2) * NPTN * NCAT
        value = updates * args.steps / (ms * 1e-3)
        e2e_value = updates * args.steps / (ms_e2e * 1e-3)
        peak, peak_src = measured_hbm_peak()
        n_p, b_p, ms_p = stats_lnl["partial"]      # the full-tree traversal launches of the lnL-only loop
        avg_ms = ms_p / max(n_p, 1)
        alg_gbps = (b_p / 1e9) / (ms_p * 1e-3) if ms_p > 0 else 0.0
        # what the traversal kernel is designed to move per launch: every node's vector + scale_num written once, one byte per tip child read
        traffic_model = nloc_all * ((NTAXA - 2) * (4 * NCAT * 8 + 2) + NTAXA)
        model_gbps = (traffic_model / 1e9) / (avg_ms * 1e-3) if avg_ms > 0 else 0.0
        traffic, traffic_src = None, "no ncu capture of this kernel committed for this shape"
        tp = os.path.join(ROOT, "profiles", "r2_trv_ncu_traffic.json")
        if os.path.exists(tp) and world == 1 and NTAXA == 100 and NPTN == 1000000:
            try:
                tj = json.load(open(tp))
                traffic = tj["dram_bytes_per_launch"]
                traffic_src = "profiles/r2_trv_ncu_traffic.json (ncu --set full of commit %s: dram__bytes_read.sum + dram__bytes_write.sum)" % tj.get("commit", "?")
            except Exception:
                pass
        kinds = {k: {"launches": v[0], "algorithmic_gb": v[1] / 1e9, "device_ms": v[2],
                     "gbps": (v[1] / 1e9) / (v[2] * 1e-3) if v[2] > 0 else None} for k, v in stats.items() if v[0]}
        line = {
            "metric": METRIC, "value": value, "unit": "updates/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "ntaxa": NTAXA, "patterns": NPTN, "ncat": NCAT, "sharding": "patterns x%d" % world, "exchange": exchange,
                       "numa_node_rank0": numa,
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
                         "achieved": model_gbps, "peak": peak, "unit": "GB/s", "frac": model_gbps / peak if peak else None,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "launches": int(n_p), "avg_launch_ms": avg_ms,
                         "bytes_per_launch": traffic_model,
                         "bytes_model": "DRAM traffic by design: every node's vector + scale_num written once, one byte per tip child read (children stay in registers / shared memory)",
                         "algorithmic_bytes_per_launch": b_p / max(n_p, 1), "achieved_algorithmic": alg_gbps,
                         "frac_algorithmic": alg_gbps / peak if peak else None,
                         "note": "frac is on the traffic the fused kernel moves; frac_algorithmic (SURVEY 8d: children read + result written per node) can exceed 1"},
            "kernels": kinds,
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            wp = measured_write_peak()
            if wp:
                line["roofline"]["write_only_peak"] = wp
                line["roofline"]["frac_of_write_only_peak"] = model_gbps / wp
                line["roofline"]["write_only_peak_source"] = "tools/write_peak.cu, this run: best of the st.global.v4.f64 streams over 8 GB (the kernel's own store instruction)"
        if oab_ms is not None:
            line["optimize_all_branches_1"] = {"ms": oab_ms, "lnl": oab_lnl, "branches": 2 * NTAXA - 3, "through": "harness driver (host-driven Newton steps, one launch per evaluation)"}
        if not args.no_cpu_baseline and world == 1:
            tmp = tempfile.mkdtemp(prefix="lk_bench_")
            try:
                fargs = headline_files(tmp, names, nwk, states)
                # the contract asks for the CPU baseline on rank 0 at N=1 only; the FULL workload, a few steps
                ref = run_harness(HARNESS_REF, fargs, CPU_STEPS, oab_reps=1, warmup=1, cwd=tmp)
                if ref is not None:
                    ups = (NTAXA - 2) * ref["nptn"] * NCAT / ref["step_s_mean"]
                    line["cpu_baseline"] = {"value": ups, "unit": "updates/s", "cores": ref["threads"], "kind": "reference",
                                            "sample": "the full workload (same alignment file): %d taxa x %d patterns, %d timed steps after 1 warm-up" % (NTAXA, ref["nptn"], ref["lnl_reps"]),
                                            "kernel": ref["kernel"], "ms_per_step": ref["step_s_mean"] * 1e3, "lnl_evals_per_s": 1.0 / ref["lnl_s_mean"],
                                            "optimize_all_branches_1_ms": ref["oab_s_mean"] * 1e3, "setup_s": ref["setup_s"]}
                else:
                    line["cpu_baseline"] = None
                # SURVEY 8(d) also asks for the "omp fma" build of the reference and for -nt 1 (bounded: one or two timed steps on the full workload)
                extra = {}
                if CPU_STEPS > 0:
                    fma = run_harness(HARNESS_FMA, fargs, max(1, CPU_STEPS // 2), warmup=1, cwd=tmp)
                    if fma is not None:
                        extra["omp_fma"] = {"value": (NTAXA - 2) * fma["nptn"] * NCAT / fma["step_s_mean"], "unit": "updates/s", "cores": fma["threads"],
                                            "kernel": fma["kernel"], "ms_per_step": fma["step_s_mean"] * 1e3, "steps": fma["lnl_reps"],
                                            "build": "oracle/Makefile.ref fma (-mavx -mfma for every unit, CMakeLists.txt:249-267)"}
                    one = run_harness(HARNESS_REF, fargs, 1, warmup=0, threads=1, cwd=tmp)
                    if one is not None:
                        extra["nt1"] = {"value": (NTAXA - 2) * one["nptn"] * NCAT / one["step_s_mean"], "unit": "updates/s", "cores": 1,
                                        "kernel": one["kernel"], "ms_per_step": one["step_s_mean"] * 1e3, "steps": one["lnl_reps"]}
                if extra and line.get("cpu_baseline"):
                    line["cpu_baseline"]["other_builds"] = extra
                if WITH_DROPIN:
                    dro = run_harness(HARNESS_CUDA, fargs, args.steps, oab_reps=1, warmup=2, cuda=True, threads=1, cwd=tmp)
                    if dro is not None:
                        dups = (NTAXA - 2) * dro["nptn"] * NCAT / dro["step_s_mean"]
                        line["dropin"] = {"what": "the reference's OWN PhyloTree::computeLikelihood + computeLikelihoodDerv (+ optimizeAllBranches(1)) on the engine: "
                                                  "IQ-TREE 1.4.3 sources + integration/iqtree-1.4.3-lkcuda.patch + integration/phylotree_cuda.cpp, option -cuda, LM_PER_NODE buffers",
                                          "value": dups, "unit": "updates/s", "ms_per_step_host_wall": dro["step_s_mean"] * 1e3,
                                          "lnl_evals_per_s": 1.0 / dro["lnl_s_mean"], "optimize_all_branches_1_ms": dro["oab_s_mean"] * 1e3,
                                          "kernels": kernel_summary(dro, None)["kernels"], "lnl": dro["lnl"], "df": dro["df"], "ddf": dro["ddf"], "setup_s": dro["setup_s"]}
                        if ref is not None:
                            par = {"lnl_reference": ref["lnl"], "lnl_engine": dro["lnl"], "lnl_rel": rel(dro["lnl"], ref["lnl"]),
                                   "df_rel": rel(dro["df"], ref["df"]), "ddf_rel": rel(dro["ddf"], ref["ddf"]),
                                   "oab1_lnl_rel": rel(dro["oab_lnl"], ref["oab_lnl"]),
                                   "what": "reference-on-engine vs the reference's AVX kernels: same alignment file (%d patterns), tree and model" % ref["nptn"],
                                   "tolerance": "lnL 1e-9 relative (north_star); df / ddf 1e-7"}
                            par["ok"] = bool(par["lnl_rel"] <= 1e-9 and par["df_rel"] <= 1e-7 and par["ddf_rel"] <= 1e-7)
                            line["parity"] = par
                            line["dropin"]["speedup_vs_reference_step"] = ref["step_s_mean"] / dro["step_s_mean"]
                            line["dropin"]["speedup_vs_reference_oab1"] = ref["oab_s_mean"] / dro["oab_s_mean"] if dro["oab_s_mean"] > 0 else None
                            if not par["ok"]:
                                failed = "parity: lnL rel %.3g, df rel %.3g, ddf rel %.3g" % (par["lnl_rel"], par["df_rel"], par["ddf_rel"])
                if WITH_CONFIGS:
                    fp64 = measured_fp64_peak()
                    line["fp64_peaks"] = fp64
                    cfgs = []
                    for cfg in config_list():
                        try:
                            cfgs.append(run_config(cfg, tmp, fp64))
                        except Exception as ex:
                            cfgs.append({"config": cfg["name"], "error": str(ex)})
                    try:
                        cfgs.append(run_partition_config(tmp, fp64))
                    except Exception as ex:
                        cfgs.append({"config": "50-partition PartitionModel", "error": str(ex)})
                    line["configs"] = cfgs
                    bad = [c["config"] for c in cfgs if c.get("parity") and not c["parity"]["ok"]]
                    if bad and not failed:
                        failed = "parity of configs: " + "; ".join(bad)
            finally:
                shutil.rmtree(tmp, ignore_errors=True)
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if failed:
        sys.stderr.write("bench.py: FAILED %s\n" % failed)
        return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())
