#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box): headline metrics, SASS opcode mix and the most-sampled instructions.
usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [n_top]"""
import collections
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 20
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed", "smsp__cycles_active.avg"]
for vals in rows[2:]:
    name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
    print("== kernel:", name[:100])
    for h, u, v in zip(hdr, units, vals):
        if h in want or h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
            if h.startswith("smsp__average_warps_issue_stalled"):
                if float(v or 0) < 0.15:
                    continue
                h = h.replace("smsp__average_warps_issue_stalled_", "stall ").replace("_per_issue_active.ratio", "")
            print("  %-70s %-10s %s" % (h, u, v))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if "Source" in r and "Address" in r)
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
ops, samp, tot, wf, data = collections.Counter(), collections.Counter(), 0, 0, []
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    s_ = r[ix["Source"]]
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", s_)
    op = m.group(2).split(".")[0] if m else "?"
    try:            # a report with several kernels repeats the header row; some columns are empty on non-executed lines
        n = int(float(r[ix["Instructions Executed"]] or 0))
        sm = int(float(r[ix["# Samples"]] or 0))
        wf += int(float(r[ix["L1 Wavefronts Shared"]] or 0)) if "L1 Wavefronts Shared" in ix else 0
    except ValueError:
        continue
    ops[op] += n; samp[op] += sm; tot += n
    data.append((sm, n, s_.strip()))
print("total warp instructions %d, shared-memory wavefronts %d" % (tot, wf))
for op, n in ops.most_common(ntop):
    print("  %-10s %12d %5.1f%%  samples %d" % (op, n, 100.0 * n / max(tot, 1), samp[op]))
print("-- most sampled instructions")
for sm, n, s_ in sorted(data, reverse=True)[:ntop]:
    print("  %6d %10d  %s" % (sm, n, s_[:100]))
