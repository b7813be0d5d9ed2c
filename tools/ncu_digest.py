#!/usr/bin/env python
"""Digest of one .ncu-rep (ncu --set full --import-source on): headline metrics, opcode mix, and the hottest SASS windows.
usage: python tools/ncu_digest.py gpurun_out/x.ncu-rep [window]"""
import csv, subprocess, sys, io

rep = sys.argv[1]
win = int(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u, v = rows[0], rows[1], rows[2]
print(v[h.index("Kernel Name")][:70], "grid", v[h.index("launch__grid_size")], "block", v[h.index("launch__block_size")])
keys = ["gpu__time_duration.sum", "launch__registers_per_thread", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_fp64.sum", "smsp__thread_inst_executed.sum"]
for a, b, c in zip(h, u, v):
    if a in keys or ("issue_stalled" in a and a.endswith("per_issue_active.ratio") and float(c or 0) > 0.3):
        print(f"  {a:90s} {b:14s} {c}")
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(sass)))
h, d = rows[1], rows[2:]
iS, iI, iT, iSrc = h.index("# Samples"), h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("Source")
tot, totI = sum(int(r[iS]) for r in d), sum(int(r[iI]) for r in d)
ops, fp64_thread, all_thread = {}, 0, 0
for r in d:
    t = r[iSrc].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    ops[op] = ops.get(op, 0) + int(r[iI])
    all_thread += int(r[iT])
    if op in ("DFMA", "DMUL", "DADD", "DSETP"): fp64_thread += int(r[iT])
print(f"SASS instrs {len(d)}  warp-inst {totI}  thread-inst {all_thread}  FP64 thread-inst {fp64_thread}")
print("  mix:", "  ".join(f"{k} {100 * x / totI:.1f}%" for k, x in sorted(ops.items(), key=lambda x: -x[1])[:18]))
if win:
    for s in range(0, len(d), win):
        ss, ii = sum(int(r[iS]) for r in d[s:s + win]), sum(int(r[iI]) for r in d[s:s + win])
        if ss > 0.01 * tot: print(f"  sass[{s}:{s + win}] samples {100 * ss / tot:5.1f}%  inst {100 * ii / totI:5.1f}%")
