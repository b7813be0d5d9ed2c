#!/usr/bin/env python
"""Per-source-line share of executed instructions and stall samples of one .ncu-rep (needs -lineinfo + --import-source on).
usage: python tools/ncu_lines.py x.ncu-rep [top]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur, agg, hdr = None, {}, None
for r in csv.reader(io.StringIO(out)):
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if r[0] == "Function Name": continue
    if r[0].isdigit() and hdr:
        try:
            i, s = int(r[hdr.index("Instructions Executed")]), int(r[hdr.index("# Samples")])
            lsb = int(r[hdr.index("stall_long_sb")]) if "stall_long_sb" in hdr else 0
            agg[(cur, int(r[0]))] = (i, s, lsb, r[1])
        except ValueError:
            pass
tot, ts = sum(v[0] for v in agg.values()), sum(v[1] for v in agg.values())
print("warp instructions", tot, "samples", ts)
for (f, l), (i, s, lsb, src) in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
    print(f"{f}:{l:4d} inst {100 * i / tot:5.2f}% samp {100 * s / ts:5.2f}% (long_sb {100 * lsb / ts:5.2f}%)  {src.strip()[:95]}")
