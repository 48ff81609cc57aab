#!/usr/bin/env python
"""Stall samples of a kernel aggregated by CUDA source line (needs -lineinfo and --import-source on).
usage: python tools/ncu_lines.py X.ncu-rep [top]   -> per file: samples per line, with the line's text"""
import csv, io, subprocess, sys
from collections import defaultdict
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur_file, hdr, total = None, None, 0
per = defaultdict(lambda: [0, 0, ""])   # (file, line) -> samples, instr, text
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = {h: i for i, h in enumerate(r)}
        continue
    if hdr is None or cur_file is None:
        continue
    if r[0].isdigit():      # a CUDA line summary row
        try:
            s = int(r[hdr["# Samples"]]); n = int(r[hdr["Instructions Executed"]] or 0)
        except (ValueError, IndexError):
            continue
        key = (cur_file, int(r[0]))
        per[key][0] += s; per[key][1] += n; per[key][2] = r[1].strip()
        total += s
print("total samples", total)
byfile = defaultdict(int)
for (f, l), (s, n, t) in per.items():
    byfile[f] += s
for f, s in sorted(byfile.items(), key=lambda kv: -kv[1]):
    print("%-40s %8d %5.1f%%" % (f, s, 100.0 * s / max(1, total)))
print()
for (f, l), (s, n, t) in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-24s %5d %8d %5.1f%% %11d  %s" % (f, l, s, 100.0 * s / max(1, total), n, t[:90]))
