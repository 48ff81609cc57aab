#!/usr/bin/env python
"""Aggregate the SASS page of an ncu report: stall samples and executed instructions by opcode, and the
hottest instructions.  usage: python tools/sass_stalls.py X.ncu-rep [kernel-regex] [top]"""
import csv, io, re, subprocess, sys
from collections import defaultdict
rep = sys.argv[1]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], stdout=subprocess.PIPE, text=True).stdout
blocks = out.split('"Kernel Name",')
for blk in blocks[1:]:
    lines = blk.splitlines()
    name = lines[0]
    if len(sys.argv) > 2 and sys.argv[2] and not re.search(sys.argv[2], name):
        continue
    rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
    head = rows[0]
    idx = {h: i for i, h in enumerate(head)}
    stall_cols = [h for h in head if h.startswith("stall_") and "Not Issued" not in h]
    byop = defaultdict(lambda: [0, 0, defaultdict(int)])
    tot_s = tot_i = 0
    recs = []
    for r in rows[1:]:
        if len(r) < len(head):
            continue
        src = r[idx["Source"]]
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
        op = m.group(2) if m else "?"
        op = ".".join(op.split(".")[:2]) if op.startswith(("LDS", "STS", "LDTM", "STTM", "LDG", "STG", "BAR", "SHFL")) else op.split(".")[0]
        s = int(r[idx["# Samples"]] or 0)
        n = int(r[idx["Instructions Executed"]] or 0)
        byop[op][0] += s
        byop[op][1] += n
        for c in stall_cols:
            v = int(r[idx[c]] or 0)
            if v:
                byop[op][2][c[6:]] += v
        tot_s += s
        tot_i += n
        recs.append((s, n, src, {c[6:]: int(r[idx[c]] or 0) for c in stall_cols if int(r[idx[c]] or 0)}))
    print("kernel:", name[:120])
    print("total samples %d, warp instructions %d" % (tot_s, tot_i))
    print("%-12s %9s %6s %12s %6s  top stall reasons" % ("opcode", "samples", "%", "warp-instr", "%"))
    for op, (s, n, st) in sorted(byop.items(), key=lambda kv: -kv[1][0])[:top]:
        reasons = ", ".join("%s %.0f%%" % (k, 100.0 * v / max(1, s)) for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:4])
        print("%-12s %9d %5.1f%% %12d %5.1f%%  %s" % (op, s, 100.0 * s / max(1, tot_s), n, 100.0 * n / max(1, tot_i), reasons))
    print("hottest instructions:")
    for s, n, src, st in sorted(recs, key=lambda x: -x[0])[:top]:
        reasons = ", ".join("%s %d" % (k, v) for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print("  %7d  %-70s %s" % (s, src[:70], reasons))
