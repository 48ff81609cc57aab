#!/usr/bin/env python
"""Split the SASS page of an ncu report into phases at marker instructions and print, per phase, the
stall samples and executed warp instructions by pipe class.
usage: python tools/sass_phases.py X.ncu-rep [kernel-regex] [min-samples]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], stdout=subprocess.PIPE, text=True).stdout
MARK = re.compile(r"\b(WARPSYNC|BAR|SYNCS|UBLKCP|ATOMS)\b")
def cls(op):
    if op.startswith(("DFMA", "DADD", "DMUL")): return "fp64"
    if op.startswith(("LDS", "STS", "SHFL", "ATOMS", "LDG", "STG", "LDL", "STL")): return "lsu"
    if op.startswith(("LDTM", "STTM")): return "tmem"
    if op.startswith(("IMAD.MOV", "MOV")): return "mov"
    return "other"
for blk in out.split('"Kernel Name",')[1:]:
    lines = blk.splitlines()
    if len(sys.argv) > 2 and sys.argv[2] and not re.search(sys.argv[2], lines[0]): continue
    rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
    head = rows[0]; idx = {h: i for i, h in enumerate(head)}
    phases = []; cur = {"start": None, "n": 0, "s": 0, "cls": {}, "mark": "entry", "st": {}}
    stall_cols = [h for h in head if h.startswith("stall_") and "Not Issued" not in h]
    tot = 0
    for r in rows[1:]:
        if len(r) < len(head): continue
        src = r[idx["Source"]]
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
        op = m.group(2) if m else "?"
        s = int(r[idx["# Samples"]] or 0); n = int(r[idx["Instructions Executed"]] or 0)
        if MARK.search(src) and cur["n"]:
            phases.append(cur)
            cur = {"start": r[idx["Address"]], "n": 0, "s": 0, "cls": {}, "mark": op, "st": {}}
        if cur["start"] is None: cur["start"] = r[idx["Address"]]
        cur["n"] += n; cur["s"] += s; tot += s
        c = cls(op); cur["cls"][c] = cur["cls"].get(c, 0) + n
        for h in stall_cols:
            v = int(r[idx[h]] or 0)
            if v: cur["st"][h[6:]] = cur["st"].get(h[6:], 0) + v
    phases.append(cur)
    mins = int(sys.argv[3]) if len(sys.argv) > 3 else tot // 200
    print("kernel:", lines[0][:100], " total samples", tot)
    print("%-8s %-9s %8s %6s %11s  fp64/lsu/tmem/mov/other (M warp-instr)   top stalls" % ("addr", "marker", "samples", "%", "warp-instr"))
    for p in phases:
        if p["s"] < mins: continue
        c = p["cls"]
        st = ", ".join("%s %.0f%%" % (k, 100.0 * v / max(1, p["s"])) for k, v in sorted(p["st"].items(), key=lambda kv: -kv[1])[:4])
        print("%-8s %-9s %8d %5.1f%% %11d  %5.0f/%5.0f/%4.0f/%5.0f/%5.0f  %s" % (p["start"][-5:], p["mark"][:9], p["s"], 100.0 * p["s"] / tot, p["n"],
              c.get("fp64", 0) / 1e6, c.get("lsu", 0) / 1e6, c.get("tmem", 0) / 1e6, c.get("mov", 0) / 1e6, c.get("other", 0) / 1e6, st))
