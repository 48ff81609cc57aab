#!/usr/bin/env python
"""DRAM bytes per launch of the hot kernels from an ncu report of `bench.py` -> profiles/traffic.json.
usage: python tools/ncu_traffic.py X.ncu-rep [out.json]"""
import csv, io, json, re, subprocess, sys
rep = sys.argv[1]
out = sys.argv[2] if len(sys.argv) > 2 else "profiles/traffic.json"
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics",
                      "dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,launch__grid_size"],
                     stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
head = rows[0]; units = rows[1]
idx = {h: i for i, h in enumerate(head)}
def val(r, name):
    v = float(r[idx[name]].replace(",", ""))
    u = units[idx[name]].lower()
    mult = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "msecond": 1.0, "usecond": 1e-3, "nsecond": 1e-6, "second": 1e3}.get(u, 1)
    return v * mult
res = {}
for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    m = re.search(r"(k_blind_rotate\w*|k_keyswitch\w*)", name)
    if not m: continue
    key = "k_blind_rotate" if m.group(1).startswith("k_blind_rotate") else m.group(1)
    d = {"dram_bytes": val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum"),
         "dram_read_bytes": val(r, "dram__bytes_read.sum"), "dram_write_bytes": val(r, "dram__bytes_write.sum"),
         "ms_under_ncu": val(r, "gpu__time_duration.sum"), "grid": int(float(r[idx["launch__grid_size"]])), "kernel": name, "report": rep}
    if key not in res or d["grid"] > res[key]["grid"]: res[key] = d
json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res, indent=1))
