#!/usr/bin/env python
"""Aggregate an `ncu --page source --print-source sass,cuda --csv` dump by CUDA source line.

    ncu -i x.ncu-rep --page source --print-source sass,cuda --csv > /tmp/src.csv
    python tools/ncu_lines.py /tmp/src.csv [top]
Prints, per (file, line): stall samples, instructions executed, shared wavefronts, top stall reasons.
"""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg = defaultdict(lambda: defaultdict(float))
src = {}
fname = "?"
hdr = None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if r[0] == "Function Name" or hdr is None or len(r) < len(hdr):
        continue
    if r[0] != "":                       # a CUDA source line (its text may be split by unquoted commas)
        try:
            key = (fname, int(r[0]))
        except ValueError:
            continue
        src[key] = ",".join(r[1:-6]).strip()
        continue
    if not r[2].startswith("0x"):
        continue
    d = agg[key]
    for h, v in zip(hdr[4:], r[4:]):
        try:
            d[h] += float(v)
        except ValueError:
            pass
tot = sum(d["# Samples"] for d in agg.values()) or 1.0
toti = sum(d["Instructions Executed"] for d in agg.values()) or 1.0
print(f"total samples {tot:.0f}, warp instructions {toti:.0f}")
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for key, d in sorted(agg.items(), key=lambda kv: -kv[1]["# Samples"])[:top]:
    st = sorted(((d[s], s[6:]) for s in stalls), reverse=True)[:3]
    print(f"{key[0]}:{key[1]:5d} {100*d['# Samples']/tot:5.1f}% smp {100*d['Instructions Executed']/toti:5.1f}% inst "
          f"shw {d['L1 Wavefronts Shared']:.2e} | " + " ".join(f"{n}={100*v/max(d['# Samples'],1):.0f}%" for v, n in st)
          + " | " + src[key][:90])
