#!/usr/bin/env python
"""Local-memory (spill) instructions by CUDA source line from `ncu --page source --print-source sass,cuda --csv`.

    python tools/ncu_local.py /tmp/src.csv [top]
"""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr = None; fname = "?"; key = None
agg = defaultdict(lambda: [0, 0.0, 0.0]); src = {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; ie = hdr.index("Instructions Executed"); isamp = hdr.index("# Samples"); continue
    if hdr is None or r[0] == "Function Name":
        continue
    if r[0] != "":
        try:
            key = (fname, int(r[0])); src[key] = ",".join(r[1:-6])[:90]
        except ValueError:
            pass
        continue
    if len(r) < len(hdr) or not r[2].startswith("0x"):
        continue
    ins = r[3]
    if "LDL" in ins or "STL" in ins:
        a = agg[(key, "LDL" if "LDL" in ins else "STL")]
        a[0] += 1; a[1] += float(r[ie] or 0); a[2] += float(r[isamp] or 0)
tot = sum(a[1] for a in agg.values())
print("total local ld/st warp instructions executed: %.3e" % tot)
for (k, kind), a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{k[0]}:{k[1]:5d} {kind} n={a[0]:3d} exec={a[1]:.2e} ({100*a[1]/tot:4.1f}%) samples={a[2]:.0f} | {src[k]}")
