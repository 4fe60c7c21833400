#!/usr/bin/env python
"""Dynamic SASS opcode mix of the first kernel in an .ncu-rep (source page).  Usage: ncu_opmix.py file.ncu-rep [rows]"""
import csv, io, subprocess, sys
from collections import defaultdict
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
si, ei, st = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
mix, samp = defaultdict(int), defaultdict(int)
tot = 0
for r in rows[2:]:
    if len(r) <= ei or not r[ei].isdigit():
        if len(r) > 1 and r[0] == "Kernel Name":
            break
        continue
    toks = r[si].split()
    op = toks[1] if toks[0].startswith("@") else toks[0]
    mix[op] += int(r[ei]); samp[op] += int(r[st]) if r[st].isdigit() else 0
    tot += int(r[ei])
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
print(f"total warp instructions {tot}  (per unit: {tot/div:.1f})")
for op in sorted(mix, key=lambda o: -mix[o])[:28]:
    print(f"{op:24s} {mix[op]:12d} {100*mix[op]/tot:6.2f}%  per-unit {mix[op]/div:9.2f}  samples {samp[op]}")
