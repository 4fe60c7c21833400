#!/usr/bin/env python
"""Totals of the warp-stall sampling columns of an `ncu --page source --csv` dump, plus the hottest SASS lines.
usage: ncu -i rep.ncu-rep --page source --csv | python tools/ncu_stalls.py [top_n]"""
import csv
import sys

rows = list(csv.reader(sys.stdin))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
body = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]
col = {n: i for i, n in enumerate(hdr)}
stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
tot = {n: sum(int(r[col[n]] or 0) for r in body) for n in stalls}
all_s = sum(tot.values())
print("samples", all_s, " instructions", len(body))
for n, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    if v:
        print(f"  {n:28s} {v:9d} {100 * v / all_s:5.1f}%")
top = int(sys.argv[1]) if len(sys.argv) > 1 else 25
ss = col["# Samples"]
print("hottest instructions:")
for r in sorted(body, key=lambda r: -int(r[ss] or 0))[:top]:
    reasons = sorted(((int(r[col[n]] or 0), n[6:]) for n in stalls), reverse=True)[:2]
    print(f"  {int(r[ss]):7d}  {r[1].strip()[:70]:70s} {reasons}")
# opcode histogram weighted by executions
ops = {}
ie = col["Instructions Executed"]
for r in body:
    op = r[1].strip().split()[0] if r[1].strip() else "?"
    if op.startswith("@"):
        op = r[1].strip().split()[1]
    ops[op] = ops.get(op, 0) + int(r[ie] or 0)
tot_i = sum(ops.values())
print("warp instructions executed", tot_i)
for op, v in sorted(ops.items(), key=lambda kv: -kv[1])[:22]:
    print(f"  {op:18s} {v:12d} {100 * v / tot_i:5.1f}%")
