#!/usr/bin/env python
"""Aggregate the per-instruction stall samples of `ncu --page source --csv` output.
  python tools/ncu_stalls.py <source.csv> [kernel-index]   -> stall reasons overall, by opcode, and the hottest instructions
"""
import collections
import csv
import sys

path = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = list(csv.reader(open(path)))
# split into kernels
kern, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        kern.append(cur)
    elif cur is not None and r and r[0] == "Address":
        cur["hdr"] = r
    elif cur is not None and cur["hdr"] and len(r) == len(cur["hdr"]):
        cur["rows"].append(r)
k = kern[which]
h = k["hdr"]
stall_cols = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
si, ii, ei = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
tot = collections.Counter()
byop = collections.defaultdict(lambda: [0, 0])
for r in k["rows"]:
    op = r[si].split()[0]
    if op.startswith("@"):
        op = r[si].split()[1]
    op = op.rstrip(";")
    byop[op][0] += int(r[ii]); byop[op][1] += int(r[ei])
    for c in stall_cols:
        tot[h[c]] += int(r[c])
S = sum(tot.values())
print("#", k["name"])
print("# samples", S, " instructions executed", sum(int(r[ei]) for r in k["rows"]))
for name, v in tot.most_common(10):
    print(f"  {name:28s} {v:8d} {100.0 * v / S:5.1f}%")
print("# by opcode: samples, executed")
for op, (s, e) in sorted(byop.items(), key=lambda kv: -kv[1][0])[:18]:
    print(f"  {op:28s} {s:8d} {100.0 * s / S:5.1f}%  {e}")
print("# hottest instructions")
for r in sorted(k["rows"], key=lambda r: -int(r[ii]))[:25]:
    top = sorted(((int(r[c]), h[c]) for c in stall_cols), reverse=True)[:2]
    print(f"  {int(r[ii]):6d} {r[si].strip()[:70]:70s} {top[0][1]}={top[0][0]} {top[1][1]}={top[1][0]}")
