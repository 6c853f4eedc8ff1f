#!/usr/bin/env python
"""Opcode mix of a kernel from `ncu -i X.ncu-rep --page source --csv --print-source sass`.
usage: sass_mix.py src.csv [launch_index]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else -1
# split into kernels
kern = []
for r in rows:
    if r and r[0] == "Kernel Name":
        kern.append({"name": r[1], "rows": []})
    elif kern:
        kern[-1]["rows"].append(r)
k = kern[which]
hdr = k["rows"][0]
ie = hdr.index("Instructions Executed"); so = hdr.index("Source"); sm = hdr.index("# Samples")
mix = collections.Counter(); samp = collections.Counter(); tot = 0; static = 0
for r in k["rows"][1:]:
    if len(r) <= ie: continue
    op = r[so].split()
    if not op: continue
    o = op[1] if op[0].startswith("@") else op[0]
    o = o.split(".")[0].rstrip(";")
    n = int(r[ie] or 0)
    mix[o] += n; tot += n; static += 1
    samp[o] += int(r[sm] or 0)
print(k["name"][:100], "static", static, "executed warp-inst", tot)
ts = sum(samp.values()) or 1
for o, n in mix.most_common(30):
    print(f"{o:10s} {n:12d} {100*n/tot:5.1f}%  samples {100*samp[o]/ts:5.1f}%")
