#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` export: stall samples by SASS opcode and the top instructions.
usage: ncu_src_summary.py src.csv [kernel-index]"""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1])))
blocks = []; cur = None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
ki = int(sys.argv[2]) if len(sys.argv) > 2 else 0
b = blocks[ki]
hdr = b["rows"][0]; data = [r for r in b["rows"][1:] if len(r) == len(hdr)]
ix = {h: i for i, h in enumerate(hdr)}
print(b["name"][:100], "instructions:", len(data))
S = ix["# Samples"]; IE = ix["Instructions Executed"]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[S] or 0) for r in data)
print("total samples", tot, " warp-inst executed", sum(int(r[IE] or 0) for r in data))
agg = collections.Counter(); cnt = collections.Counter(); st = collections.Counter()
for r in data:
    src = r[ix["Source"]].split()
    op = src[0] if src else "?"
    if op.startswith("@") and len(src) > 1:
        op = src[1]
    parts = op.split(".")
    op = parts[0] + ("." + ".".join(parts[1:3]) if parts[0] in ("LD", "ST", "LDG", "STG", "LDL", "STL", "LDS", "STS", "BAR", "SYNCS", "LDC") else "")
    agg[op] += int(r[S] or 0); cnt[op] += int(r[IE] or 0)
    for h in stalls:
        st[h] += int(r[ix[h]] or 0)
print("stall reasons:", ", ".join(f"{h[6:]} {100*v/max(tot,1):.1f}%" for h, v in st.most_common(8)))
print("by opcode (samples%, warp-insts):")
for op, v in agg.most_common(24):
    print(f"  {op:24s} {100*v/tot:5.1f}%  {cnt[op]:>12d}")
print("top instructions:")
top = sorted(data, key=lambda r: -int(r[S] or 0))[:28]
for r in top:
    rs = {h[6:]: int(r[ix[h]] or 0) for h in stalls}
    m = max(rs, key=rs.get)
    print(f"  {r[ix['Address']][-6:]} {100*int(r[S])/tot:4.1f}% {m:10s} {r[ix['Source']][:90]}")
