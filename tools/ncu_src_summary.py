#!/usr/bin/env python3
"""Summarise an `ncu --page source --csv` export: stall reasons, instruction mix by opcode, hottest SASS lines."""
import csv, sys, collections, re
path = sys.argv[1]
rows = list(csv.reader(open(path)))
# find header row
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
def num(x):
    try: return float(x)
    except: return 0.0
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter()
for r in data:
    for s in stalls: tot[s] += num(r[col[s]])
allsamp = sum(tot.values())
print("== stall reasons (all samples) ==")
for s, v in tot.most_common(10): print(f"{s:28s} {v:10.0f} {100*v/allsamp:5.1f}%")
mix = collections.Counter(); samp = collections.Counter()
for r in data:
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[col["Source"]])
    op = m.group(2).split(".")[0] if m else "?"
    mix[op] += num(r[col["Instructions Executed"]]); samp[op] += num(r[col["# Samples"]])
ti = sum(mix.values()); ts = sum(samp.values())
print("== instruction mix (warp-level executed) / samples ==")
for op, v in mix.most_common(25): print(f"{op:12s} {v:14.0f} {100*v/ti:5.1f}%   samples {100*samp[op]/ts:5.1f}%")
print("total inst", ti)
print("== hottest SASS lines ==")
top = sorted(data, key=lambda r: -num(r[col["# Samples"]]))[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]
for r in top:
    st = sorted(((num(r[col[s]]), s) for s in stalls), reverse=True)[:2]
    print(f"{r[col['Address']][-5:]} {num(r[col['# Samples']]):7.0f} {r[col['Source']][:70]:70s} {st[0][1]}:{st[0][0]:.0f} {st[1][1]}:{st[1][0]:.0f}")
