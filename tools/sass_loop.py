#!/usr/bin/env python3
"""Opcode histogram of the walk loop of aji_tile_kernel<9,7> in a built library: tools/sass_loop.py lib.so
The loop is taken as the smallest backward-branch region containing >= 8 LDS.64."""
import re, subprocess, sys, collections
lib = sys.argv[1]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
# split functions
funcs = re.split(r"\n\s*Function : ", out)
f = next(x for x in funcs if x.startswith("_ZN5pfaai15aji_tile_kernelILi9ELi7EEEvNS_10TileParamsE"))
ins = []
for line in f.split("\n"):
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
addr_index = {a: i for i, (a, _) in enumerate(ins)}
best = None
for i, (a, t) in enumerate(ins):
    m = re.search(r"BRA\s+(?:`\()?(0x[0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt < a and tgt in addr_index:
            j = addr_index[tgt]
            body = ins[j:i + 1]
            n64 = sum(1 for _, x in body if "LDS.64" in x)
            if n64 >= 8 and (best is None or len(body) < len(best)): best = body
if best is None: sys.exit("loop not found")
hist = collections.Counter()
for _, t in best:
    t = re.sub(r"^@!?U?P\d+\s+", "", t)
    hist[t.split()[0].split(".")[0]] += 1
ALU = {"LOP3", "SHF", "LEA", "IADD3", "ISETP", "SEL", "VIADD", "VIMNMX", "PRMT", "POPC", "LOP", "IADD", "MOV", "PLOP3", "ISETP"}
print("loop instructions:", len(best), " LDS.64:", sum(1 for _, x in best if "LDS.64" in x))
print("ALU-pipe-ish:", sum(v for k, v in hist.items() if k in ALU), " IMAD:", hist["IMAD"])
print(dict(hist.most_common()))
if len(sys.argv) > 2:
    for a, t in best: print(hex(a), t)
