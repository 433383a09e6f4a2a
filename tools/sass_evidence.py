#!/usr/bin/env python3
"""SASS evidence of the Blackwell-specific paths in libpfaai_gpu.so (no GPU needed): opcode counts per kernel and excerpts around
the tensor-core issue, the bulk copies, the tensor-memory loads / stores, the cp.async ring and the walk's hit path.
Usage: python tools/sass_evidence.py [lib] > profiles/<tag>_sass_evidence.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "parfastaai_b200/libpfaai_gpu.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout.splitlines()
pat = re.compile(r"\b(UTCIMMA|UTCBAR|UTCATOMSWS|UBLKCP|LDTM|STTM|SYNCS|LDGSTS|ATOMS|REDG?|POPC|LOP3)[A-Z0-9_.]*")
fn, counts, first = None, collections.defaultdict(collections.Counter), {}
for i, line in enumerate(sass):
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        continue
    m = pat.search(line)
    if m and fn and "/*" in line:
        op = m.group(0).split(".")[0]
        counts[fn][op] += 1
        first.setdefault((fn, op), i)
print(f"== cuobjdump -sass {lib}: Blackwell-specific / hot-path opcodes per kernel ==")
for f in sorted(counts):
    c = counts[f]
    if any(k in c for k in ("UTCIMMA", "UBLKCP", "LDTM", "LDGSTS")) or "aji_" in f:
        print(f"{f}\n    " + ", ".join(f"{k} x{v}" for k, v in sorted(c.items())))
def excerpt(title, fnsub, op, before=3, after=5):
    for (f, o), i in first.items():
        if fnsub in f and o == op:
            print(f"\n== {title}: first {op} in {f} ==")
            for line in sass[max(0, i - 2 * before): i + 2 * after]:
                if "/*" in line and not re.match(r"\s*/\* 0x", line):
                    print(line.rstrip()[:140])
            return
excerpt("tensor-core issue (tcgen05.mma kind::i8)", "aji_mma_kernel", "UTCIMMA")
excerpt("tcgen05.commit", "aji_mma_kernel", "UTCBAR")
excerpt("bulk async copy (cp.async.bulk) of an operand stage", "aji_mma_kernel", "UBLKCP")
excerpt("bulk async copy of a dictionary stage", "aji_tile_kernel", "UBLKCP")
excerpt("tensor-memory load of the fp64 sums", "aji_tile_kernel", "LDTM")
excerpt("tensor-memory store", "aji_tile_kernel", "STTM")
excerpt("cp.async ring of the sparse row items", "aji_mma_kernel", "LDGSTS")
excerpt("sparse walk: counter increment", "aji_mma_kernel", "ATOMS")
