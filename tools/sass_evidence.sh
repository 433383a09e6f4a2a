#!/bin/bash
# SASS evidence of the Blackwell-specific paths in libpfaai_gpu.so (no GPU needed): opcode counts per kernel and excerpts
# around the tensor-core issue, the bulk copies and the tensor-memory loads/stores.  Usage: tools/sass_evidence.sh > profiles/<tag>_sass_evidence.txt
set -u
LIB=${1:-parfastaai_b200/libpfaai_gpu.so}
TMP=$(mktemp)
cuobjdump -sass "$LIB" > "$TMP"
echo "== cuobjdump -sass $LIB: Blackwell-specific opcodes per kernel =="
awk '/Function :/ {fn=$3} /UTCIMMA|UTCBAR|UBLKCP|LDTM|STTM|SYNCS|LDGSTS|ATOMS|REDS|RED\./ {match($0,/(UTCIMMA[A-Z0-9._]*|UTCBAR[A-Z0-9._]*|UBLKCP[A-Z0-9._]*|LDTM[A-Z0-9._]*|STTM[A-Z0-9._]*|SYNCS[A-Z0-9._]*|LDGSTS[A-Z0-9._]*|ATOMS[A-Z0-9._]*|RED[A-Z0-9._]*)/,m); if (m[1]!="") c[fn" "m[1]]++} END {for (k in c) print c[k], k}' "$TMP" | sort -k2,2 -k1,1nr
for pat in UTCIMMA UTCBAR "UBLKCP" "LDTM" "LDGSTS"; do
  echo; echo "== first site of $pat (with 4 lines of context) =="
  grep -n -m1 -B4 -A4 "$pat" "$TMP" | sed 's/^\([0-9]*\)[:-]//' | cut -c1-150
done
echo; echo "== the sparse walk's hit path in aji_mma_kernel (LDS.U16 rank / entry, ATOMS/RED on the counters) =="
awk '/Function : .*aji_mma_kernel.*32590/ {on=1} /Function :/ && !/aji_mma_kernel.*32590/ {on=0} on && /LDS.U16|ATOMS|POPC|REDS/ {print}' "$TMP" | head -24 | cut -c1-150
rm -f "$TMP"
