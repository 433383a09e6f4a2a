#!/bin/bash
# full ncu capture of the index-build kernels: tools/ncu_build_kernels.sh <tag>
set -u
OUT=gpurun_out/$1; mkdir -p $OUT
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_onesweep|k_dict_build|k_make_items|k_pack_fill" -c 6 -o $OUT/build_full $CMD > $OUT/ncu.log 2>&1
echo "ncu rc=$?"; ls -la $OUT
