#!/bin/bash
# 1 x B200: ncu capture (source level) of the hybrid kernel on the sparse family at config-4 size
set -u
OUT=gpurun_out/${TAG:-r2o}; mkdir -p $OUT
timeout 300 python bench.py --genomes 10000 --path mma --tau 1024 --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:aji_mma -s 1 -c 1 -o $OUT/mma_sparse_full \
    python bench.py --genomes 10000 --path mma --tau 1024 --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/ncu.log 2>&1
echo "ncu rc=$?"; ls -la $OUT
